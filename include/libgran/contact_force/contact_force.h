// contact_force_functor: Luding frictional contact between two equal spheres -- host-side parameter holder.
// Same constructor as the reference (include/libgran/contact_force/contact_force.h:10-30).  The force law
// (:57-147) runs on the GPU (libgran_b200/csrc/physics.cuh: contact_force / shear_spring); the dense N x N spring
// store (:53,156) is replaced by the device's sparse per-ordered-pair slot table.
#ifndef LIBGRAN_CONTACT_FORCE_H
#define LIBGRAN_CONTACT_FORCE_H

#include <cstddef>

#include <libgran/backend/lgb200_backend.h>

namespace lgb200 {
template <typename real_t>
inline lg_contact_params make_contact_params(real_t k, real_t gamma_n, real_t k_t, real_t gamma_t, real_t mu_s, real_t phi_d,
                                             real_t k_r, real_t gamma_r, real_t mu_r, real_t phi_r, real_t k_o, real_t gamma_o,
                                             real_t mu_o, real_t phi_o, real_t dt) {
    lg_contact_params p;
    p.k = k; p.gamma_n = gamma_n;
    p.k_t = k_t; p.gamma_t = gamma_t; p.mu_s = mu_s; p.phi_d = phi_d;
    p.k_r = k_r; p.gamma_r = gamma_r; p.mu_r = mu_r; p.phi_r = phi_r;
    p.k_o = k_o; p.gamma_o = gamma_o; p.mu_o = mu_o; p.phi_o = phi_o;
    p.dt = dt;
    return p;
}
}  // namespace lgb200

template <typename field_value_t, typename real_t>
struct contact_force_functor {
    contact_force_functor(std::size_t n_part, real_t k, real_t gamma_n, real_t k_t, real_t gamma_t, real_t mu_s, real_t phi_d,
                          real_t k_r, real_t gamma_r, real_t mu_r, real_t phi_r, real_t k_o, real_t gamma_o, real_t mu_o,
                          real_t phi_o, real_t r_part, real_t mass, real_t inertia, real_t dt, field_value_t /*field_zero*/,
                          real_t /*real_zero*/)
        : n_part(n_part),
          contact(lgb200::make_contact_params(k, gamma_n, k_t, gamma_t, mu_s, phi_d, k_r, gamma_r, mu_r, phi_r, k_o, gamma_o,
                                              mu_o, phi_o, dt)) {
        constants.r_part = r_part; constants.have_r = true;
        constants.mass = mass; constants.have_mass = true;
        constants.inertia = inertia; constants.have_inertia = true;
    }

    // ---- backend protocol ----
    static constexpr bool lgb200_binary_model = true;
    lgb200::particle_constants lgb200_constants() const { return constants; }
    std::size_t lgb200_n_part() const { return n_part; }
    void lgb200_install(lg_system* s) const { LGB200_CHECK(lg_set_contact(s, &contact)); }
    lg_contact_params const& params() const { return contact; }

private:
    std::size_t n_part;
    lg_contact_params contact;
    lgb200::particle_constants constants;
};

#endif  // LIBGRAN_CONTACT_FORCE_H
