// surface_contact_force_functor: particle - wall Luding contact -- parameter holder
// (reference constructor: include/libgran/contact_force/surface_contact_force.h:10-30; law :57-145 ->
// libgran_b200/csrc/physics.cuh: wall_contact_force; per-particle spring table :53 -> device wspr[facet]).
#ifndef LIBGRAN_SURFACE_CONTACT_FORCE_H
#define LIBGRAN_SURFACE_CONTACT_FORCE_H

#include <cstddef>

#include <libgran/contact_force/contact_force.h>

template <typename field_value_t, typename real_t>
struct surface_contact_force_functor {
    surface_contact_force_functor(std::size_t n_part, real_t k, real_t gamma_n, real_t k_t, real_t gamma_t, real_t mu_s,
                                  real_t phi_d, real_t k_r, real_t gamma_r, real_t mu_r, real_t phi_r, real_t k_o,
                                  real_t gamma_o, real_t mu_o, real_t phi_o, real_t r_part, real_t mass, real_t inertia,
                                  real_t dt, field_value_t /*field_zero*/, real_t /*real_zero*/)
        : n_part(n_part),
          contact(lgb200::make_contact_params(k, gamma_n, k_t, gamma_t, mu_s, phi_d, k_r, gamma_r, mu_r, phi_r, k_o, gamma_o,
                                              mu_o, phi_o, dt)) {
        constants.r_part = r_part; constants.have_r = true;
        constants.mass = mass; constants.have_mass = true;
        constants.inertia = inertia; constants.have_inertia = true;
    }

    static constexpr bool lgb200_surface_contact = true;
    lgb200::particle_constants lgb200_constants() const { return constants; }
    lg_contact_params const& params() const { return contact; }

private:
    std::size_t n_part;
    lg_contact_params contact;
    lgb200::particle_constants constants;
};

#endif  // LIBGRAN_SURFACE_CONTACT_FORCE_H
