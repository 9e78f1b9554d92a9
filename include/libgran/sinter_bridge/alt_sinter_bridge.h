// alt_sinter_functor: bonds that reuse the contact kinematics (normal + sliding / rolling / torsion springs, no Coulomb
// limit, attractive as well as repulsive) with an embedded contact model for the non-bonded pairs -- parameter holder.
// Public surface as in the reference (include/libgran/sinter_bridge/alt_sinter_bridge.h): constructor :18-89.
//
// The constructor runs the bond scan where the reference runs it (:65-88: |dist - 2 r| < critical_separation in
// lexicographic i < j order, a bond that would close a cycle is skipped) through lg_alt_bond_table_create (host code in
// liblgb200.so).  The bonded law (:92-131) and its spring update (:141-160) run on the GPU
// (libgran_b200/csrc/physics.cuh: alt_bonded_force).  Used with the plain rotational_step_handler (test/alt_sintered.cpp:100).
#ifndef LIBGRAN_ALT_SINTER_BRIDGE_H
#define LIBGRAN_ALT_SINTER_BRIDGE_H

#include <cstddef>
#include <memory>
#include <vector>

#include <libgran/contact_force/contact_force.h>

template <typename field_value_t, typename real_t>
struct alt_sinter_functor {
    alt_sinter_functor(std::size_t n_part, std::vector<field_value_t> x0, real_t k, real_t gamma_n, real_t k_t, real_t gamma_t,
                       real_t k_r, real_t gamma_r, real_t k_o, real_t gamma_o, real_t r_part, real_t /*mass*/, real_t /*inertia*/,
                       real_t dt, field_value_t /*field_zero*/, real_t /*real_zero*/, real_t critical_separation,
                       contact_force_functor<field_value_t, real_t> contact_force)
        : n_part(n_part), bond{k, gamma_n, k_t, gamma_t, k_r, gamma_r, k_o, gamma_o, dt}, contact_force(std::move(contact_force)) {
        std::vector<double> x;
        lgb200::pack(x0, x);
        lg_bond_table* raw = nullptr;
        LGB200_CHECK(lg_alt_bond_table_create(static_cast<int64_t>(x0.size()), x.data(), r_part, critical_separation, &raw));
        table.reset(raw, [](lg_bond_table* t) { lg_bond_table_destroy(t); });
    }

    // number of bonds (unordered pairs) that survived cycle prevention
    std::size_t n_bonds() const {
        int64_t b = 0;
        LGB200_CHECK(lg_bond_table_count(table.get(), &b));
        return static_cast<std::size_t>(b);
    }

    // ---- backend protocol ----
    static constexpr bool lgb200_binary_model = true;
    lgb200::particle_constants lgb200_constants() const { return contact_force.lgb200_constants(); }
    void lgb200_install(lg_system* s) const { LGB200_CHECK(lg_set_alt_bond_table(s, &contact_force.params(), bond, table.get())); }

private:
    std::size_t n_part;
    double bond[9];  // k, gamma_n, k_t, gamma_t, k_r, gamma_r, k_o, gamma_o, dt
    contact_force_functor<field_value_t, real_t> contact_force;  // non-bonded pairs (reference :95-96)
    std::shared_ptr<lg_bond_table> table;
};

#endif  // LIBGRAN_ALT_SINTER_BRIDGE_H
