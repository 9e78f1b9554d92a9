// sinter_functor / sinter_step_handler: sinter-bridge bonded springs with an embedded contact model.
// Public surface as in the reference (include/libgran/sinter_bridge/sinter_bridge.h): constructor :101-126,
// get_step_handler<container>() :228-232, sinter_step_handler<container, value, real> :14-94.
//
// The constructor runs the bond scan where the reference runs it -- immediately, consuming the caller's random
// unit vector generator in lexicographic (i < j) pair order (:148-211) -- through lg_bond_table_create (host code
// in liblgb200.so; a uniform grid replaces the O(N^2) loop, the dense N x N tables become 2 records per bond).
// The bonded force (:235-259) and the connector motion of the step handler (:24-42, :55-75) run on the GPU.
#ifndef LIBGRAN_SINTER_BRIDGE_H
#define LIBGRAN_SINTER_BRIDGE_H

#include <cstddef>
#include <functional>
#include <memory>
#include <vector>

#include <libgran/contact_force/contact_force.h>
#include <libtimestep/rotational_step_handler/rotational_step_handler.h>

// Step handler that also carries the five bridge connectors of every bond with their particle.  On the device
// this is the connector stage of the fused step kernel, armed by installing the bond table; the object is the
// token that ties a system to the functor the table came from.
template <typename field_container_t, typename field_value_t, typename real_t>
struct sinter_step_handler : rotational_step_handler<field_container_t, field_value_t> {
    static constexpr int lgb200_step_handler_kind = lgb200::STEP_SINTER_CONNECTORS;
    explicit sinter_step_handler(const lg_bond_table* table = nullptr) : table(table) {}
    const lg_bond_table* table;
};

template <typename field_value_t, typename real_t>
struct sinter_functor {
    sinter_functor(std::size_t n_part, real_t k_d, real_t gamma_d, real_t k, real_t gamma_n, real_t k_t, real_t gamma_t,
                   real_t mu_s, real_t phi_d, real_t k_r, real_t gamma_r, real_t mu_r, real_t phi_r, real_t k_o, real_t gamma_o,
                   real_t mu_o, real_t phi_o, real_t r_part, real_t mass, real_t inertia, real_t dt, field_value_t field_zero,
                   real_t real_zero, typename std::vector<field_value_t>::const_iterator x0_begin_itr,
                   std::function<field_value_t()> random_unit_vector_generator, real_t critical_separation)
        : n_part(n_part), k_d(k_d), gamma_d(gamma_d),
          contact_force(n_part, k, gamma_n, k_t, gamma_t, mu_s, phi_d, k_r, gamma_r, mu_r, phi_r, k_o, gamma_o, mu_o, phi_o,
                        r_part, mass, inertia, dt, field_zero, real_zero) {
        std::vector<double> x(3 * n_part);
        for (std::size_t i = 0; i < n_part; i++)
            for (int c = 0; c < 3; c++) x[3 * i + c] = static_cast<double>((*(x0_begin_itr + i))[c]);
        lg_bond_table* raw = nullptr;
        LGB200_CHECK(lg_bond_table_create(static_cast<int64_t>(n_part), x.data(), r_part, critical_separation, &draw,
                                          &random_unit_vector_generator, &raw));
        table.reset(raw, [](lg_bond_table* t) { lg_bond_table_destroy(t); });
    }

    // Create an instance of sinter_step_handler to be used with the granular system
    template <typename field_container_t>
    sinter_step_handler<field_container_t, field_value_t, real_t> get_step_handler() {
        return sinter_step_handler<field_container_t, field_value_t, real_t>(table.get());
    }

    // number of bonds (unordered pairs) found at construction
    std::size_t n_bonds() const {
        int64_t b = 0;
        LGB200_CHECK(lg_bond_table_count(table.get(), &b));
        return static_cast<std::size_t>(b);
    }

    // ---- backend protocol ----
    static constexpr bool lgb200_binary_model = true;
    static constexpr bool lgb200_needs_connector_step_handler = true;
    lgb200::particle_constants lgb200_constants() const { return contact_force.lgb200_constants(); }
    std::size_t lgb200_n_part() const { return n_part; }
    void lgb200_install(lg_system* s) const {
        LGB200_CHECK(lg_set_bond_table(s, &contact_force.params(), k_d, gamma_d, table.get()));
    }
    const lg_bond_table* lgb200_table() const { return table.get(); }

private:
    static void draw(void* user, double out[3]) {
        auto& gen = *static_cast<std::function<field_value_t()>*>(user);
        field_value_t v = gen();
        for (int c = 0; c < 3; c++) out[c] = static_cast<double>(v[c]);
    }

    std::size_t n_part;
    double k_d, gamma_d;
    contact_force_functor<field_value_t, real_t> contact_force;  // embedded contact model (reference :134-135)
    std::shared_ptr<lg_bond_table> table;
};

#endif  // LIBGRAN_SINTER_BRIDGE_H
