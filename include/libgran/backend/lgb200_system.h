// Shared implementation of granular_system_b200 and granular_system_neighbor_list_b200: owns the device handle,
// installs the force models found in the containers, mirrors the six state vectors lazily.
// Reference counterparts: libtimestep/rotational_system/rotational_system.h:41-130 (state, getters, do_step),
// rotational_binary_system_omp.h:60-91 / rotational_binary_system_neighbors_omp.h:69-116 (pair loops, neighbor list),
// include/libgran/granular_system/granular_system.h:25-86.
#ifndef LIBGRAN_LGB200_SYSTEM_H
#define LIBGRAN_LGB200_SYSTEM_H

#include <array>
#include <cstdint>
#include <tuple>
#include <type_traits>
#include <utility>
#include <vector>

#include <libgran/backend/lgb200_backend.h>
#include <libtimestep/rotational_step_handler/rotational_step_handler.h>

namespace lgb200 {

template <typename field_value_t, typename real_t,
          template <typename, typename, typename, typename, template <typename, typename> typename> typename integrator_t,
          template <typename, typename> typename step_handler_t, typename binary_force_functor_container_t,
          typename unary_force_functor_container_t>
class device_system {
public:
    typedef std::vector<field_value_t> field_container_t;
    typedef step_handler_t<field_container_t, field_value_t> step_handler_instance_t;
    static constexpr std::size_t n_unary =
        std::tuple_size<decltype(unary_force_functor_container_t::unary_force_functors)>::value;
    static constexpr std::size_t n_binary =
        std::tuple_size<decltype(binary_force_functor_container_t::binary_force_functors)>::value;

    device_system(device_system const&) = delete;
    device_system& operator=(device_system const&) = delete;

    ~device_system() { lg_destroy(handle); }

    // do_step(dt): reference rotational_system.h:89-91 -> rotational_velocity_verlet_half.h:43-76
    void do_step(real_t dt) {
        sync_facets();
        LGB200_CHECK(lg_do_step(handle, static_cast<double>(dt)));
        invalidate();
    }

    // The driver loop `for k: if (k % rebuild_every == 0) update_neighbor_list(); do_step(dt);` (and
    // facet.update_positions(dt) after each step when move_facets) queued without returning to the host between
    // steps.  Extension; results are identical to calling the members one by one.
    void run(real_t dt, std::int64_t n_steps, std::int64_t rebuild_every = 0, bool move_facets = false) {
        sync_facets();
        LGB200_CHECK(lg_run(handle, static_cast<double>(dt), n_steps, rebuild_every, move_facets ? 1 : 0));
        if (move_facets) pull_facets();
        invalidate();
    }

    field_container_t const& get_x() const { return mirror(LG_X); }
    field_container_t const& get_v() const { return mirror(LG_V); }
    field_container_t const& get_a() const { return mirror(LG_A); }
    field_container_t const& get_theta() const { return mirror(LG_THETA); }
    field_container_t const& get_omega() const { return mirror(LG_OMEGA); }
    field_container_t const& get_alpha() const { return mirror(LG_ALPHA); }

    void reset_acceleration_buffers() {
        LGB200_CHECK(lg_reset_acceleration_buffers(handle));
        invalidate();
    }

    // test/compute_energy.cpp:7-58 and test/mass_distribution.cpp:7-26 as device reductions:
    // {KE, translational KE, rotational KE, |P|, |L|, com x, com y, com z, radius of gyration}
    std::array<double, 9> diagnostics() const {
        std::array<double, 9> out{};
        LGB200_CHECK(lg_diagnostics(handle, out.data()));
        return out;
    }

    lg_system* native_handle() const { return handle; }
    void synchronize() const { LGB200_CHECK(lg_synchronize(handle)); }

protected:
    device_system(bool neighbor_list, long n_part_arg, real_t r_verlet, field_container_t x0, field_container_t v0,
                  field_container_t theta0, field_container_t omega0, step_handler_instance_t& step_handler,
                  binary_force_functor_container_t binary, unary_force_functor_container_t unary)
        : binary_force_functors(std::move(binary)), unary_force_functors(std::move(unary)) {
        static_assert(requires { integrator_t<field_container_t, field_value_t, real_t, int, step_handler_t>::lgb200_integrator_kind; },
                      "libgran-b200: this integrator has no device implementation (rotational_velocity_verlet_half and "
                      "rotational_forward_euler only)");
        static_assert(requires { step_handler_instance_t::lgb200_step_handler_kind; },
                      "libgran-b200: user-written step handlers cannot run on the device "
                      "(rotational_step_handler, sinter_step_handler and fixed_particle_step_handler over either; no CPU fallback)");
        static_assert(n_binary + n_unary > 0, "at least one force functor must be provided");
        if (x0.size() != v0.size() || x0.size() != theta0.size() || x0.size() != omega0.size())
            throw SizeMismatchException("Size mismatch in arguments passed to granular system constructor: x0, v0, theta0, "
                                        "omega0 must have the same size", __FILE__, __LINE__);
        if (neighbor_list && n_part_arg >= 0 && static_cast<std::size_t>(n_part_arg) != x0.size())
            throw SizeMismatchException("n_part does not match the size of x0", __FILE__, __LINE__);

        particle_constants pc;
        std::apply([&pc](auto&... f) { (pc.merge(f.lgb200_constants()), ...); }, binary_force_functors.binary_force_functors);
        std::apply([&pc](auto&... f) { (pc.merge(f.lgb200_constants()), ...); }, unary_force_functors.unary_force_functors);
        if (!pc.have_r || !pc.have_mass) throw Lgb200BackendError(LG_EINVAL, "no force functor supplies r_part and mass");
        if (!pc.have_inertia) pc.inertia = 2.0 / 5.0 * pc.mass * pc.r_part * pc.r_part;  // Hamaker-only systems: no torque arises

        n = static_cast<std::int64_t>(x0.size());
        LGB200_CHECK(lg_create(n, pc.r_part, pc.mass, pc.inertia, -1, &handle));
        try {
            std::vector<double> bx, bv, bt, bw;
            pack(x0, bx); pack(v0, bv); pack(theta0, bt); pack(omega0, bw);
            LGB200_CHECK(lg_set_state(handle, n, bx.data(), bv.data(), bt.data(), bw.data()));
            bool needs_connectors = false;
            std::apply(
                [&](auto&... f) {
                    (
                        [&](auto& g) {
                            using T = std::decay_t<decltype(g)>;
                            static_assert(requires { T::lgb200_binary_model; },
                                          "libgran-b200: this binary force functor has no device implementation "
                                          "(contact_force_functor, hamaker_functor, sinter_functor; no CPU fallback)");
                            g.lgb200_install(handle);
                            if constexpr (requires { T::lgb200_needs_connector_step_handler; }) needs_connectors = true;
                        }(f),
                        ...);
                },
                binary_force_functors.binary_force_functors);
            if (needs_connectors && step_handler_instance_t::lgb200_step_handler_kind != STEP_SINTER_CONNECTORS)
                throw Lgb200BackendError(LG_EINVAL, "sinter_functor requires the step handler returned by its get_step_handler()");
            constexpr int integrator_kind = integrator_t<field_container_t, field_value_t, real_t, int, step_handler_t>::lgb200_integrator_kind;
            if constexpr (integrator_kind != 0) LGB200_CHECK(lg_set_integrator(handle, integrator_kind));
            if constexpr (requires { step_handler_instance_t::lgb200_fixes_a_particle; })
                LGB200_CHECK(lg_set_fixed_particle(handle, static_cast<std::int64_t>(step_handler.lgb200_fixed_particle())));
            (void)step_handler;
            std::size_t k = 0;
            std::apply(
                [&](auto&... f) {
                    (
                        [&](auto& g) {
                            using T = std::decay_t<decltype(g)>;
                            static_assert(requires { T::lgb200_unary_model; },
                                          "libgran-b200: this unary force functor has no device implementation "
                                          "(triangular_facet and central_attraction_functor only; no CPU fallback)");
                            facet_ids[k++] = g.lgb200_install(handle);
                        }(f),
                        ...);
                },
                unary_force_functors.unary_force_functors);
            if (neighbor_list) LGB200_CHECK(lg_enable_neighbor_list(handle, static_cast<double>(r_verlet)));
        } catch (...) {
            lg_destroy(handle);
            handle = nullptr;
            throw;
        }
    }

    void invalidate() {
        for (bool& f : fresh) f = false;
    }

    lg_system* handle = nullptr;
    std::int64_t n = 0;

private:
    field_container_t const& mirror(int field) const {
        if (!fresh[field]) {
            scratch.resize(3 * static_cast<std::size_t>(n));
            LGB200_CHECK(lg_get_field(handle, field, scratch.data()));
            unpack(scratch, host[field]);
            fresh[field] = true;
        }
        return host[field];
    }

    void sync_facets() {
        std::size_t k = 0;
        std::apply([&](auto&... f) { (f.lgb200_sync(handle, facet_ids[k++]), ...); }, unary_force_functors.unary_force_functors);
    }

    // after lg_run moved the facets on the device side, bring the holders' vertices up to date
    void pull_facets() {
        std::size_t k = 0;
        std::apply(
            [&](auto&... f) {
                (
                    [&](auto& g) {
                        const int id = facet_ids[k++];
                        if constexpr (requires(double* q) { g.lgb200_set_vertices(q); }) {
                            double vtx[9];
                            LGB200_CHECK(lg_facet_get_vertices(handle, id, vtx));
                            g.lgb200_set_vertices(vtx);
                        }
                    }(f),
                    ...);
            },
            unary_force_functors.unary_force_functors);
    }

    binary_force_functor_container_t binary_force_functors;
    unary_force_functor_container_t unary_force_functors;
    std::array<int, (n_unary > 0 ? n_unary : 1)> facet_ids{};
    mutable std::array<field_container_t, 6> host;
    mutable std::array<bool, 6> fresh{};
    mutable std::vector<double> scratch;
};

}  // namespace lgb200

#endif  // LIBGRAN_LGB200_SYSTEM_H
