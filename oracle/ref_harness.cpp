// ref_harness.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// Drives the UNMODIFIED reference headers (/root/reference/include/libgran/**,
// /root/reference/deps/libtimestep/include/**, /root/reference/deps/eigen) in lock-step behind a
// small C API so the parity tests and bench.py's reference arm can run the reference's own CPU
// implementation of the hot path on arbitrary inputs.  Nothing of the reference is copied into
// this repository: the headers are included from where they lie and the resulting object code is
// written to oracle/_ref/ (git-ignored).  Build recipe: oracle/Makefile.
//
// The driver loop follows test/hamaker_neighbor.cpp:119-124 (rebuild every `rebuild_every` steps,
// then do_step) and test/sintered.cpp:106-134 (sinter functor + its step handler).
//
// One translation unit is compiled per binary-container combination (-DREF_BINARY=0..4) to keep
// compile times parallel:
//   0: contact            1: contact + hamaker      2: sinter
//   3: sinter + hamaker   4: hamaker
//   5: alt_sinter         6: contact + limited_range_hamaker      7: contact + simplified_hamaker
// and each instantiates {all-pairs, neighbor list} x {no unary force, one triangular facet}.
// -DREF_INTEGRATOR=1 builds the same combination with rotational_forward_euler instead of
// rotational_velocity_verlet_half (factory ref_make_binary_fe_<k>).

#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <forward_list>
#include <functional>
#include <memory>
#include <random>
#include <tuple>
#include <utility>
#include <vector>
#include <algorithm>
#include <numeric>
#include <exception>
#include <string>
#include <iterator>
#include <type_traits>
#include <iostream>
#include <omp.h>

#include <Eigen/Eigen>

// The reference keeps the neighbor rows, the spring table and the connector table private with no
// accessors (rotational_binary_system_neighbors_omp.h:118-122, contact_force.h:149-156,
// sinter_bridge.h:261-269).  The harness needs to read them for parity, so the reference headers
// (and only those; every std/Eigen header is already included above) are parsed with the access
// specifiers opened up.  This does not change layout or behaviour.
#define private public
#define protected public
// include order as in test/surface.cpp:20-27 (the pair operator+ of utility/operator_overloads.h
// must be visible before triangular_facet.h; surface_hamaker_force.h needs hamaker_force.h first)
#include <libgran/contact_force/contact_force.h>
#include <libgran/contact_force/surface_contact_force.h>
#if REF_BINARY == 6
#include <libgran/hamaker_force/limited_range_hamaker_force.h>  // defines square / cube itself: cannot be co-included
#else
#include <libgran/hamaker_force/hamaker_force.h>
#endif
#if REF_BINARY == 7
#include <libgran/hamaker_force/simplified_hamaker_force.h>
#endif
#include <libgran/hamaker_force/surface_hamaker_force.h>
#include <libgran/sinter_bridge/sinter_bridge.h>
#include <libgran/sinter_bridge/alt_sinter_bridge.h>
#include <libgran/granular_system/granular_system.h>
#undef binary_system_implementation
#include <libgran/granular_system/granular_system_neighbor_list.h>
#include <libgran/surface_force/triangular_facet.h>
#undef private
#undef protected

#include "dem_config.h"
#include "ref_iface.h"

#ifndef REF_BINARY
#error "compile with -DREF_BINARY=0..7"
#endif
#ifndef REF_INTEGRATOR
#define REF_INTEGRATOR 0
#endif
#if REF_INTEGRATOR == 1
#define REF_INTEGRATOR_T rotational_forward_euler
#else
#define REF_INTEGRATOR_T rotational_velocity_verlet_half
#endif

namespace {

using V3 = Eigen::Vector3d;
using contact_t = contact_force_functor<V3, double>;
#if REF_BINARY == 6
using hamaker_t = limited_range_hamaker_functor<V3, double>;
#elif REF_BINARY == 7
using hamaker_t = simplified_hamaker_functor<V3, double>;
#else
using hamaker_t = hamaker_functor<V3, double>;
#endif
using alt_sinter_t = alt_sinter_functor<V3, double>;
using sinter_t = sinter_functor<V3, double>;
using surf_contact_t = surface_contact_force_functor<V3, double>;
using surf_hamaker_t = surface_hamaker_functor<V3, double>;
using facet_t = triangular_facet<V3, double, surf_contact_t, surf_hamaker_t>;

template <typename C, typename V>
using sinter_step_handler_double = sinter_step_handler<C, V, double>;

#if REF_BINARY == 0
constexpr int kMask = DEM_MODEL_CONTACT;
using binary_t = binary_force_functor_container<V3, double, contact_t>;
#elif REF_BINARY == 1
constexpr int kMask = DEM_MODEL_CONTACT | DEM_MODEL_HAMAKER;
using binary_t = binary_force_functor_container<V3, double, contact_t, hamaker_t>;
#elif REF_BINARY == 2
constexpr int kMask = DEM_MODEL_SINTER;
using binary_t = binary_force_functor_container<V3, double, sinter_t>;
#elif REF_BINARY == 3
constexpr int kMask = DEM_MODEL_SINTER | DEM_MODEL_HAMAKER;
using binary_t = binary_force_functor_container<V3, double, sinter_t, hamaker_t>;
#elif REF_BINARY == 4
constexpr int kMask = DEM_MODEL_HAMAKER;
using binary_t = binary_force_functor_container<V3, double, hamaker_t>;
#elif REF_BINARY == 5
constexpr int kMask = DEM_MODEL_ALT_SINTER;
using binary_t = binary_force_functor_container<V3, double, alt_sinter_t>;
#elif REF_BINARY == 6 || REF_BINARY == 7
constexpr int kMask = DEM_MODEL_CONTACT | DEM_MODEL_HAMAKER;
using binary_t = binary_force_functor_container<V3, double, contact_t, hamaker_t>;
#endif
constexpr bool kSinter = (kMask & DEM_MODEL_SINTER) != 0;
constexpr bool kAlt = (kMask & DEM_MODEL_ALT_SINTER) != 0;

using unary_none_t = unary_force_functor_container<V3, double>;
using unary_facet_t = unary_force_functor_container<V3, double, facet_t>;

template <bool NEIGH, typename unary_t>
struct system_sel;
#if REF_BINARY == 2 || REF_BINARY == 3
template <typename unary_t>
struct system_sel<false, unary_t> {
    using type = granular_system<V3, double, REF_INTEGRATOR_T, sinter_step_handler_double, binary_t, unary_t>;
    using handler = sinter_step_handler_double<std::vector<V3>, V3>;
};
template <typename unary_t>
struct system_sel<true, unary_t> {
    using type = granular_system_neighbor_list<V3, double, REF_INTEGRATOR_T, sinter_step_handler_double, binary_t, unary_t>;
    using handler = sinter_step_handler_double<std::vector<V3>, V3>;
};
#else
template <typename unary_t>
struct system_sel<false, unary_t> {
    using type = granular_system<V3, double, REF_INTEGRATOR_T, rotational_step_handler, binary_t, unary_t>;
    using handler = rotational_step_handler<std::vector<V3>, V3>;
};
template <typename unary_t>
struct system_sel<true, unary_t> {
    using type = granular_system_neighbor_list<V3, double, REF_INTEGRATOR_T, rotational_step_handler, binary_t, unary_t>;
    using handler = rotational_step_handler<std::vector<V3>, V3>;
};
#endif

std::vector<V3> to_vec(const double* p, int64_t n) {
    std::vector<V3> out((size_t)n);
    for (int64_t i = 0; i < n; i++) out[(size_t)i] = V3(p[3 * i], p[3 * i + 1], p[3 * i + 2]);
    return out;
}

template <bool NEIGH, bool FACET>
struct RefImpl final : RefSystem {
    using unary_t = std::conditional_t<FACET, unary_facet_t, unary_none_t>;
    using sel = system_sel<NEIGH, unary_t>;
    using system_t = typename sel::type;
    using handler_t = typename sel::handler;

    dem_config cfg;
    std::mt19937_64 mt;
    std::uniform_real_distribution<double> dist{-1.0, 1.0};
    std::unique_ptr<contact_t> contact;
    std::unique_ptr<hamaker_t> hamaker;
    std::unique_ptr<sinter_t> sinter;
    std::unique_ptr<alt_sinter_t> alt;
    std::unique_ptr<surf_contact_t> surf_contact;
    std::unique_ptr<surf_hamaker_t> surf_hamaker;
    std::unique_ptr<facet_t> facet;
    std::unique_ptr<handler_t> handler;
    std::unique_ptr<system_t> system;
    int64_t step_count = 0;

    // test/sintered.cpp:33-39
    V3 random_unit_vector() {
        V3 vec;
        do {
            vec = {dist(mt), dist(mt), dist(mt)};
        } while (vec.norm() == 0);
        return vec.normalized();
    }

    RefImpl(const dem_config& c, const double* x0, const double* v0, const double* th0, const double* om0)
        : cfg(c), mt(c.rng_seed) {
        const size_t n = (size_t)c.n;
        const double* p = c.contact;
        const V3 zero = V3::Zero();
        std::vector<V3> x = to_vec(x0, c.n), v = to_vec(v0, c.n), th = to_vec(th0, c.n), om = to_vec(om0, c.n);

        if constexpr ((kMask & DEM_MODEL_CONTACT) != 0 || kAlt)
            contact = std::make_unique<contact_t>(n, p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8], p[9], p[10],
                                                  p[11], p[12], p[13], c.r_part, c.mass, c.inertia, p[14], zero, 0.0);
        if constexpr ((kMask & DEM_MODEL_HAMAKER) != 0) {
#if REF_BINARY == 6
            hamaker = std::make_unique<hamaker_t>(c.hamaker_A, c.hamaker_h0, c.hamaker_cutoff, c.r_part, c.mass, zero, 0.0);
#else
            hamaker = std::make_unique<hamaker_t>(c.hamaker_A, c.hamaker_h0, c.r_part, c.mass, zero, 0.0);
#endif
        }
        if constexpr (kAlt) {  // test/alt_sintered.cpp:111-113
            const double* b = c.alt_bond;
            alt = std::make_unique<alt_sinter_t>(n, x, b[0], b[1], b[2], b[3], b[4], b[5], b[6], b[7], c.r_part, c.mass,
                                                 c.inertia, p[14], zero, 0.0, c.critical_separation, *contact);
        }
        if constexpr (kSinter)
            sinter = std::make_unique<sinter_t>(n, c.k_d, c.gamma_d, p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8],
                                                p[9], p[10], p[11], p[12], p[13], c.r_part, c.mass, c.inertia, p[14], zero,
                                                0.0, x.begin(), [this]() { return random_unit_vector(); },
                                                c.critical_separation);
        if constexpr (FACET) {
            const double* s = c.surf_contact;
            surf_contact = std::make_unique<surf_contact_t>(n, s[0], s[1], s[2], s[3], s[4], s[5], s[6], s[7], s[8], s[9],
                                                            s[10], s[11], s[12], s[13], c.r_part, c.mass, c.inertia, s[14],
                                                            zero, 0.0);
            surf_hamaker = std::make_unique<surf_hamaker_t>(c.surf_A, c.surf_h0, c.r_part, c.mass, zero, 0.0);
            const double* f = c.facet_vertices;
            std::tuple<V3, V3, V3> verts{V3(f[0], f[1], f[2]), V3(f[3], f[4], f[5]), V3(f[6], f[7], f[8])};
            facet = std::make_unique<facet_t>(V3(c.facet_v[0], c.facet_v[1], c.facet_v[2]), verts, *surf_contact,
                                              *surf_hamaker);
        }

        if constexpr (kSinter)
            handler = std::make_unique<handler_t>(sinter->template get_step_handler<std::vector<V3>>());
        else
            handler = std::make_unique<handler_t>();

        auto make_binary = [&]() -> binary_t {
#if REF_BINARY == 0
            return binary_t{*contact};
#elif REF_BINARY == 1
            return binary_t{*contact, *hamaker};
#elif REF_BINARY == 2
            return binary_t{*sinter};
#elif REF_BINARY == 3
            return binary_t{*sinter, *hamaker};
#elif REF_BINARY == 4
            return binary_t{*hamaker};
#elif REF_BINARY == 5
            return binary_t{*alt};
#else
            return binary_t{*contact, *hamaker};
#endif
        };
        auto make_unary = [&]() -> unary_t {
            if constexpr (FACET)
                return unary_t{*facet};
            else
                return unary_t{};
        };

        if constexpr (NEIGH)
            system = std::make_unique<system_t>((long)c.n, c.r_verlet, std::move(x), std::move(v), std::move(th),
                                                std::move(om), 0.0, zero, 0.0, *handler, make_binary(), make_unary());
        else
            system = std::make_unique<system_t>(std::move(x), std::move(v), std::move(th), std::move(om), 0.0, zero, 0.0,
                                                *handler, make_binary(), make_unary());
    }

    void update_neighbor_list() override {
        if constexpr (NEIGH) system->update_neighbor_list();
    }

    void do_step(double dt) override { system->do_step(dt); }

    void run(double dt, int64_t n_steps, int64_t rebuild_every) override {
        for (int64_t s = 0; s < n_steps; s++, step_count++) {
            if constexpr (NEIGH)
                if (rebuild_every > 0 && step_count % rebuild_every == 0) system->update_neighbor_list();
            system->do_step(dt);
        }
    }

    void facet_update_positions(double dt) override {
        if constexpr (FACET) facet->update_positions(dt);
    }

    void get_field(int field, double* out) const override {
        const std::vector<V3>* src = nullptr;
        switch (field) {
            case DEM_X: src = &system->get_x(); break;
            case DEM_V: src = &system->get_v(); break;
            case DEM_A: src = &system->get_a(); break;
            case DEM_THETA: src = &system->get_theta(); break;
            case DEM_OMEGA: src = &system->get_omega(); break;
            case DEM_ALPHA: src = &system->get_alpha(); break;
            default: return;
        }
        std::memcpy(out, src->data(), src->size() * sizeof(V3));
    }

    int64_t n_pairs() const override {
        if constexpr (NEIGH) {
            int64_t p = 0;
            for (auto const& row : system->neighbor_list) p += (int64_t)row.size();
            return p;
        } else {
            return cfg.n * (cfg.n - 1);
        }
    }

    void get_rows(int64_t* row_ptr, int64_t* cols) const override {
        int64_t p = 0;
        for (int64_t i = 0; i < cfg.n; i++) {
            row_ptr[i] = p;
            if constexpr (NEIGH) {
                for (long j : system->neighbor_list[(size_t)i]) cols[p++] = j;
            } else {
                for (int64_t j = 0; j < cfg.n; j++)
                    if (j != i) cols[p++] = j;
            }
        }
        row_ptr[cfg.n] = p;
    }

    // Hidden per-ordered-pair contact springs (contact_force.h:156): 9 doubles for pair (i, j).
    void get_contact_springs(int64_t i, int64_t j, double* out9) const override {
        const contact_t* cf = nullptr;
        if constexpr (kAlt) {  // bonded pairs keep their springs in the alt functor's own table (alt_sinter_bridge.h:146)
            if (alt->bonded_contacts[(size_t)(i * cfg.n + j)]) {
                auto const& rec = alt->contact_springs[(size_t)(i * cfg.n + j)];
                std::memcpy(out9 + 0, std::get<0>(rec).data(), 24);
                std::memcpy(out9 + 3, std::get<1>(rec).data(), 24);
                std::memcpy(out9 + 6, std::get<2>(rec).data(), 24);
                return;
            }
            cf = &alt->contact_force;
        } else if constexpr (kSinter)
            cf = &sinter->contact_force;
        else if constexpr ((kMask & DEM_MODEL_CONTACT) != 0)
            cf = contact.get();
        if (!cf) {
            std::fill(out9, out9 + 9, 0.0);
            return;
        }
        auto const& rec = cf->contact_springs[(size_t)(i * cfg.n + j)];
        std::memcpy(out9 + 0, std::get<0>(rec).data(), 24);
        std::memcpy(out9 + 3, std::get<1>(rec).data(), 24);
        std::memcpy(out9 + 6, std::get<2>(rec).data(), 24);
    }

    // Bond table of the sinter functor (sinter_bridge.h:139-141): returns 1 if (i, j) is bonded and
    // copies the 5 connectors of i in bond (i, j) and the 5 rest lengths.
    int get_bond(int64_t i, int64_t j, double* conn15, double* len5) const override {
        if constexpr (kSinter) {
            size_t idx = (size_t)(i * cfg.n + j);
            if (!sinter->bonded_contacts[idx]) return 0;
            for (int k = 0; k < 5; k++) std::memcpy(conn15 + 3 * k, sinter->spring_connectors[idx][(size_t)k].data(), 24);
            std::memcpy(len5, sinter->spring_lengths[idx].data(), 40);
            return 1;
        } else if constexpr (kAlt) {  // no connectors in this model: only the flag
            std::fill(conn15, conn15 + 15, 0.0);
            std::fill(len5, len5 + 5, 0.0);
            return alt->bonded_contacts[(size_t)(i * cfg.n + j)] ? 1 : 0;
        } else {
            (void)i; (void)j; (void)conn15; (void)len5;
            return 0;
        }
    }

    void get_facet_vertices(double* out9) const override {
        if constexpr (FACET) {
            auto const& vt = facet->get_vertices();
            std::memcpy(out9 + 0, std::get<0>(vt).data(), 24);
            std::memcpy(out9 + 3, std::get<1>(vt).data(), 24);
            std::memcpy(out9 + 6, std::get<2>(vt).data(), 24);
        } else {
            std::fill(out9, out9 + 9, 0.0);
        }
    }
};

}  // namespace

#define REF_CONCAT2(a, b) a##b
#define REF_CONCAT(a, b) REF_CONCAT2(a, b)

// One factory per translation unit; ref_api.cpp dispatches on the binary-model mask.
#if REF_INTEGRATOR == 1
#define REF_FACTORY REF_CONCAT(ref_make_binary_fe_, REF_BINARY)
#else
#define REF_FACTORY REF_CONCAT(ref_make_binary_, REF_BINARY)
#endif
RefSystem* REF_FACTORY(const dem_config* c, const double* x0, const double* v0,
                                                     const double* th0, const double* om0) {
    const bool facet = (c->model_mask & DEM_MODEL_FACET) != 0;
    if (c->use_neighbor_list) {
        if (facet) return new RefImpl<true, true>(*c, x0, v0, th0, om0);
        return new RefImpl<true, false>(*c, x0, v0, th0, om0);
    }
    if (facet) return new RefImpl<false, true>(*c, x0, v0, th0, om0);
    return new RefImpl<false, false>(*c, x0, v0, th0, om0);
}
