// Generic driver over the header-template API (include/libgran/**): reads a case file written by
// tests/test_cpp_host.py, builds functors / containers / step handler / system exactly as a libgran driver
// program does (cf. reference test/sintered.cpp:106-127, test/hamaker_neighbor.cpp:87-117, test/surface.cpp:93-111),
// runs the reference's driver loop and writes the six state vectors.  The Python test compares them with the
// CPU oracle bit for bit.
//
// case file (little-endian doubles unless noted):
//   int64 n, model_mask, use_nl, n_steps, rebuild_every, move_facets, n_unit_vectors
//   double r_part, mass, inertia, r_verlet, dt, contact[15], hamaker_A, hamaker_h0, k_d, gamma_d, crit,
//          facet_vertices[9], facet_v[3], surf_contact[15], surf_A, surf_h0
//   double x[3n], v[3n], theta[3n], omega[3n], unit_vectors[3 * n_unit_vectors]
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <stdexcept>
#include <vector>

#include "vec3.h"

#include <libgran/contact_force/contact_force.h>
#include <libgran/contact_force/surface_contact_force.h>
#include <libgran/granular_system/granular_system.h>
#include <libgran/granular_system/granular_system_neighbor_list.h>
#include <libgran/hamaker_force/hamaker_force.h>
#include <libgran/hamaker_force/surface_hamaker_force.h>
#include <libgran/sinter_bridge/sinter_bridge.h>
#include <libgran/surface_force/triangular_facet.h>

#include "case_io.h"

using contact_t = contact_force_functor<vec3, double>;
using hamaker_t = hamaker_functor<vec3, double>;
using sinter_t = sinter_functor<vec3, double>;
using facet_t = triangular_facet<vec3, double, surface_contact_force_functor<vec3, double>, surface_hamaker_functor<vec3, double>>;
template <typename C, typename V>
using sinter_step_handler_double = sinter_step_handler<C, V, double>;

// the driver loop of test/hamaker_neighbor.cpp:120-124 (+ facet motion as a driver would add it)
template <bool NL, typename System, typename... Facets>
static void drive(Case const& c, System& sys, Facets&... facets) {
    for (int64_t k = 0; k < c.n_steps; k++) {
        if constexpr (NL)
            if (c.rebuild_every > 0 && k % c.rebuild_every == 0) sys.update_neighbor_list();
        sys.do_step(c.dt);
        if (c.move_facets) (facets.update_positions(c.dt), ...);
    }
}

template <bool NL, template <typename, typename> typename Handler, typename BinaryC, typename UnaryC, typename... Facets>
static void build_and_run(Case const& c, Handler<std::vector<vec3>, vec3>& handler, BinaryC binary, UnaryC unary, const char* out,
                          Facets&... facets) {
    if constexpr (NL) {
        granular_system_neighbor_list_b200<vec3, double, rotational_velocity_verlet_half, Handler, BinaryC, UnaryC> sys(
            (long)c.n, c.r_verlet, c.x, c.v, c.th, c.w, 0.0, vec3::Zero(), 0.0, handler, binary, unary);
        drive<true>(c, sys, facets...);
        dump(out, sys);
    } else {
        granular_system_b200<vec3, double, rotational_velocity_verlet_half, Handler, BinaryC, UnaryC> sys(
            c.x, c.v, c.th, c.w, 0.0, vec3::Zero(), 0.0, handler, binary, unary);
        drive<false>(c, sys, facets...);
        dump(out, sys);
    }
}

template <bool NL, template <typename, typename> typename Handler, typename... Binary>
static void with_unary(Case const& c, Handler<std::vector<vec3>, vec3>& handler, const char* out, Binary&... b) {
    binary_force_functor_container<vec3, double, Binary...> binary{b...};
    if (c.mask & M_FACET) {
        const double* s = c.sc;
        surface_contact_force_functor<vec3, double> sc(c.n, s[0], s[1], s[2], s[3], s[4], s[5], s[6], s[7], s[8], s[9], s[10], s[11],
                                                       s[12], s[13], c.r_part, c.mass, c.inertia, s[14], vec3::Zero(), 0.0);
        surface_hamaker_functor<vec3, double> sh(c.sA, c.sh0, c.r_part, c.mass, vec3::Zero(), 0.0);
        facet_t facet(vec3(c.fvel[0], c.fvel[1], c.fvel[2]),
                      {vec3(c.fv[0], c.fv[1], c.fv[2]), vec3(c.fv[3], c.fv[4], c.fv[5]), vec3(c.fv[6], c.fv[7], c.fv[8])}, sc, sh);
        unary_force_functor_container<vec3, double, facet_t> unary{facet};
        build_and_run<NL, Handler>(c, handler, binary, unary, out, facet);
    } else {
        unary_force_functor_container<vec3, double> unary;
        build_and_run<NL, Handler>(c, handler, binary, unary, out);
    }
}

template <bool NL>
static void with_binary(Case const& c, const char* out) {
    const double* p = c.contact;
    if (c.mask & M_SINTER) {
        size_t next = 0;
        std::function<vec3()> gen = [&]() {
            if (next >= c.uv.size()) throw std::runtime_error("ran out of unit vectors");
            return c.uv[next++];
        };
        sinter_t sinter(c.n, c.k_d, c.gamma_d, p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8], p[9], p[10], p[11], p[12], p[13],
                        c.r_part, c.mass, c.inertia, p[14], vec3::Zero(), 0.0, c.x.begin(), gen, c.crit);
        auto handler = sinter.get_step_handler<std::vector<vec3>>();
        if (c.mask & M_HAMAKER) {
            hamaker_t hamaker(c.A, c.h0, c.r_part, c.mass, vec3::Zero(), 0.0);
            with_unary<NL, sinter_step_handler_double>(c, handler, out, sinter, hamaker);
        } else {
            with_unary<NL, sinter_step_handler_double>(c, handler, out, sinter);
        }
    } else {
        contact_t contact(c.n, p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8], p[9], p[10], p[11], p[12], p[13], c.r_part,
                          c.mass, c.inertia, p[14], vec3::Zero(), 0.0);
        rotational_step_handler<std::vector<vec3>, vec3> handler;
        if (c.mask & M_HAMAKER) {
            hamaker_t hamaker(c.A, c.h0, c.r_part, c.mass, vec3::Zero(), 0.0);
            with_unary<NL, rotational_step_handler>(c, handler, out, contact, hamaker);
        } else {
            with_unary<NL, rotational_step_handler>(c, handler, out, contact);
        }
    }
}

int main(int argc, char** argv) {
    if (argc < 3) {
        fprintf(stderr, "usage: template_driver case.bin out.bin | --size-mismatch\n");
        return 2;
    }
    try {
        Case c = load(argv[1]);
        if (argc > 3 && std::string(argv[3]) == "--size-mismatch") {
            c.v.pop_back();  // the constructor must throw SizeMismatchException (rotational_system.h:69-70)
            try {
                with_binary<false>(c, argv[2]);
            } catch (SizeMismatchException const& e) {
                printf("SizeMismatchException: %s\n", e.what());
                return 0;
            }
            return 3;
        }
        if (c.use_nl) with_binary<true>(c, argv[2]); else with_binary<false>(c, argv[2]);
    } catch (std::exception const& e) {
        fprintf(stderr, "template_driver: %s\n", e.what());
        return 1;
    }
    return 0;
}
