// Driver over the header templates of the SURVEY.md section 8f variants, written the way the reference's tutorials and
// test/alt_sintered.cpp write theirs: reads a case file (tests/cpp/case_io.h), builds the functors named by `mode`, runs
// the driver loop, writes the six state vectors.  tests/test_cpp_host.py compares them with the CPU oracle bit for bit.
//   variants_driver case.bin out.bin euler                      rotational_forward_euler, contact + Hamaker
//   variants_driver case.bin out.bin fixed <index>              fixed_particle_step_handler over rotational_step_handler
//   variants_driver case.bin out.bin limited <r_cutoff>         contact + limited_range_hamaker_functor
//   variants_driver case.bin out.bin simplified                 contact + simplified_hamaker_functor
//   variants_driver case.bin out.bin alt <8 bond coefficients>  alt_sinter_functor (test/alt_sintered.cpp:103-127)
//   variants_driver case.bin out.bin central <k_a>              contact + Hamaker + central_attraction_functor (tutorial 3)
#include <cstdlib>
#include <cstring>
#include <string>

#include "case_io.h"

#include <libgran/central_force/central_attraction.h>
#include <libgran/contact_force/contact_force.h>
#include <libgran/granular_system/granular_system.h>
#include <libgran/granular_system/granular_system_neighbor_list.h>
#include <libgran/hamaker_force/hamaker_force.h>
#include <libgran/hamaker_force/limited_range_hamaker_force.h>
#include <libgran/hamaker_force/simplified_hamaker_force.h>
#include <libgran/sinter_bridge/alt_sinter_bridge.h>
#include <libgran/step_handler/fixed_particle_step_handler.h>
#include <libtimestep/rotational_integrator/rotational_forward_euler.h>
#include <libtimestep/rotational_integrator/rotational_velocity_verlet_half.h>

using field_container_t = std::vector<vec3>;
using contact_t = contact_force_functor<vec3, double>;
using hamaker_t = hamaker_functor<vec3, double>;
using empty_unary_t = unary_force_functor_container<vec3, double>;

template <typename field_container_t, typename field_value_t>
using fixed_handler_t = fixed_particle_step_handler<field_container_t, field_value_t, rotational_step_handler>;

template <typename System>
static void loop(Case const& c, System& sys) {
    for (int64_t k = 0; k < c.n_steps; k++) {
        if constexpr (requires { sys.update_neighbor_list(); })
            if (c.rebuild_every > 0 && k % c.rebuild_every == 0) sys.update_neighbor_list();
        sys.do_step(c.dt);
    }
}

template <template <typename, typename, typename, typename, template <typename, typename> typename> typename Integrator,
          template <typename, typename> typename Handler, typename BinaryC, typename UnaryC>
static void run(Case const& c, Handler<field_container_t, vec3>& handler, BinaryC binary, UnaryC unary, const char* out) {
    if (c.use_nl) {
        granular_system_neighbor_list_b200<vec3, double, Integrator, Handler, BinaryC, UnaryC> sys(
            (long)c.n, c.r_verlet, c.x, c.v, c.th, c.w, 0.0, vec3::Zero(), 0.0, handler, binary, unary);
        loop(c, sys);
        dump(out, sys);
    } else {
        granular_system_b200<vec3, double, Integrator, Handler, BinaryC, UnaryC> sys(c.x, c.v, c.th, c.w, 0.0, vec3::Zero(), 0.0, handler,
                                                                                     binary, unary);
        loop(c, sys);
        dump(out, sys);
    }
}

int main(int argc, char** argv) {
    if (argc < 4) {
        fprintf(stderr, "usage: variants_driver case.bin out.bin mode [args]\n");
        return 2;
    }
    try {
        const Case c = load(argv[1]);
        const char* out = argv[2];
        const std::string mode = argv[3];
        const double* p = c.contact;
        contact_t contact(c.n, p[0], p[1], p[2], p[3], p[4], p[5], p[6], p[7], p[8], p[9], p[10], p[11], p[12], p[13], c.r_part, c.mass,
                          c.inertia, p[14], vec3::Zero(), 0.0);
        hamaker_t hamaker(c.A, c.h0, c.r_part, c.mass, vec3::Zero(), 0.0);
        rotational_step_handler<field_container_t, vec3> plain;
        empty_unary_t no_unary;
        if (mode == "euler") {
            binary_force_functor_container<vec3, double, contact_t, hamaker_t> binary{contact, hamaker};
            run<rotational_forward_euler, rotational_step_handler>(c, plain, binary, no_unary, out);
        } else if (mode == "fixed" && argc > 4) {
            fixed_handler_t<field_container_t, vec3> handler((size_t)std::atol(argv[4]), plain);
            binary_force_functor_container<vec3, double, contact_t, hamaker_t> binary{contact, hamaker};
            run<rotational_velocity_verlet_half, fixed_handler_t>(c, handler, binary, no_unary, out);
        } else if (mode == "limited" && argc > 4) {
            limited_range_hamaker_functor<vec3, double> lim(c.A, c.h0, std::atof(argv[4]), c.r_part, c.mass, vec3::Zero(), 0.0);
            binary_force_functor_container<vec3, double, contact_t, limited_range_hamaker_functor<vec3, double>> binary{contact, lim};
            run<rotational_velocity_verlet_half, rotational_step_handler>(c, plain, binary, no_unary, out);
        } else if (mode == "simplified") {
            simplified_hamaker_functor<vec3, double> simp(c.A, c.h0, c.r_part, c.mass, vec3::Zero(), 0.0);
            binary_force_functor_container<vec3, double, contact_t, simplified_hamaker_functor<vec3, double>> binary{contact, simp};
            run<rotational_velocity_verlet_half, rotational_step_handler>(c, plain, binary, no_unary, out);
        } else if (mode == "alt" && argc > 11) {
            double b[8];
            for (int k = 0; k < 8; k++) b[k] = std::atof(argv[4 + k]);
            alt_sinter_functor<vec3, double> alt(c.n, c.x, b[0], b[1], b[2], b[3], b[4], b[5], b[6], b[7], c.r_part, c.mass, c.inertia,
                                                 p[14], vec3::Zero(), 0.0, c.crit, contact);
            binary_force_functor_container<vec3, double, alt_sinter_functor<vec3, double>> binary{alt};
            run<rotational_velocity_verlet_half, rotational_step_handler>(c, plain, binary, no_unary, out);
        } else if (mode == "central" && argc > 4) {
            central_attraction_functor<vec3, double> central(c.r_part, std::atof(argv[4]), c.mass, vec3::Zero());
            central.update_center_of_mass(c.x);
            binary_force_functor_container<vec3, double, contact_t, hamaker_t> binary{contact, hamaker};
            unary_force_functor_container<vec3, double, central_attraction_functor<vec3, double>> unary{central};
            run<rotational_velocity_verlet_half, rotational_step_handler>(c, plain, binary, unary, out);
        } else {
            fprintf(stderr, "variants_driver: unknown mode %s\n", mode.c_str());
            return 2;
        }
    } catch (std::exception const& e) {
        fprintf(stderr, "variants_driver: %s\n", e.what());
        return 1;
    }
    return 0;
}
