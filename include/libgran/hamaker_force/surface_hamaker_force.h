// surface_hamaker_functor: particle - half-space van der Waals attraction -- parameter holder
// (reference constructor: include/libgran/hamaker_force/surface_hamaker_force.h:10-15; law :18-30 ->
// libgran_b200/csrc/physics.cuh: wall_hamaker_accel).  Used as a member of triangular_facet.
#ifndef LIBGRAN_SURFACE_HAMAKER_FORCE_H
#define LIBGRAN_SURFACE_HAMAKER_FORCE_H

#include <libgran/backend/lgb200_backend.h>

template <typename field_value_t, typename real_t>
struct surface_hamaker_functor {
    surface_hamaker_functor(real_t A, real_t h0, real_t r_part, real_t mass, field_value_t /*field_zero*/, real_t /*real_zero*/)
        : A(A), h0(h0) {
        constants.r_part = r_part; constants.have_r = true;
        constants.mass = mass; constants.have_mass = true;
    }

    static constexpr bool lgb200_surface_hamaker = true;
    lgb200::particle_constants lgb200_constants() const { return constants; }
    double hamaker_constant() const { return A; }
    double saturation_distance() const { return h0; }

private:
    double A, h0;
    lgb200::particle_constants constants;
};

#endif  // LIBGRAN_SURFACE_HAMAKER_FORCE_H
