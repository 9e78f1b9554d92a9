// hamaker_functor: sphere-sphere van der Waals attraction with saturation distance h0 -- parameter holder.
// Constructor as in the reference (include/libgran/hamaker_force/hamaker_force.h:20-26); the law (:28-43) is
// libgran_b200/csrc/physics.cuh: hamaker_accel.
#ifndef LIBGRAN_HAMAKER_FORCE_H
#define LIBGRAN_HAMAKER_FORCE_H

#include <libgran/backend/lgb200_backend.h>

template <typename field_value_t, typename real_t>
struct hamaker_functor {
    hamaker_functor(real_t A, real_t h0, real_t r_part, real_t mass, field_value_t /*field_zero*/, real_t /*real_zero*/)
        : A(A), h0(h0) {
        constants.r_part = r_part; constants.have_r = true;
        constants.mass = mass; constants.have_mass = true;
    }

    static constexpr bool lgb200_binary_model = true;
    lgb200::particle_constants lgb200_constants() const { return constants; }
    void lgb200_install(lg_system* s) const { LGB200_CHECK(lg_set_hamaker(s, A, h0)); }

private:
    double A, h0;
    lgb200::particle_constants constants;
};

#endif  // LIBGRAN_HAMAKER_FORCE_H
