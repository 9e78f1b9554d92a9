// limited_range_hamaker_functor: hamaker_functor that returns zero beyond a centre distance r_cutoff -- parameter holder.
// Constructor as in the reference (include/libgran/hamaker_force/limited_range_hamaker_force.h:20-27); the law (:29-47,
// cutoff test :34-35) is libgran_b200/csrc/physics.cuh: hamaker_accel(kind = HAMAKER_LIMITED_RANGE).
#ifndef LIBGRAN_LIMITED_RANGE_HAMAKER_FORCE_H
#define LIBGRAN_LIMITED_RANGE_HAMAKER_FORCE_H

#include <libgran/backend/lgb200_backend.h>

template <typename field_value_t, typename real_t>
struct limited_range_hamaker_functor {
    limited_range_hamaker_functor(real_t A, real_t h0, real_t r_cutoff, real_t r_part, real_t mass, field_value_t /*field_zero*/,
                                  real_t /*real_zero*/)
        : A(A), h0(h0), r_cutoff(r_cutoff) {
        constants.r_part = r_part; constants.have_r = true;
        constants.mass = mass; constants.have_mass = true;
    }
    static constexpr bool lgb200_binary_model = true;
    lgb200::particle_constants lgb200_constants() const { return constants; }
    void lgb200_install(lg_system* s) const {
        LGB200_CHECK(lg_set_hamaker(s, A, h0));
        LGB200_CHECK(lg_set_hamaker_variant(s, LG_HAMAKER_LIMITED_RANGE, r_cutoff));
    }

private:
    double A, h0, r_cutoff;
    lgb200::particle_constants constants;
};

#endif  // LIBGRAN_LIMITED_RANGE_HAMAKER_FORCE_H
