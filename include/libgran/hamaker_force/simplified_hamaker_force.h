// simplified_hamaker_functor: F = 2 A r / h^2 with saturation distance h0 -- parameter holder.
// Constructor as in the reference (include/libgran/hamaker_force/simplified_hamaker_force.h:10-16); the law (:18-31) is
// libgran_b200/csrc/physics.cuh: hamaker_accel(kind = HAMAKER_SIMPLIFIED).
#ifndef LIBGRAN_SIMPLIFIED_HAMAKER_FORCE_H
#define LIBGRAN_SIMPLIFIED_HAMAKER_FORCE_H

#include <libgran/backend/lgb200_backend.h>

template <typename field_value_t, typename real_t>
struct simplified_hamaker_functor {
    simplified_hamaker_functor(real_t A, real_t h0, real_t r_part, real_t mass, field_value_t /*field_zero*/, real_t /*real_zero*/)
        : A(A), h0(h0) {
        constants.r_part = r_part; constants.have_r = true;
        constants.mass = mass; constants.have_mass = true;
    }
    static constexpr bool lgb200_binary_model = true;
    lgb200::particle_constants lgb200_constants() const { return constants; }
    void lgb200_install(lg_system* s) const {
        LGB200_CHECK(lg_set_hamaker(s, A, h0));
        LGB200_CHECK(lg_set_hamaker_variant(s, LG_HAMAKER_SIMPLIFIED, 0.0));
    }

private:
    double A, h0;
    lgb200::particle_constants constants;
};

#endif  // LIBGRAN_SIMPLIFIED_HAMAKER_FORCE_H
