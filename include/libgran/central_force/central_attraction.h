// central_attraction_functor: the unary model the reference's tutorial builds (Writerside/topics/Custom-unary-force-model.md:
// 44-131, full listing :466-500) -- a_i = k_a / mass * (x_i - x0), alpha_i = 0, with x0 the centre of mass the driver
// refreshes through update_center_of_mass(x).  Holder: the sum is done on the host in index order, exactly as the tutorial's
// loop (:115-121); the force runs on the GPU (lg_set_central_attraction; kernels_force.cu).
#ifndef LIBGRAN_CENTRAL_ATTRACTION_H
#define LIBGRAN_CENTRAL_ATTRACTION_H

#include <vector>

#include <libgran/backend/lgb200_backend.h>

template <typename field_value_t, typename real_t>
struct central_attraction_functor {
    central_attraction_functor(central_attraction_functor const&) = delete;
    central_attraction_functor(real_t r, real_t k_a, real_t mass, field_value_t zero_field)
        : k_a(k_a), zero_field(zero_field), x0(zero_field) {
        constants.r_part = r; constants.have_r = true;
        constants.mass = mass; constants.have_mass = true;
    }

    void update_center_of_mass(std::vector<field_value_t> const& x) {
        field_value_t x0_new = zero_field;
        for (auto const& point : x) x0_new += point;
        x0 = x0_new / real_t(x.size());
    }

    // ---- backend protocol (unary) ----
    static constexpr bool lgb200_unary_model = true;
    lgb200::particle_constants lgb200_constants() const { return constants; }
    int lgb200_install(lg_system* s) const {
        lgb200_sync(s, -1);
        return -1;
    }
    void lgb200_sync(lg_system* s, int) const {
        double c[3];
        lgb200::pack1(x0, c);
        LGB200_CHECK(lg_set_central_attraction(s, k_a, c));
    }

private:
    double k_a;
    field_value_t zero_field, x0;
    lgb200::particle_constants constants;
};

#endif  // LIBGRAN_CENTRAL_ATTRACTION_H
