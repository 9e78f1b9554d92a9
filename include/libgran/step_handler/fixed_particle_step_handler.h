// fixed_particle_step_handler: the step handler the reference's tutorial builds (Writerside/topics/Custom-step-handler.md:
// 48-146, full listing :340-416) -- increment_v of one particle is dropped, everything else goes to the base step handler.
// On the device this is the `fixed` index of the integrate / half-kick kernels (lg_set_fixed_particle); the class keeps the
// tutorial's constructor and members so that its driver (:176-310) compiles against this repository's headers.
#ifndef LIBGRAN_FIXED_PARTICLE_STEP_HANDLER_H
#define LIBGRAN_FIXED_PARTICLE_STEP_HANDLER_H

#include <cstddef>
#include <utility>

template <typename field_container_t, typename field_value_t,
          template <typename _field_container_t, typename _field_value_t> typename base_step_handler_t>
struct fixed_particle_step_handler {
    typedef base_step_handler_t<field_container_t, field_value_t> base_t;
    static constexpr int lgb200_step_handler_kind = base_t::lgb200_step_handler_kind;
    static constexpr bool lgb200_fixes_a_particle = true;

    fixed_particle_step_handler(std::size_t fixed_particle, base_t base_step_handler)
        : fixed_particle(fixed_particle), base_step_handler(std::move(base_step_handler)) {}

    std::size_t lgb200_fixed_particle() const { return fixed_particle; }

    template <typename... iterators_t>
    void increment_v(std::size_t n, field_value_t const& dv, iterators_t... itrs) {
        if (n == fixed_particle) return;  // Custom-step-handler.md:95-97
        base_step_handler.increment_v(n, dv, itrs...);
    }
    template <typename... iterators_t>
    void increment_x(std::size_t n, field_value_t const& dx, iterators_t... itrs) { base_step_handler.increment_x(n, dx, itrs...); }
    template <typename... iterators_t>
    void increment_theta(std::size_t n, field_value_t const& dtheta, iterators_t... itrs) { base_step_handler.increment_theta(n, dtheta, itrs...); }
    template <typename... iterators_t>
    void increment_omega(std::size_t n, field_value_t const& domega, iterators_t... itrs) { base_step_handler.increment_omega(n, domega, itrs...); }

private:
    const std::size_t fixed_particle;
    base_t base_step_handler;
};

#endif  // LIBGRAN_FIXED_PARTICLE_STEP_HANDLER_H
