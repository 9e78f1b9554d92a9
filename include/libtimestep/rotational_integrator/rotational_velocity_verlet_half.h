// rotational_velocity_verlet_half: integrator selector with the reference's template signature
// (reference: deps/libtimestep/include/libtimestep/rotational_integrator/rotational_velocity_verlet_half.h:15-43),
// so that `granular_system_b200<V, double, rotational_velocity_verlet_half, ...>` reads like the reference.
// The scheme itself (first-call half kick back :45-60, v += a dt; x += v dt; omega += alpha dt; theta += omega dt
// :63-73, closing force evaluation :75) is executed on the device by lg_do_step (libgran_b200/csrc/api.cu: step()).
#ifndef LIBTIMESTEP_ROTATIONAL_VELOCITY_VERLET_HALF_H
#define LIBTIMESTEP_ROTATIONAL_VELOCITY_VERLET_HALF_H

namespace lgb200 {
enum integrator_kind : int { INTEGRATOR_VELOCITY_VERLET_HALF = 0 };
}  // namespace lgb200

template <typename field_container_t, typename field_value_t, typename real_t, typename functor_t,
          template <typename _field_container_t, typename _field_value_t> typename step_handler_t>
class rotational_velocity_verlet_half {
public:
    static constexpr int lgb200_integrator_kind = lgb200::INTEGRATOR_VELOCITY_VERLET_HALF;
};

#endif  // LIBTIMESTEP_ROTATIONAL_VELOCITY_VERLET_HALF_H
