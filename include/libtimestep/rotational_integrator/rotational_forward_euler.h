// rotational_forward_euler: integrator selector with the reference's template signature
// (reference: deps/libtimestep/include/libtimestep/rotational_integrator/rotational_forward_euler.h:15-42).
// The scheme (:44-62: update_acceleration(); x += v dt + 0.5 a dt dt with v copied before its own increment; v += a dt;
// theta += omega dt + 0.5 alpha dt dt; omega += alpha dt) is executed on the device by lg_do_step after
// lg_set_integrator(LG_INTEGRATOR_FORWARD_EULER) (libgran_b200/csrc/kernels_integrate.cu: euler_particles_kernel).
#ifndef LIBTIMESTEP_ROTATIONAL_FORWARD_EULER_H
#define LIBTIMESTEP_ROTATIONAL_FORWARD_EULER_H

template <typename field_container_t, typename field_value_t, typename real_t, typename functor_t,
          template <typename _field_container_t, typename _field_value_t> typename step_handler_t>
class rotational_forward_euler {
public:
    static constexpr int lgb200_integrator_kind = 1;  // LG_INTEGRATOR_FORWARD_EULER
};

#endif  // LIBTIMESTEP_ROTATIONAL_FORWARD_EULER_H
