// granular_system_b200: the all-pairs granular system on the B200 backend.
// Same template parameters and constructor as the reference's granular_system
// (include/libgran/granular_system/granular_system.h:25-64); a reference driver such as test/sintered.cpp switches
// backend by renaming the type `granular_system<...>` to `granular_system_b200<...>` (or by compiling with
// -DLIBGRAN_B200_DROP_IN, which makes `granular_system` itself name this class).
//
// What differs by design: force functors are parameter holders (their operator() does not exist on the host),
// user-written functors / step handlers are rejected at compile time (no CPU fallback), get_*() return a host
// mirror refreshed on demand, and the pair loop (reference rotational_binary_system_omp.h:69-90) is the CUDA step
// kernel behind lg_do_step.
#ifndef LIBGRAN_GRANULAR_SYSTEM_H
#define LIBGRAN_GRANULAR_SYSTEM_H

#include <libgran/backend/lgb200_system.h>
#include <libgran/granular_system/binary_force_container.h>
#include <libgran/granular_system/unary_force_container.h>
#include <libgran/utility/operator_overloads.h>
#include <libtimestep/rotational_integrator/rotational_velocity_verlet_half.h>
#include <libtimestep/rotational_step_handler/rotational_step_handler.h>

template <typename field_value_t, typename real_t,
          template <typename _field_container_t, typename _field_value_t, typename _real_t, typename _functor_t,
                    template <typename __field_container_t, typename __field_value_t> typename _step_handler_t>
          typename integrator_t,
          template <typename _field_container_t, typename _field_value_t> typename step_handler_t,
          typename binary_force_functor_container_t, typename unary_force_functor_container_t>
class granular_system_b200
    : public lgb200::device_system<field_value_t, real_t, integrator_t, step_handler_t, binary_force_functor_container_t,
                                   unary_force_functor_container_t> {
    typedef lgb200::device_system<field_value_t, real_t, integrator_t, step_handler_t, binary_force_functor_container_t,
                                  unary_force_functor_container_t> base_t;

public:
    typedef std::vector<field_value_t> field_container_t;

    granular_system_b200(field_container_t x0, field_container_t v0, field_container_t theta0, field_container_t omega0,
                         real_t /*t0*/, field_value_t /*field_zero*/, real_t /*real_zero*/,
                         step_handler_t<field_container_t, field_value_t>& step_handler,
                         binary_force_functor_container_t binary_force_functors,
                         unary_force_functor_container_t unary_force_functors)
        : base_t(false, -1, real_t(0), std::move(x0), std::move(v0), std::move(theta0), std::move(omega0), step_handler,
                 std::move(binary_force_functors), std::move(unary_force_functors)) {}
};

#ifdef LIBGRAN_B200_DROP_IN
template <typename field_value_t, typename real_t,
          template <typename, typename, typename, typename, template <typename, typename> typename> typename integrator_t,
          template <typename, typename> typename step_handler_t, typename binary_force_functor_container_t,
          typename unary_force_functor_container_t>
using granular_system = granular_system_b200<field_value_t, real_t, integrator_t, step_handler_t,
                                             binary_force_functor_container_t, unary_force_functor_container_t>;
#endif

#endif  // LIBGRAN_GRANULAR_SYSTEM_H
