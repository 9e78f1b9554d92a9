// rotational_step_handler: the default step handler of libtimestep
// (reference: deps/libtimestep/include/libtimestep/rotational_step_handler/rotational_step_handler.h:19-68).
// With the B200 backend the four increments run inside the fused step kernel
// (libgran_b200/csrc/kernels_force_tiled.cu, finalize stage); this class keeps the reference's members so
// that user code which calls them on its own buffers still compiles, and tags the handler for the backend.
#ifndef LIBTIMESTEP_ROTATIONAL_STEP_HANDLER_H
#define LIBTIMESTEP_ROTATIONAL_STEP_HANDLER_H

#include <cstddef>

namespace lgb200 {
// what the device integrate stage has to do besides x, v, theta, omega
enum step_handler_kind : int { STEP_PLAIN = 0, STEP_SINTER_CONNECTORS = 1 };
}  // namespace lgb200

template <typename field_container_t, typename field_value_t>
struct rotational_step_handler {
    static constexpr int lgb200_step_handler_kind = lgb200::STEP_PLAIN;

    void increment_x(std::size_t n, field_value_t const& dx, typename field_container_t::iterator x_begin_itr,
                     typename field_container_t::const_iterator, typename field_container_t::const_iterator,
                     typename field_container_t::const_iterator, typename field_container_t::const_iterator,
                     typename field_container_t::const_iterator) const {
        *(x_begin_itr + n) += dx;
    }
    void increment_v(std::size_t n, field_value_t const& dv, typename field_container_t::const_iterator,
                     typename field_container_t::iterator v_begin_itr, typename field_container_t::const_iterator,
                     typename field_container_t::const_iterator, typename field_container_t::const_iterator,
                     typename field_container_t::const_iterator) const {
        *(v_begin_itr + n) += dv;
    }
    void increment_theta(std::size_t n, field_value_t const& dtheta, typename field_container_t::const_iterator,
                         typename field_container_t::const_iterator, typename field_container_t::const_iterator,
                         typename field_container_t::iterator theta_begin_itr, typename field_container_t::const_iterator,
                         typename field_container_t::const_iterator) const {
        *(theta_begin_itr + n) += dtheta;
    }
    void increment_omega(std::size_t n, field_value_t const& domega, typename field_container_t::const_iterator,
                         typename field_container_t::const_iterator, typename field_container_t::const_iterator,
                         typename field_container_t::const_iterator, typename field_container_t::iterator omega_begin_itr,
                         typename field_container_t::const_iterator) const {
        *(omega_begin_itr + n) += domega;
    }
};

#endif  // LIBTIMESTEP_ROTATIONAL_STEP_HANDLER_H
