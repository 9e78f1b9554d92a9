// granular_system_neighbor_list_b200: the Verlet-list granular system on the B200 backend.
// Same template parameters and constructor as the reference's granular_system_neighbor_list
// (include/libgran/granular_system/granular_system_neighbor_list.h:25-64: n_part and r_verlet come first);
// test/hamaker_neighbor.cpp switches backend by renaming the type (or with -DLIBGRAN_B200_DROP_IN).
// update_neighbor_list() (reference rotational_binary_system_neighbors_omp.h:102-116, O(N^2)) is a cell-list
// rebuild on the device that yields the same rows: { j != i : |x_i - x_j| < r_verlet }, ascending j.  As in the
// reference the list is empty until the first call (:66) and is never refreshed automatically.
#ifndef LIBGRAN_GRANULAR_SYSTEM_NEIGHBOR_LIST_H
#define LIBGRAN_GRANULAR_SYSTEM_NEIGHBOR_LIST_H

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
class granular_system_neighbor_list_b200
    : public lgb200::device_system<field_value_t, real_t, integrator_t, step_handler_t, binary_force_functor_container_t,
                                   unary_force_functor_container_t> {
    typedef lgb200::device_system<field_value_t, real_t, integrator_t, step_handler_t, binary_force_functor_container_t,
                                  unary_force_functor_container_t> base_t;

public:
    typedef std::vector<field_value_t> field_container_t;

    granular_system_neighbor_list_b200(long n_part, real_t r_verlet, field_container_t x0, field_container_t v0,
                                       field_container_t theta0, field_container_t omega0, real_t /*t0*/,
                                       field_value_t /*field_zero*/, real_t /*real_zero*/,
                                       step_handler_t<field_container_t, field_value_t>& step_handler,
                                       binary_force_functor_container_t binary_force_functors,
                                       unary_force_functor_container_t unary_force_functors)
        : base_t(true, n_part, r_verlet, std::move(x0), std::move(v0), std::move(theta0), std::move(omega0), step_handler,
                 std::move(binary_force_functors), std::move(unary_force_functors)) {}

    void update_neighbor_list() {
        LGB200_CHECK(lg_update_neighbor_list(this->handle));
        this->invalidate();  // a pending closing evaluation is flushed by the rebuild; a / alpha mirrors may be stale
    }

    // the rows of the current list flattened to CSR (the reference keeps them private, :118-122)
    void get_neighbor_rows(std::vector<std::int64_t>& row_ptr, std::vector<std::int64_t>& cols) const {
        std::int64_t pairs = 0;
        LGB200_CHECK(lg_get_neighbor_count(this->handle, &pairs));
        row_ptr.assign(static_cast<std::size_t>(this->n) + 1, 0);
        cols.assign(static_cast<std::size_t>(pairs), 0);
        LGB200_CHECK(lg_get_neighbor_rows(this->handle, row_ptr.data(), cols.data()));
    }
};

#ifdef LIBGRAN_B200_DROP_IN
template <typename field_value_t, typename real_t,
          template <typename, typename, typename, typename, template <typename, typename> typename> typename integrator_t,
          template <typename, typename> typename step_handler_t, typename binary_force_functor_container_t,
          typename unary_force_functor_container_t>
using granular_system_neighbor_list =
    granular_system_neighbor_list_b200<field_value_t, real_t, integrator_t, step_handler_t, binary_force_functor_container_t,
                                       unary_force_functor_container_t>;
#endif

#endif  // LIBGRAN_GRANULAR_SYSTEM_NEIGHBOR_LIST_H
