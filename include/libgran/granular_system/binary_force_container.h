// binary_force_functor_container: tuple of references to the binary force models of a system, folded left to
// right per ordered pair (reference: include/libgran/granular_system/binary_force_container.h:17-34).  With the
// B200 backend the fold is a compile-time model mask of the step kernel; this class only carries the references.
#ifndef LIBGRAN_BINARY_FORCE_CONTAINER_H
#define LIBGRAN_BINARY_FORCE_CONTAINER_H

#include <tuple>
#include <vector>

template <typename field_value_t, typename real_t, typename... binary_force_functors_t>
struct binary_force_functor_container {
    typedef std::vector<field_value_t> field_container_t;

    binary_force_functor_container(binary_force_functors_t&... force_functors) : binary_force_functors{force_functors...} {}

    std::tuple<binary_force_functors_t&...> binary_force_functors;
};

#endif  // LIBGRAN_BINARY_FORCE_CONTAINER_H
