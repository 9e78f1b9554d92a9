// unary_force_functor_container: tuple of references to the unary force models (facets) of a system
// (reference: include/libgran/granular_system/unary_force_container.h:17-34).  May be empty, which disables the
// unary stage (reference granular_system.h:47).
#ifndef LIBGRAN_UNARY_FORCE_CONTAINER_H
#define LIBGRAN_UNARY_FORCE_CONTAINER_H

#include <tuple>
#include <vector>

template <typename field_value_t, typename real_t, typename... unary_force_functors_t>
struct unary_force_functor_container {
    typedef std::vector<field_value_t> field_container_t;

    unary_force_functor_container(unary_force_functors_t&... force_functors) : unary_force_functors{force_functors...} {}

    std::tuple<unary_force_functors_t&...> unary_force_functors;
};

#endif  // LIBGRAN_UNARY_FORCE_CONTAINER_H
