// triangular_facet: unary force model holding surface force functors -- host-side holder of a (moving) triangle.
// Public surface as in the reference (include/libgran/surface_force/triangular_facet.h:18-69): constructor
// (v_facet, vertices, functors...), public v_facet, update_positions(dt), get_vertices(), get_unit_normal().
// Closest-point search and the surface force fold (:32-51, :73-137) run on the GPU
// (libgran_b200/csrc/physics.cuh: facet_closest_point; kernels_force_tiled.cu: unary_forces).  As in the
// reference the unit normal and the area are fixed at construction (:21-28) -- update_positions only translates.
#ifndef LIBGRAN_TRIANGULAR_FACET_H
#define LIBGRAN_TRIANGULAR_FACET_H

#include <cmath>
#include <tuple>
#include <utility>
#include <vector>

#include <libgran/backend/lgb200_backend.h>

template <typename field_value_t, typename real_t, typename... surface_force_functors_t>
struct triangular_facet {
    typedef std::vector<field_value_t> field_container_t;

    explicit triangular_facet(field_value_t v_facet, std::tuple<field_value_t, field_value_t, field_value_t> vertices,
                              surface_force_functors_t&... force_functors)
        : v_facet(std::move(v_facet)), vertices(std::move(vertices)), surface_force_functors{force_functors...} {
        static_assert(sizeof...(surface_force_functors_t) > 0, "at least one surface force functor must be provided");
        // unit normal for get_unit_normal(); the device recomputes it from the vertices it is given at registration
        double a[3], b[3], c[3];
        lgb200::pack1(std::get<0>(this->vertices), a);
        lgb200::pack1(std::get<1>(this->vertices), b);
        lgb200::pack1(std::get<2>(this->vertices), c);
        const double u[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, w[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
        double n[3] = {u[1] * w[2] - u[2] * w[1], u[2] * w[0] - u[0] * w[2], u[0] * w[1] - u[1] * w[0]};
        const double z = (n[0] * n[0] + n[1] * n[1]) + n[2] * n[2];
        if (z > 0.0) {
            const double len = std::sqrt(z);
            for (double& q : n) q /= len;
        }
        plane_unit_normal = field_value_t(n[0], n[1], n[2]);
    }

    field_value_t const& get_unit_normal() const { return plane_unit_normal; }
    std::tuple<field_value_t, field_value_t, field_value_t> const& get_vertices() const { return vertices; }

    // Move the facet according to its velocity (reference :62-66); the owning system pushes the new vertices
    // to the device before its next step.
    void update_positions(real_t dt) {
        std::get<0>(vertices) += v_facet * dt;
        std::get<1>(vertices) += v_facet * dt;
        std::get<2>(vertices) += v_facet * dt;
    }

    field_value_t v_facet;  // velocity of the facet (public data member in the reference, :69)

    // ---- backend protocol ----
    static constexpr bool lgb200_unary_model = true;
    lgb200::particle_constants lgb200_constants() const {
        lgb200::particle_constants pc;
        std::apply([&pc](auto&... f) { (pc.merge(f.lgb200_constants()), ...); }, surface_force_functors);
        return pc;
    }
    // registers the facet with a system; returns the facet id
    int lgb200_install(lg_system* s) const {
        const lg_contact_params* sc = nullptr;
        int has_hamaker = 0;
        double A = 0.0, h0 = 0.0;
        std::apply(
            [&](auto&... f) {
                (
                    [&](auto& g) {
                        using T = std::decay_t<decltype(g)>;
                        if constexpr (requires { T::lgb200_surface_contact; }) {
                            sc = &g.params();
                        } else if constexpr (requires { T::lgb200_surface_hamaker; }) {
                            has_hamaker = 1;
                            A = g.hamaker_constant();
                            h0 = g.saturation_distance();
                        } else {
                            static_assert(sizeof(T) == 0,
                                          "libgran-b200: this surface force functor has no device implementation "
                                          "(there is no CPU fallback)");
                        }
                    }(f),
                    ...);
            },
            surface_force_functors);
        double vtx[9], vf[3];
        lgb200_pack_vertices(vtx);
        lgb200::pack1(v_facet, vf);
        int id = -1;
        LGB200_CHECK(lg_add_facet(s, vtx, vf, sc, has_hamaker, A, h0, &id));
        return id;
    }
    // pushes the current vertices and velocity (both user-mutable) before a step
    void lgb200_sync(lg_system* s, int id) const {
        double vtx[9], vf[3];
        lgb200_pack_vertices(vtx);
        lgb200::pack1(v_facet, vf);
        LGB200_CHECK(lg_facet_set_vertices(s, id, vtx));
        LGB200_CHECK(lg_facet_set_velocity(s, id, vf));
    }

    void lgb200_set_vertices(const double vtx[9]) {
        std::get<0>(vertices) = field_value_t(vtx[0], vtx[1], vtx[2]);
        std::get<1>(vertices) = field_value_t(vtx[3], vtx[4], vtx[5]);
        std::get<2>(vertices) = field_value_t(vtx[6], vtx[7], vtx[8]);
    }

private:
    void lgb200_pack_vertices(double out[9]) const {
        lgb200::pack1(std::get<0>(vertices), out);
        lgb200::pack1(std::get<1>(vertices), out + 3);
        lgb200::pack1(std::get<2>(vertices), out + 6);
    }

    std::tuple<field_value_t, field_value_t, field_value_t> vertices;
    std::tuple<surface_force_functors_t&...> surface_force_functors;
    field_value_t plane_unit_normal;
};

#endif  // LIBGRAN_TRIANGULAR_FACET_H
