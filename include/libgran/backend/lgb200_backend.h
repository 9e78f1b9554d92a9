// Glue between the header templates of include/libgran/** and the C ABI (include/lgb200/lgb200.h).
// Nothing here computes: it converts field_value_t <-> packed doubles and turns ABI error codes into exceptions.
#ifndef LIBGRAN_LGB200_BACKEND_H
#define LIBGRAN_LGB200_BACKEND_H

#include <cstddef>
#include <string>
#include <vector>

#include <lgb200/lgb200.h>
#include <libtimestep/exception/exception.h>

namespace lgb200 {

// field_value_t contract: constructible from three doubles, components readable through operator[]
// (Eigen::Vector3d satisfies it).
template <typename field_value_t>
inline void pack(std::vector<field_value_t> const& src, std::vector<double>& dst) {
    dst.resize(3 * src.size());
    for (std::size_t i = 0; i < src.size(); i++)
        for (int c = 0; c < 3; c++) dst[3 * i + c] = static_cast<double>(src[i][c]);
}

template <typename field_value_t>
inline void unpack(std::vector<double> const& src, std::vector<field_value_t>& dst) {
    dst.resize(src.size() / 3);
    for (std::size_t i = 0; i < dst.size(); i++) dst[i] = field_value_t(src[3 * i], src[3 * i + 1], src[3 * i + 2]);
}

template <typename field_value_t>
inline void pack1(field_value_t const& v, double out[3]) {
    for (int c = 0; c < 3; c++) out[c] = static_cast<double>(v[c]);
}

inline void check(int rc, const char* file, int line) {
    if (rc == LG_OK) return;
    std::string msg = lg_last_error();
    if (rc == LG_ESIZE) throw SizeMismatchException(msg, file, line);
    throw Lgb200BackendError(rc, msg + " [" + file + ":" + std::to_string(line) + "]");
}

// parameters every reference functor receives and the backend needs once per system
struct particle_constants {
    double r_part = 0.0, mass = 0.0, inertia = 0.0;
    bool have_r = false, have_mass = false, have_inertia = false;
    void merge(particle_constants const& o) {
        if (o.have_r && !have_r) { r_part = o.r_part; have_r = true; }
        if (o.have_mass && !have_mass) { mass = o.mass; have_mass = true; }
        if (o.have_inertia && !have_inertia) { inertia = o.inertia; have_inertia = true; }
    }
};

}  // namespace lgb200

#define LGB200_CHECK(expr) ::lgb200::check((expr), __FILE__, __LINE__)

#endif  // LIBGRAN_LGB200_BACKEND_H
