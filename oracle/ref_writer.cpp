// ref_writer.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// C entry points over the UNMODIFIED dump functions of the reference (writer.cpp:13-170, compiled where it lies under
// /root/reference by oracle/Makefile: _ref/libref_writer.so): the byte-for-byte checker of the product's dump pipeline
// (libgran_b200/csrc/dump_impl.cuh) and the generator of tests/golden/vtk/* (tests/golden/make_golden_vtk.py).
// The reference numbers its files with static counters that start at 1 (writer.cpp:9-11): one call per kind and process
// writes <dir>/particles_1.vtk, <dir>/necks_1.vtk, <dir>/spring_connectors_1.vtk, the next one _2, ...
#include <array>
#include <cstdint>
#include <string>
#include <vector>

#include <Eigen/Eigen>

#include "writer.h"  // $(REF)/writer.h

static std::vector<Eigen::Vector3d> vecs(const double* p, int64_t n) {
    std::vector<Eigen::Vector3d> out((size_t)n);
    for (int64_t i = 0; i < n; i++) out[(size_t)i] = Eigen::Vector3d(p[3 * i], p[3 * i + 1], p[3 * i + 2]);
    return out;
}

extern "C" {

void refw_write_particles(const char* dir, int64_t n, const double* x, const double* theta, double r_part) {
    write_particles(dir, vecs(x, n), vecs(theta, n), r_part);
}

// bonded: n * n bytes, 1 = bonded pair.  The reference's argument is `enabled_contacts` = NOT bonded (writer.cpp:28,39,137).
void refw_write_necks(const char* dir, int64_t n, const double* x, double r_part, const uint8_t* bonded) {
    std::vector<bool> enabled((size_t)(n * n));
    for (int64_t k = 0; k < n * n; k++) enabled[(size_t)k] = !bonded[k];
    write_necks(dir, vecs(x, n), r_part, enabled);
}

// connectors: n * n * 15 doubles (dense, row-major (i, j)), only bonded entries are read
void refw_write_spring_connectors(const char* dir, int64_t n, double r_part, const double* connectors, const uint8_t* bonded) {
    std::vector<std::array<Eigen::Vector3d, 5>> conn((size_t)(n * n));
    std::vector<bool> enabled((size_t)(n * n));
    for (int64_t k = 0; k < n * n; k++) {
        enabled[(size_t)k] = !bonded[k];
        for (int q = 0; q < 5; q++)
            conn[(size_t)k][q] = Eigen::Vector3d(connectors[15 * k + 3 * q], connectors[15 * k + 3 * q + 1], connectors[15 * k + 3 * q + 2]);
    }
    write_spring_connectors(dir, r_part, conn, enabled);
}

}  // extern "C"
