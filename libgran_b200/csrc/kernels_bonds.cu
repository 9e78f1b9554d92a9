// kernels_bonds.cu -- device-side bond detection and bridge construction.
//
// Replaces the O(N^2) double loop of sinter_functor's constructor (LG/sinter_bridge/sinter_bridge.h:148-211): candidate
// partners come from the cell list (rows of radius 2 r + crit built by the neighbor kernels, ascending j), the predicate and
// the construction are the reference's (bond_build.cuh).  The reference consumes one random unit vector per bond in
// lexicographic (i < j) pair order; here bond number k -- its rank in that order, from two prefix sums -- takes vector k of
// the caller's array.  A vector collinear with its bond axis (which the reference would redraw, :171-175, shifting every
// later draw) raises a flag; the host then falls back to the sequential scan.
//
// One thread per particle i writes every ordered record (i, j) of i: partner, mirror index, five connectors, rest lengths.
#include "bond_build.cuh"
#include "system.h"

namespace lg {

namespace {

__device__ __forceinline__ bool bonded_pair(const double* x, int a, int b, double r, double crit) {
    const int lo = a < b ? a : b, hi = a < b ? b : a;
    return bridge_candidate(load3(x + 3 * (size_t)lo), load3(x + 3 * (size_t)hi), r, crit);
}

// deg[i] = bonded partners of i, up[i] = those with j > i
__global__ void __launch_bounds__(128) bond_count_kernel(int n, const double* x, const uint32_t* row_ptr, const int32_t* cols, double r,
                                                         double crit, uint32_t* deg, uint32_t* up) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t d = 0, u = 0;
    for (uint32_t p = row_ptr[i]; p < row_ptr[i + 1]; p++) {
        const int j = cols[p];
        if (bonded_pair(x, i, j, r, crit)) {
            d++;
            u += j > i ? 1u : 0u;
        }
    }
    deg[i] = d;
    up[i] = u;
}

// bptr / upoff: exclusive scans of deg / up
__global__ void __launch_bounds__(128) bond_fill_kernel(int n, const double* x, const uint32_t* row_ptr, const int32_t* cols, double r,
                                                        double crit, const uint32_t* bptr, const uint32_t* upoff, const double* uv,
                                                        int64_t n_uv, int32_t* bj, int32_t* brev, double* bconn, double* blen, int* flags) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t slot = bptr[i];
    for (uint32_t p = row_ptr[i]; p < row_ptr[i + 1]; p++) {
        const int j = cols[p];
        if (!bonded_pair(x, i, j, r, crit)) continue;
        const int lo = i < j ? i : j, hi = i < j ? j : i;
        // rank of the bond (lo, hi) in lexicographic order, and position of the mirror record (j, i) in row j
        uint32_t rank = upoff[lo], mirror = bptr[j];
        for (uint32_t q = row_ptr[lo]; q < row_ptr[lo + 1]; q++) {
            const int k = cols[q];
            if (k > lo && k < hi && bonded_pair(x, lo, k, r, crit)) rank++;
        }
        for (uint32_t q = row_ptr[j]; q < row_ptr[j + 1]; q++) {
            const int k = cols[q];
            if (k < i && bonded_pair(x, j, k, r, crit)) mirror++;
        }
        double ci[15], cj[15];
        bool ok = (int64_t)rank < n_uv;
        if (!ok) {
            atomicOr(flags, 1);  // more bonds than unit vectors
        } else {
            const d3 tv = load3(uv + 3 * (size_t)rank);
            ok = build_bridge(load3(x + 3 * (size_t)lo), load3(x + 3 * (size_t)hi), tv, r, ci, cj);
            if (!ok) atomicOr(flags, 2);  // collinear draw: the reference would redraw
        }
        const double* mine = i == lo ? ci : cj;
        bj[slot] = j;
        brev[slot] = (int32_t)mirror;
        for (int k = 0; k < 15; k++) bconn[(size_t)slot * 15 + k] = ok ? mine[k] : 0.0;
        for (int k = 0; k < 5; k++) blen[(size_t)slot * 5 + k] = k < 4 ? 2.0 * r : 0.0;
        slot++;
    }
}

}  // namespace

void launch_bond_count(int n, const double* x, const uint32_t* row_ptr, const int32_t* cols, double r, double crit, uint32_t* deg,
                       uint32_t* up, cudaStream_t st) {
    bond_count_kernel<<<(n + 127) / 128, 128, 0, st>>>(n, x, row_ptr, cols, r, crit, deg, up);
    g_launch_count++;
}

void launch_bond_fill(int n, const double* x, const uint32_t* row_ptr, const int32_t* cols, double r, double crit, const uint32_t* bptr,
                      const uint32_t* upoff, const double* uv, int64_t n_uv, int32_t* bj, int32_t* brev, double* bconn, double* blen,
                      int* flags, cudaStream_t st) {
    bond_fill_kernel<<<(n + 127) / 128, 128, 0, st>>>(n, x, row_ptr, cols, r, crit, bptr, upoff, uv, n_uv, bj, brev, bconn, blen, flags);
    g_launch_count++;
}

}  // namespace lg
