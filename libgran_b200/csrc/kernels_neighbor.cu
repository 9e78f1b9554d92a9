// kernels_neighbor.cu -- cell-list rebuild of the neighbor rows (sm_100a).
//
// Replaces the O(N^2) update_neighbor_list() of
// LT/rotational_system/rotational_binary_system_neighbors_omp.h:102-116:
//     row(i) = { j != i : (x[i] - x[j]).norm() < r_verlet },  ascending j
// by: bounding box -> cell index + histogram -> exclusive prefix scan -> counting-sort scatter ->
// per-particle 27-cell sweep (count) -> scan -> sweep again (fill) -> per-row sort by particle id.
// The cell grid is only a candidate filter (cell edge > r_verlet); membership is decided by the
// reference's exact predicate: sqrt((dx*dx + dy*dy) + dz*dz) < r_verlet, FP64, no FMA contraction
// (this file is compiled with --fmad=false), strict '<' -- so the pair set is bit-identical.
#include "system.h"

#include <cstdlib>
#include <mutex>

namespace lg {

// ---- bounding box --------------------------------------------------------------------------------
constexpr int kBboxThreads = 256;
constexpr int kBboxMaxBlocks = 1024;

__device__ __forceinline__ void warp_minmax(double* mn, double* mx) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            mn[k] = fmin(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], off));
            mx[k] = fmax(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], off));
        }
    }
}

__global__ void __launch_bounds__(kBboxThreads) bbox_partial_kernel(int n, const double* x, double* partial) {
    double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const double c = x[3 * (size_t)i + k];
            mn[k] = fmin(mn[k], c);
            mx[k] = fmax(mx[k], c);
        }
    }
    warp_minmax(mn, mx);
    __shared__ double s[kBboxThreads / 32][6];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0)
        for (int k = 0; k < 3; k++) {
            s[warp][k] = mn[k];
            s[warp][3 + k] = mx[k];
        }
    __syncthreads();
    if (warp == 0) {
        const bool have = lane < kBboxThreads / 32;
        for (int k = 0; k < 3; k++) {
            mn[k] = have ? s[lane][k] : 1e300;
            mx[k] = have ? s[lane][3 + k] : -1e300;
        }
        warp_minmax(mn, mx);
        if (lane == 0)
            for (int k = 0; k < 3; k++) {
                partial[blockIdx.x * 6 + k] = mn[k];
                partial[blockIdx.x * 6 + 3 + k] = mx[k];
            }
    }
}

__global__ void __launch_bounds__(kBboxThreads) bbox_final_kernel(int n_partial, const double* partial, double* out) {
    double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
    for (int b = threadIdx.x; b < n_partial; b += blockDim.x)
        for (int k = 0; k < 3; k++) {
            mn[k] = fmin(mn[k], partial[b * 6 + k]);
            mx[k] = fmax(mx[k], partial[b * 6 + 3 + k]);
        }
    warp_minmax(mn, mx);
    __shared__ double s[kBboxThreads / 32][6];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0)
        for (int k = 0; k < 3; k++) {
            s[warp][k] = mn[k];
            s[warp][3 + k] = mx[k];
        }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int wv = 1; wv < kBboxThreads / 32; wv++)
            for (int k = 0; k < 3; k++) {
                s[0][k] = fmin(s[0][k], s[wv][k]);
                s[0][3 + k] = fmax(s[0][3 + k], s[wv][3 + k]);
            }
        for (int k = 0; k < 6; k++) out[k] = s[0][k];
    }
}

// d_scratch: 6 * kBboxMaxBlocks doubles
void launch_bbox(int n, const double* x, double* d_scratch, double* d_out, cudaStream_t st) {
    int blocks = (n + kBboxThreads - 1) / kBboxThreads;
    if (blocks > kBboxMaxBlocks) blocks = kBboxMaxBlocks;
    if (blocks < 1) blocks = 1;
    bbox_partial_kernel<<<blocks, kBboxThreads, 0, st>>>(n, x, d_scratch);
    bbox_final_kernel<<<1, kBboxThreads, 0, st>>>(blocks, d_scratch, d_out);
    g_launch_count += 2;
}

// ---- exclusive prefix scan of u32 (three-phase, hand-written) ----------------------------------------
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;  // 2048 elements per block

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total) {
    __shared__ uint32_t warp_sums[kScanThreads / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t ws = (lane < kScanThreads / 32) ? warp_sums[lane] : 0u;
        uint32_t wincl = ws;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, wincl, off);
            if (lane >= off) wincl += t;
        }
        if (lane < kScanThreads / 32) warp_sums[lane] = wincl - ws;  // exclusive offsets of the warps
        if (lane == kScanThreads / 32 - 1) *total = wincl;
    }
    __syncthreads();
    const uint32_t res = warp_sums[warp] + (incl - v);
    __syncthreads();
    return res;
}

__global__ void __launch_bounds__(kScanThreads) scan_tile_kernel(uint32_t* data, size_t n, uint32_t* tile_sums) {
    __shared__ uint32_t total;
    const size_t base = (size_t)blockIdx.x * kScanTile + (size_t)threadIdx.x * kScanItems;
    uint32_t item[kScanItems];
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < kScanItems; k++) {
        item[k] = (base + k < n) ? data[base + k] : 0u;
        sum += item[k];
    }
    uint32_t run = block_exclusive_scan(sum, &total);
#pragma unroll
    for (int k = 0; k < kScanItems; k++) {
        if (base + k < n) data[base + k] = run;
        run += item[k];
    }
    if (threadIdx.x == 0 && tile_sums != nullptr) tile_sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kScanThreads) scan_add_kernel(uint32_t* data, size_t n, const uint32_t* tile_offsets) {
    const uint32_t off = tile_offsets[blockIdx.x];
    const size_t base = (size_t)blockIdx.x * kScanTile + (size_t)threadIdx.x * kScanItems;
#pragma unroll
    for (int k = 0; k < kScanItems; k++)
        if (base + k < n) data[base + k] += off;
}

size_t scan_scratch_elems(size_t n) {
    size_t total = 0;
    while (n > (size_t)kScanTile) {
        n = (n + kScanTile - 1) / kScanTile;
        total += n;
    }
    return total + 1;
}

void launch_exclusive_scan_u32(uint32_t* data, size_t n, uint32_t* scratch, cudaStream_t st) {
    if (n == 0) return;
    const size_t tiles = (n + kScanTile - 1) / kScanTile;
    if (tiles == 1) {
        scan_tile_kernel<<<1, kScanThreads, 0, st>>>(data, n, nullptr);
        g_launch_count++;
        return;
    }
    scan_tile_kernel<<<(unsigned)tiles, kScanThreads, 0, st>>>(data, n, scratch);
    launch_exclusive_scan_u32(scratch, tiles, scratch + tiles, st);
    scan_add_kernel<<<(unsigned)tiles, kScanThreads, 0, st>>>(data, n, scratch);
    g_launch_count += 2;
}

// ---- cell list -------------------------------------------------------------------------------------
__device__ __forceinline__ void cell_coords(const NeighborGrid& g, d3 p, int c[3]) {
    const double q[3] = {(p.x - g.lo[0]) / g.cell, (p.y - g.lo[1]) / g.cell, (p.z - g.lo[2]) / g.cell};
#pragma unroll
    for (int k = 0; k < 3; k++) {
        int t = (int)floor(q[k]);
        t = t < 0 ? 0 : t;
        c[k] = t >= g.dim[k] ? g.dim[k] - 1 : t;
    }
}
__device__ __forceinline__ int cell_linear(const NeighborGrid& g, int cx, int cy, int cz) {
    return (cx * g.dim[1] + cy) * g.dim[2] + cz;
}

__global__ void __launch_bounds__(256)
    cell_count_kernel(int n, const double* x, const NeighborGrid g, int32_t* cell_of, uint32_t* cell_count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int c[3];
    cell_coords(g, load3(x + 3 * (size_t)i), c);
    const int cl = cell_linear(g, c[0], c[1], c[2]);
    cell_of[i] = cl;
    atomicAdd(&cell_count[cl], 1u);  // integer histogram: the result does not depend on arrival order
}

__global__ void __launch_bounds__(256) cell_fill_kernel(int n, const int32_t* cell_of, const uint32_t* cell_start,
                                                        uint32_t* cell_cursor, int32_t* order) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int cl = cell_of[i];
    const uint32_t pos = atomicAdd(&cell_cursor[cl], 1u);
    order[cell_start[cl] + pos] = i;  // order inside a cell is arbitrary; rows are sorted by id afterwards
}

void launch_cell_count(int n, const double* x, NeighborGrid g, int32_t* cell_of, uint32_t* cell_count, cudaStream_t st) {
    cell_count_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, x, g, cell_of, cell_count);
    g_launch_count++;
}
void launch_cell_fill(int n, const int32_t* cell_of, const uint32_t* cell_start, uint32_t* cell_cursor, int32_t* order,
                      cudaStream_t st) {
    cell_fill_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, cell_of, cell_start, cell_cursor, order);
    g_launch_count++;
}

// The reference predicate, term order as Eigen evaluates (x[i] - x[j]).norm().
__device__ __forceinline__ bool within_verlet(d3 xi, d3 xj, double r_verlet) {
    const d3 d = xi - xj;
    return sqrt((d.x * d.x + d.y * d.y) + d.z * d.z) < r_verlet;
}

constexpr int kSweepRowCap = 64;  // hit-list entries per particle (second version of the rebuild, below)

// ONLY_LONG (fill pass): leave the rows that fit the hit list to row_fill_hits_kernel (below)
template <bool FILL, bool ONLY_LONG>
__global__ void __launch_bounds__(128)
    row_sweep_kernel(int n, const double* x, const NeighborGrid g, const uint32_t* cell_start, const int32_t* order,
                     double r_verlet, uint32_t* row_count, const uint32_t* row_ptr, int32_t* cols, const uint8_t* owned) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (FILL && ONLY_LONG && row_ptr[i + 1] - row_ptr[i] <= (uint32_t)kSweepRowCap) return;
    if (owned != nullptr && !owned[i]) {  // ghost particle (slab decomposition): its forces belong to another rank
        if (!FILL) row_count[i] = 0u;
        return;
    }
    const d3 xi = load3(x + 3 * (size_t)i);
    int c[3];
    cell_coords(g, xi, c);
    uint32_t cnt = 0;
    const uint32_t base = FILL ? row_ptr[i] : 0u;
    for (int dx = -1; dx <= 1; dx++) {
        const int cx = c[0] + dx;
        if (cx < 0 || cx >= g.dim[0]) continue;
        for (int dy = -1; dy <= 1; dy++) {
            const int cy = c[1] + dy;
            if (cy < 0 || cy >= g.dim[1]) continue;
            // the three z-neighbours are contiguous in the cell order: one range
            const int z0 = max(c[2] - 1, 0), z1 = min(c[2] + 1, g.dim[2] - 1);
            const uint32_t s0 = cell_start[cell_linear(g, cx, cy, z0)];
            const uint32_t s1 = cell_start[cell_linear(g, cx, cy, z1) + 1];
            for (uint32_t s = s0; s < s1; s++) {
                const int j = order[s];
                if (j == i) continue;
                if (within_verlet(xi, load3(x + 3 * (size_t)j), r_verlet)) {
                    if (FILL) cols[base + cnt] = j;
                    cnt++;
                }
            }
        }
    }
    if (!FILL) {
        row_count[i] = cnt;
    } else {
        // insertion sort of the row by particle id (rows are short; done once per rebuild)
        int32_t* row = cols + base;
        for (uint32_t a = 1; a < cnt; a++) {
            const int32_t key = row[a];
            int b = (int)a - 1;
            while (b >= 0 && row[b] > key) {
                row[b + 1] = row[b];
                b--;
            }
            row[b + 1] = key;
        }
    }
}

// ---- rows, second version (round 2) ------------------------------------------------------------------------------------
// What the first version paid for (ncu launch lists of configs 3, 4, 5: the fill pass cost 1.5-2.4 x the count pass): a second
// full sweep of the 27 cells, an IEEE square root per candidate, dependent order[] -> x[] gathers one candidate at a time, an
// insertion sort of the row IN GLOBAL MEMORY and 4-byte column stores scattered over 32 rows per warp instruction.  Now:
//   count pass  row_count_hits_kernel: the sweep, kSweepBatch candidates in flight, no square root outside a thin band, and the
//               ids it finds go to a hit list hits[k * n + i] (k < kSweepRowCap; entry k of every particle side by side);
//   fill pass   row_fill_hits_kernel: NO sweep -- reads the hit list, sorts the row in shared memory and writes the columns of a
//               warp's 32 consecutive rows as one coalesced span;  rows longer than kSweepRowCap (dense inputs with a wide
//               skin) are left to the first version's kernel, restricted to those rows.
// The predicate is still the reference's: sqrt((dx*dx + dy*dy) + dz*dz) < r_verlet decides every pair whose d^2 lies within
// 1e-9 (relative) of r_verlet^2; outside that band the comparison of d^2 alone gives the same answer (sqrt is monotone and
// correctly rounded: d2 < r^2 (1 - 1e-9) implies fl(sqrt(d2)) <= r (1 - 5e-10)(1 + 2^-53) < r, and symmetrically above).
constexpr int kSweepThreads = 128;
constexpr int kSweepBatch = 4;    // candidates in flight per thread

int neighbor_hit_list_entries() { return kSweepRowCap; }

struct VerletTest {
    double r, lo2, hi2;
};
__device__ __forceinline__ bool within_verlet_banded(const VerletTest& t, d3 xi, d3 xj) {
    const d3 d = xi - xj;
    const double d2 = (d.x * d.x + d.y * d.y) + d.z * d.z;
    if (d2 < t.lo2) return true;
    if (d2 >= t.hi2) return false;
    return sqrt(d2) < t.r;
}

__global__ void __launch_bounds__(kSweepThreads, 7)
    row_count_hits_kernel(int n, const double* __restrict__ x, const NeighborGrid g, const uint32_t* __restrict__ cell_start,
                          const int32_t* __restrict__ order, double r_verlet, uint32_t* __restrict__ row_count,
                          int32_t* __restrict__ hits, const uint8_t* __restrict__ owned) {
    const int i = blockIdx.x * kSweepThreads + threadIdx.x;
    if (i >= n) return;
    if (owned != nullptr && !owned[i]) {  // ghost particle (slab decomposition): its forces belong to another rank
        row_count[i] = 0u;
        return;
    }
    VerletTest vt;
    vt.r = r_verlet;
    vt.lo2 = (r_verlet * r_verlet) * (1.0 - 1e-9);
    vt.hi2 = (r_verlet * r_verlet) * (1.0 + 1e-9);
    const d3 xi = load3(x + 3 * (size_t)i);
    int c[3];
    cell_coords(g, xi, c);
    uint32_t cnt = 0;
    for (int dx = -1; dx <= 1; dx++) {
        const int cx = c[0] + dx;
        if (cx < 0 || cx >= g.dim[0]) continue;
        for (int dy = -1; dy <= 1; dy++) {
            const int cy = c[1] + dy;
            if (cy < 0 || cy >= g.dim[1]) continue;
            // the three z-neighbours are contiguous in the cell order: one range
            const int z0 = max(c[2] - 1, 0), z1 = min(c[2] + 1, g.dim[2] - 1);
            const uint32_t s0 = cell_start[cell_linear(g, cx, cy, z0)];
            const uint32_t s1 = cell_start[cell_linear(g, cx, cy, z1) + 1];
            for (uint32_t s = s0; s < s1; s += kSweepBatch) {
                int j[kSweepBatch];
                d3 xj[kSweepBatch];
#pragma unroll
                for (int u = 0; u < kSweepBatch; u++) j[u] = s + u < s1 ? order[s + u] : i;  // i = "skip"
#pragma unroll
                for (int u = 0; u < kSweepBatch; u++) xj[u] = load3(x + 3 * (size_t)j[u]);
#pragma unroll
                for (int u = 0; u < kSweepBatch; u++) {
                    if (j[u] == i || !within_verlet_banded(vt, xi, xj[u])) continue;
                    if (hits != nullptr && cnt < (uint32_t)kSweepRowCap) hits[(size_t)cnt * (size_t)n + (size_t)i] = j[u];
                    cnt++;
                }
            }
        }
    }
    row_count[i] = cnt;
}

// Staging slot of entry k of thread t's row: column t of a [kSweepRowCap][kSweepThreads] array, rotated by k inside the
// warp's 32 columns -- a warp's lanes hit 32 different banks both when every lane touches entry k of its OWN row (loading,
// sorting) and when consecutive lanes read consecutive entries of ONE row (the span write).
__device__ __forceinline__ int sweep_slot(int k, int t) { return k * kSweepThreads + (t & ~31) + ((t + k) & 31); }

__global__ void __launch_bounds__(kSweepThreads, 6)
    row_fill_hits_kernel(int n, const uint32_t* __restrict__ row_ptr, const int32_t* __restrict__ hits,
                         int32_t* __restrict__ cols) {
    __shared__ int32_t sbuf[kSweepRowCap * kSweepThreads];
    __shared__ uint32_t sbase[kSweepThreads];
    __shared__ uint32_t scnt[kSweepThreads];
    const int tid = threadIdx.x;
    const int i = blockIdx.x * kSweepThreads + tid;
    const uint32_t base = row_ptr[i < n ? i : n];
    uint32_t cnt = i < n ? row_ptr[i + 1] - base : 0u;
    if (cnt > (uint32_t)kSweepRowCap) cnt = 0u;  // a long row: not in the hit list, written by row_sweep_kernel<true, true>
    for (int k = 0; k < (int)cnt; k++) sbuf[sweep_slot(k, tid)] = hits[(size_t)k * (size_t)n + (size_t)i];
    for (int a = 1; a < (int)cnt; a++) {  // insertion sort by particle id
        const int32_t key = sbuf[sweep_slot(a, tid)];
        int b = a - 1;
        while (b >= 0) {
            const int32_t vb = sbuf[sweep_slot(b, tid)];
            if (vb <= key) break;
            sbuf[sweep_slot(b + 1, tid)] = vb;
            b--;
        }
        sbuf[sweep_slot(b + 1, tid)] = key;
    }
    // The rows of a warp's 32 consecutive particles are one contiguous span of cols: write it with consecutive lanes on
    // consecutive columns.  Column q belongs to the last row whose base is <= q (empty rows share their successor's base).
    sbase[tid] = base;
    scnt[tid] = cnt;
    __syncwarp();
    const int lane = tid & 31, w0 = tid & ~31;
    const uint32_t span0 = sbase[w0];
    const uint32_t span1 = __shfl_sync(0xffffffffu, i < n ? row_ptr[i + 1] : base, 31);
    for (uint32_t q = span0 + (uint32_t)lane; q < span1; q += 32u) {
        int r = 0;
#pragma unroll
        for (int step = 16; step > 0; step >>= 1)
            if (sbase[w0 + r + step] <= q) r += step;
        const uint32_t k = q - sbase[w0 + r];
        if (k < scnt[w0 + r]) cols[q] = sbuf[sweep_slot((int)k, w0 + r)];
    }
}

static bool rebuild_v1() {  // LGB200_REBUILD_V1=1: the round-1 kernels only, kept for A/B timing (scripts/rebuild_ab.py)
    static const bool v1 = [] {
        const char* e = std::getenv("LGB200_REBUILD_V1");
        return e != nullptr && e[0] == '1';
    }();
    return v1;
}

// hits: n * neighbor_hit_list_entries() ints of scratch shared by the two passes, or nullptr (then the fill pass sweeps again)
void launch_row_count(int n, const double* x, NeighborGrid g, const uint32_t* cell_start, const int32_t* order,
                      double r_verlet, uint32_t* row_count, int32_t* hits, const uint8_t* owned, cudaStream_t st) {
    if (rebuild_v1())
        row_sweep_kernel<false, false><<<(n + 127) / 128, 128, 0, st>>>(n, x, g, cell_start, order, r_verlet, row_count,
                                                                        nullptr, nullptr, owned);
    else
        row_count_hits_kernel<<<(n + kSweepThreads - 1) / kSweepThreads, kSweepThreads, 0, st>>>(
            n, x, g, cell_start, order, r_verlet, row_count, hits, owned);
    g_launch_count++;
}
void launch_row_fill(int n, const double* x, NeighborGrid g, const uint32_t* cell_start, const int32_t* order,
                     double r_verlet, const uint32_t* row_ptr, int32_t* cols, const int32_t* hits, const uint8_t* owned,
                     cudaStream_t st) {
    if (rebuild_v1() || hits == nullptr) {
        row_sweep_kernel<true, false><<<(n + 127) / 128, 128, 0, st>>>(n, x, g, cell_start, order, r_verlet, nullptr, row_ptr,
                                                                       cols, owned);
        g_launch_count++;
        return;
    }
    row_fill_hits_kernel<<<(n + kSweepThreads - 1) / kSweepThreads, kSweepThreads, 0, st>>>(n, row_ptr, hits, cols);
    row_sweep_kernel<true, true><<<(n + 127) / 128, 128, 0, st>>>(n, x, g, cell_start, order, r_verlet, nullptr, row_ptr, cols,
                                                                  owned);
    g_launch_count += 2;
}

// Capacity check of a rebuild that runs without a host round trip: d_total[0] <- number of pairs, d_total[1] <- 1 and
// every row emptied when they do not fit into `cap` column slots (the host learns it from the error flag at its next
// synchronisation; nothing is written out of bounds and later kernels see empty rows).
__global__ void pair_guard_kernel(const uint32_t* row_ptr, int n, uint32_t cap, uint32_t* d_total, int* err) {
    const uint32_t total = row_ptr[n];
    d_total[0] = total;
    d_total[1] = total > cap ? 1u : 0u;
    if (total > cap) atomicOr(err, ERR_PAIR_CAPACITY);
}
__global__ void __launch_bounds__(256) pair_clamp_kernel(uint32_t* row_ptr, int n, const uint32_t* d_total) {
    if (d_total[1] == 0u) return;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i <= n) row_ptr[i] = 0u;
}
void launch_pair_guard(int n, uint32_t* row_ptr, uint32_t cap, uint32_t* d_total, int* err, cudaStream_t st) {
    pair_guard_kernel<<<1, 1, 0, st>>>(row_ptr, n, cap, d_total, err);
    pair_clamp_kernel<<<(n + 256) / 256, 256, 0, st>>>(row_ptr, n, d_total);
    g_launch_count += 2;
}

// all-pairs loop (rotational_binary_system_omp.h:72-75): row(i) = every j != i, ascending
__global__ void __launch_bounds__(256) allpairs_fill_kernel(int n, uint32_t* row_ptr, int32_t* cols) {
    const int i = blockIdx.x;
    const uint32_t base = (uint32_t)i * (uint32_t)(n - 1);
    if (threadIdx.x == 0) {
        row_ptr[i] = base;
        if (i == n - 1) row_ptr[n] = (uint32_t)n * (uint32_t)(n - 1);
    }
    for (int j = threadIdx.x; j < n; j += blockDim.x)
        if (j != i) cols[base + (uint32_t)(j < i ? j : j - 1)] = j;
}
void launch_allpairs_fill(int n, uint32_t* row_ptr, int32_t* cols, cudaStream_t st) {
    allpairs_fill_kernel<<<n, 256, 0, st>>>(n, row_ptr, cols);
    g_launch_count++;
}

// ---- tags: where does the history of pair p live? ----------------------------------------------------
// Carries the sparse contact-spring records across a rebuild: a record keyed by partner id survives if the
// partner is still in the row; otherwise it is dropped (the reference would have reset it on the last
// visited non-overlapping evaluation before the pair left the list).
// tag[0 .. row_ptr[n]) <- kTagNone with 16-byte stores (the first version did it row by row inside retag_kernel: one
// scattered byte store per pair)
__global__ void __launch_bounds__(256) tag_clear_kernel(const uint32_t* row_ptr, int n, uint8_t* tag) {
    const uint32_t total = row_ptr[n];
    const uint32_t fill = 0x01010101u * (uint32_t)(uint8_t)kTagNone;
    const uint4 v = make_uint4(fill, fill, fill, fill);
    const uint32_t n16 = total >> 4;
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < n16; k += gridDim.x * blockDim.x)
        reinterpret_cast<uint4*>(tag)[k] = v;
    if (blockIdx.x == 0 && threadIdx.x < (total & 15u)) tag[(n16 << 4) + threadIdx.x] = (uint8_t)kTagNone;
}

template <bool CLEAR_ROWS>
__global__ void __launch_bounds__(128)
    retag_kernel(int n, const uint32_t* row_ptr, const int32_t* cols, uint8_t* tag, int32_t* sgid, double* spr, int S,
                 const uint32_t* bptr, const int32_t* bj, int* err) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t p0 = row_ptr[i], p1 = row_ptr[i + 1];
    if (CLEAR_ROWS)
        for (uint32_t p = p0; p < p1; p++) tag[p] = (uint8_t)kTagNone;
    if (sgid != nullptr) {
        for (int q = 0; q < S; q++) {
            const int32_t partner = sgid[(size_t)i * S + q];
            if (partner < 0) continue;
            // binary search of partner in the sorted row
            uint32_t lo = p0, hi = p1;
            while (lo < hi) {
                const uint32_t mid = (lo + hi) >> 1;
                if (cols[mid] < partner) lo = mid + 1; else hi = mid;
            }
            if (lo < p1 && cols[lo] == partner)
                tag[lo] = (uint8_t)q;
            else
                sgid[(size_t)i * S + q] = -1;
        }
    }
    if (bptr != nullptr) {
        const uint32_t b0 = bptr[i], b1 = bptr[i + 1];
        for (uint32_t b = b0; b < b1; b++) {
            const int32_t partner = bj[b];
            uint32_t lo = p0, hi = p1;
            while (lo < hi) {
                const uint32_t mid = (lo + hi) >> 1;
                if (cols[mid] < partner) lo = mid + 1; else hi = mid;
            }
            if (lo < p1 && cols[lo] == partner) {
                if (tag[lo] != kTagNone) {  // a bonded pair must never own a contact slot
                    sgid[(size_t)i * S + tag[lo]] = -1;
                }
                tag[lo] = (uint8_t)(kTagBond + (b - b0));
            }
            // a bonded partner outside the row is simply not visited, as in the reference
        }
    }
    (void)spr;
    (void)err;
}

void launch_retag(int n, const uint32_t* row_ptr, const int32_t* cols, uint8_t* tag, int32_t* sgid, double* spr, int S,
                  const uint32_t* bptr, const int32_t* bj, int* err, cudaStream_t st) {
    if (n <= 0) return;
    if (rebuild_v1()) {
        retag_kernel<true><<<(n + 127) / 128, 128, 0, st>>>(n, row_ptr, cols, tag, sgid, spr, S, bptr, bj, err);
        g_launch_count++;
        return;
    }
    tag_clear_kernel<<<148 * 4, 256, 0, st>>>(row_ptr, n, tag);
    retag_kernel<false><<<(n + 127) / 128, 128, 0, st>>>(n, row_ptr, cols, tag, sgid, spr, S, bptr, bj, err);
    g_launch_count += 2;
}

__global__ void __launch_bounds__(256) count_slots_kernel(size_t total, const int32_t* sgid, unsigned long long* out) {
    unsigned long long c = 0;
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += (size_t)gridDim.x * blockDim.x)
        c += sgid[k] >= 0 ? 1ull : 0ull;
    for (int off = 16; off > 0; off >>= 1) c += __shfl_xor_sync(0xffffffffu, c, off);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(out, c);
}
void launch_count_slots(int n, const int32_t* sgid, int S, unsigned long long* d_out, cudaStream_t st) {
    cudaMemsetAsync(d_out, 0, sizeof(unsigned long long), st);
    if (n <= 0 || sgid == nullptr) return;
    const size_t total = (size_t)n * S;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 2048) blocks = 2048;
    count_slots_kernel<<<blocks, 256, 0, st>>>(total, sgid, d_out);
    g_launch_count++;
}

// ---- diagnostics (test/compute_energy.cpp:7-58, test/mass_distribution.cpp:7-26) --------------------
// pass 0 sums: [0] m v.v/2  [1] I w.w/2  [2..4] sum v  [5..7] sum I w + x cross m v  [8..10] sum x
// pass 1 sums: [11] sum |x - com|^2
constexpr int kDiagThreads = 256;
constexpr int kDiagMaxBlocks = 512;
constexpr int kDiagSlots = 12;

__global__ void __launch_bounds__(kDiagThreads)
    diag_partial_kernel(int n, const double* x, const double* v, const double* w, double mass, double inertia,
                        const double* com, double* partial, int pass, const uint8_t* owned) {
    double s[kDiagSlots];
#pragma unroll
    for (int k = 0; k < kDiagSlots; k++) s[k] = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        if (owned != nullptr && !owned[i]) continue;  // slab decomposition: ghosts belong to another rank's sums
        const d3 xi = load3(x + 3 * (size_t)i);
        if (pass == 0) {
            const d3 vi = load3(v + 3 * (size_t)i), wi = load3(w + 3 * (size_t)i);
            s[0] += mass * dot(vi, vi) / 2.0;
            s[1] += inertia * dot(wi, wi) / 2.0;
            s[2] += vi.x; s[3] += vi.y; s[4] += vi.z;
            const d3 l = inertia * wi + cross(xi, mass * vi);
            s[5] += l.x; s[6] += l.y; s[7] += l.z;
            s[8] += xi.x; s[9] += xi.y; s[10] += xi.z;
        } else {
            const d3 d = xi - mk3(com[0], com[1], com[2]);
            s[11] += dot(d, d);
        }
    }
    __shared__ double sh[kDiagThreads / 32][kDiagSlots];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < kDiagSlots; k++) {
        double t = s[k];
        for (int off = 16; off > 0; off >>= 1) t += __shfl_xor_sync(0xffffffffu, t, off);
        if (lane == 0) sh[warp][k] = t;
    }
    __syncthreads();
    if (threadIdx.x < kDiagSlots) {
        double t = 0.0;
        for (int wv = 0; wv < kDiagThreads / 32; wv++) t += sh[wv][threadIdx.x];
        partial[blockIdx.x * kDiagSlots + threadIdx.x] = t;
    }
}

__global__ void diag_final_kernel(int n_partial, const double* partial, double* out, int pass) {
    const int k = threadIdx.x;
    if (k >= kDiagSlots) return;
    if ((pass == 0) != (k < 11)) return;
    double t = 0.0;
    for (int b = 0; b < n_partial; b++) t += partial[b * kDiagSlots + k];  // fixed order: deterministic
    out[k] = t;
}

// d_partials: kDiagSlots * kDiagMaxBlocks doubles; d_out: 16 doubles
void launch_diag_sums(int n, const double* x, const double* v, const double* w, double mass, double inertia,
                      const double* com, double* d_partials, double* d_out, int pass, const uint8_t* owned, cudaStream_t st) {
    int blocks = (n + kDiagThreads - 1) / kDiagThreads;
    if (blocks > kDiagMaxBlocks) blocks = kDiagMaxBlocks;
    if (blocks < 1) blocks = 1;
    diag_partial_kernel<<<blocks, kDiagThreads, 0, st>>>(n, x, v, w, mass, inertia, com, d_partials, pass, owned);
    diag_final_kernel<<<1, 32, 0, st>>>(blocks, d_partials, d_out, pass);
    g_launch_count += 2;
}

// ---- the same sums without any allocation (lg_diag_sums_static) ------------------------------------------------------
// For callers that may not call cudaMalloc (graph capture, serving loops): the per-block partial sums of
// diag_partial_kernel live in a file-scope device array instead of a buffer owned by an lg_system.  Every GPU that loads
// the library holds these kDiagSlots * kDiagMaxBlocks doubles (48 KiB) until exit.  All calls on a device share the array:
// two calls that overlap on the GPU (different streams) race on it.
static __device__ __align__(256) double g_diag_static_scratch[kDiagSlots * kDiagMaxBlocks];
constexpr int kDiagStaticMaxN = 1 << 26;  // stated limit of the entry point (67 108 864 particles)

constexpr int kStaticMaxDevices = 64;
static double* g_diag_static_ptr[kStaticMaxDevices];  // device address of the array, per device; null until first use
static std::mutex g_diag_static_mutex;

// The array's name is a host-side stand-in in host code: resolve its device address once per device, with that device
// current (the caller's current device is the one the kernels launch on).
static cudaError_t diag_static_scratch(double** out, int* device) {
    cudaError_t e = cudaGetDevice(device);
    if (e != cudaSuccess) return e;
    if (*device < 0 || *device >= kStaticMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_diag_static_mutex);
    if (g_diag_static_ptr[*device] == nullptr) {
        double* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_diag_static_scratch);
        if (e != cudaSuccess) return e;
        g_diag_static_ptr[*device] = p;
    }
    *out = g_diag_static_ptr[*device];
    return cudaSuccess;
}

int diag_static_max_particles() { return kDiagStaticMaxN; }

// Launches diag_partial_kernel + diag_final_kernel on `st` with the static scratch; no cudaMalloc / cudaFree anywhere.
// Returns a CUDA error code (cudaSuccess = 0); *what names the call that failed.
int launch_diag_sums_static(int n, const double* x, const double* v, const double* w, double mass, double inertia,
                            const double* com, double* d_out, int pass, cudaStream_t st, const char** what) {
    double* scratch = nullptr;
    int device = -1;
    cudaError_t e = diag_static_scratch(&scratch, &device);
    if (e != cudaSuccess) {
        *what = "cudaGetSymbolAddress(g_diag_static_scratch)";
        return (int)e;
    }
    int blocks = (n + kDiagThreads - 1) / kDiagThreads;
    if (blocks > kDiagMaxBlocks) blocks = kDiagMaxBlocks;
    if (blocks < 1) blocks = 1;
    // the array keeps the previous call's partials: clear the part this call reads (every slot is also rewritten by
    // diag_partial_kernel, so this is belt and braces against a launch that fails part-way)
    e = cudaMemsetAsync(scratch, 0, (size_t)blocks * kDiagSlots * sizeof(double), st);
    if (e != cudaSuccess) {
        *what = "cudaMemsetAsync(static scratch)";
        return (int)e;
    }
    diag_partial_kernel<<<blocks, kDiagThreads, 0, st>>>(n, x, v, w, mass, inertia, com, scratch, pass, nullptr);
    diag_final_kernel<<<1, 32, 0, st>>>(blocks, scratch, d_out, pass);
    g_launch_count += 2;
    e = cudaGetLastError();
    if (e != cudaSuccess) *what = "diag_partial_kernel / diag_final_kernel launch";
    return (int)e;
}

// ---- spatial renumbering (lg_set_reorder) ------------------------------------------------
// Particles are renumbered along a Morton (Z-order) curve over cubic cells of edge `cell` (>= one particle diameter, at most
// 256 cells per axis), so that the neighbours of consecutive particles are themselves (mostly) consecutive and the staged
// position window of the row kernel holds them.  Counting sort by cell; inside a cell ascending caller id, so the numbering
// is a pure function of the positions (deterministic).
__device__ __forceinline__ uint32_t spread3(uint32_t v) {  // 8 bits -> every third bit
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
__device__ __forceinline__ uint32_t morton_key(const double* x, int i, double lx, double ly, double lz, double inv_cell) {
    const int cx = min(255, max(0, (int)floor((x[3 * (size_t)i] - lx) * inv_cell)));
    const int cy = min(255, max(0, (int)floor((x[3 * (size_t)i + 1] - ly) * inv_cell)));
    const int cz = min(255, max(0, (int)floor((x[3 * (size_t)i + 2] - lz) * inv_cell)));
    return spread3((uint32_t)cx) | (spread3((uint32_t)cy) << 1) | (spread3((uint32_t)cz) << 2);
}
__global__ void __launch_bounds__(256) morton_count_kernel(int n, const double* x, double lx, double ly, double lz, double inv_cell,
                                                           uint32_t* key, uint32_t* count) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t k = morton_key(x, i, lx, ly, lz, inv_cell);
    key[i] = k;
    atomicAdd(&count[k], 1u);
}
__global__ void __launch_bounds__(256) morton_fill_kernel(int n, const uint32_t* key, const uint32_t* start, uint32_t* cursor,
                                                          int32_t* tmp) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t k = key[i];
    tmp[start[k] + atomicAdd(&cursor[k], 1u)] = i;  // arrival order inside a cell: fixed up by morton_rank_kernel
}
// position of particle i = start of its cell + number of particles of the cell with a smaller id
__global__ void __launch_bounds__(256) morton_rank_kernel(int n, const uint32_t* key, const uint32_t* start, const int32_t* tmp,
                                                          int32_t* perm, int32_t* inv) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t k = key[i];
    const uint32_t b = start[k], e = start[k + 1];
    uint32_t rank = 0;
    for (uint32_t q = b; q < e; q++) rank += tmp[q] < i ? 1u : 0u;
    perm[b + rank] = i;   // new -> caller
    inv[i] = (int32_t)(b + rank);  // caller -> new
}
// dst[k] = src[idx[k]] for 3-vectors
__global__ void __launch_bounds__(256) gather3_kernel(int n, const int32_t* idx, const double* src, double* dst) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const size_t q = 3 * (size_t)idx[k];
    dst[3 * (size_t)k] = src[q];
    dst[3 * (size_t)k + 1] = src[q + 1];
    dst[3 * (size_t)k + 2] = src[q + 2];
}

size_t morton_cells() { return (size_t)1 << 24; }

// key: n u32; count / cursor: morton_cells() + 1 u32 each (zeroed here); tmp, perm, inv: n i32
void launch_morton_order(int n, const double* x, const double lo[3], double cell, uint32_t* key, uint32_t* count, uint32_t* cursor,
                         uint32_t* scan_scratch, int32_t* tmp, int32_t* perm, int32_t* inv, cudaStream_t st) {
    const size_t nc1 = morton_cells() + 1;
    cudaMemsetAsync(count, 0, nc1 * sizeof(uint32_t), st);
    cudaMemsetAsync(cursor, 0, nc1 * sizeof(uint32_t), st);
    const int threads = 256, blocks = (n + threads - 1) / threads;
    morton_count_kernel<<<blocks, threads, 0, st>>>(n, x, lo[0], lo[1], lo[2], 1.0 / cell, key, count);
    launch_exclusive_scan_u32(count, nc1, scan_scratch, st);
    morton_fill_kernel<<<blocks, threads, 0, st>>>(n, key, count, cursor, tmp);
    morton_rank_kernel<<<blocks, threads, 0, st>>>(n, key, count, tmp, perm, inv);
    g_launch_count += 3;
}

void launch_gather3(int n, const int32_t* idx, const double* src, double* dst, cudaStream_t st) {
    if (n <= 0) return;
    gather3_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, idx, src, dst);
    g_launch_count++;
}

}  // namespace lg
