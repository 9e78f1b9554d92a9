// kernels_force_tiled.cu -- the production step kernels (sm_100a).  Two organisations of the same fused step:
//   step_tiled_kernel  pair-parallel force evaluation over a shared-memory tile with an ordered per-particle
//                      reduction; every model mask, both math modes, force-only or fused.  Below.
//   step_rows_kernel   one thread per bond record + one thread per particle, no tile; relaxed arithmetic, force-only or
//                      fused, launched for short rows (aggregates).  Further down, with its own header comment.
//
// Either one replaces, in ONE launch per time step,
//   LT/rotational_system/rotational_binary_system_neighbors_omp.h:78-98   (pair loop, ascending j)
//   LG/granular_system/binary_force_container.h:20-32                     (functor fold per pair)
//   LG/contact_force/contact_force.h:57-147, LG/hamaker_force/hamaker_force.h:28-43,
//   LG/sinter_bridge/sinter_bridge.h:214-259                              (force laws, via physics.cuh)
//   LT/rotational_integrator/rotational_velocity_verlet_half.h:63-73      (integrate, MODE_FUSED)
//   LG/sinter_bridge/sinter_bridge.h:24-42,55-75                          (connector update, MODE_FUSED)
//
// Organisation (one CTA of LG_TB threads owns `rows_per_block` consecutive particles):
//   stage   Everything the CTA reads that is CONTIGUOUS in HBM -- x, v, omega, theta of its particles, its slice of
//           the row / bond pointers, the columns and tags of its first tile of pairs, the connectors, rest lengths
//           and mirror indices of its particles' bond records -- is brought into shared memory by bulk asynchronous
//           copies (cp.async.bulk, the TMA engine) that ONE thread issues at kernel entry against two mbarriers.
//           The dependent chain row pointer -> column -> tag -> mirror index -> partner connector, which used to be
//           five serialized global round trips per CTA (profiles/README.md r01d), becomes two: the bulk stage and
//           the gathers of x_j (and of the few partner connectors owned by another CTA).
//   per tile of up to LG_TILE consecutive neighbor pairs of those rows:
//     phase A  one thread per PAIR: geometry, Hamaker, contact test (the cheap, uniform work);
//              pairs that need the expensive laws (bonded, overlapping) are compacted into a list with
//              warp ballots so that phase B runs with full lanes instead of diverging inside phase A
//     phase B  one thread per HEAVY pair: sinter-bridge springs or Luding contact with its history springs
//     phase C  one thread per (PARTICLE, component): adds its pairs' contributions from shared memory in
//              ascending-j order -- the reference's accumulation order, so a / alpha are bit-identical to it; no
//              atomics, deterministic
//   finalize unary (facet) forces, then either write a, alpha (MODE_FORCE) or integrate v, x, omega, theta
//            into the next-step buffers and move the particle's bridge connectors (MODE_FUSED), with
//            coalesced stores staged through shared memory.
// Particle state is never modified in place while other CTAs may still read it: x, v, omega and the
// connectors are double-buffered (A.x / A.x_next ...), swapped by the host after the launch.
#include "step_common.cuh"

namespace lg {

namespace {

// Tunables (compile-time; -DLG_TB=... builds an experimental library, see scripts/build_variants.sh)
#ifndef LG_TB
#define LG_TB 256
#endif
#ifndef LG_TILE
#define LG_TILE 512
#endif
#ifndef LG_RMAX
#define LG_RMAX 52
#endif
#ifndef LG_CCAP
#define LG_CCAP 112
#endif
#ifndef LG_MINB
#define LG_MINB 4
#endif
#ifndef LG_MINB_EXACT
#define LG_MINB_EXACT 3  // the reference-order arithmetic needs ~80 registers: 3 CTAs/SM measured faster than 4 with spills
#endif
#ifndef LG_BU
#define LG_BU 5  // unroll factor of the five bridge springs in the relaxed-arithmetic path
#endif
constexpr int TB = LG_TB;      // threads per CTA
constexpr int TILE = LG_TILE;  // pairs per tile
constexpr int RMAX = LG_RMAX;  // max particles per CTA (<= TB)
constexpr int CCAP = LG_CCAP;  // bond records per CTA staged in shared memory (more: the CTA reads them from global)
constexpr int TS = TILE + 1;   // padded component stride of the tile (bank spread for phase C)
constexpr int BU = LG_BU;     // unroll factor of the five bridge springs (relaxed-arithmetic path)
static_assert(RMAX <= TB && RMAX <= 256, "one staging pass per CTA; local rows are stored in a byte");
static_assert(TILE % TB == 0, "phase A handles TILE / TB pairs per thread");
static_assert(15 * RMAX <= 6 * TS, "the output staging area reuses the tile");
static_assert(RMAX % 4 == 0 && TILE % 16 == 0 && CCAP % 16 == 0, "staged windows are multiples of 16 bytes");

enum : int { KIND_LIGHT = 0, KIND_BOND = 1, KIND_CONTACT = 2 };

// A staged array is a 16-byte aligned WINDOW around the bytes the CTA needs (cp.async.bulk wants 16-byte aligned
// addresses and sizes): up to 15 bytes before and after are copied along; `lead` says where the payload starts.
// Every device buffer ends with >= 16 bytes of slack (api.cu: DevBuf), so the window never leaves the allocation.
struct alignas(16) Smem {
    double tile[6 * TS];             // a.x a.y a.z alpha.x alpha.y alpha.z, structure-of-arrays over the tile
    double wx[3 * RMAX + 4];         // windows: x, v, omega, theta of the CTA's particles
    double wv[3 * RMAX + 4];
    double ww[3 * RMAX + 4];
    double wth[3 * RMAX + 4];
    double wconn[15 * CCAP + 4];     // window: connectors of the CTA's bond records
    double wlen[5 * CCAP + 4];       // window: their rest lengths
    double acc[RMAX * 6];            // running a, alpha of the CTA's particles ([row][component])
    uint32_t wrow[RMAX + 12];        // window: row_ptr[i0 .. i0 + nrows]
    uint32_t wbond[RMAX + 12];       // window: bptr[i0 .. i0 + nrows]
    int32_t wcols[TILE + 8];         // window: columns of the first tile
    int32_t wrev[CCAP + 8];          // window: mirror record of every staged bond
    uint16_t heavy[TILE];
    uint8_t wtag[TILE + 32];         // window: tags of the first tile (snapshot at kernel entry)
    uint8_t srow[TILE];              // local row of every pair of the tile (written by the row owners)
    uint8_t brow[CCAP];              // local row of the CTA's staged bond records (connector stage)
    unsigned long long bar[2];       // mbarriers: [0] state + pointers, [1] pair / bond data
    int cnt[2];
};

template <int MASK, int MODE, bool FACETS, bool FAST>
__global__ void __launch_bounds__(TB, FAST ? LG_MINB : LG_MINB_EXACT) step_tiled_kernel(const __grid_constant__ ForceArgs A) {
    constexpr bool kContact = (MASK & (MODEL_CONTACT | MODEL_SINTER)) != 0;
    constexpr bool kSinter = (MASK & MODEL_SINTER) != 0;
    constexpr bool kHamaker = (MASK & MODEL_HAMAKER) != 0;
    constexpr int PPT = TILE / TB;  // pairs per thread and tile in phase A

    extern __shared__ __align__(16) unsigned char smem_raw[];
    Smem& sm = *reinterpret_cast<Smem*>(smem_raw);

    const int tid = threadIdx.x, lane = tid & 31;
    const int R = A.rows_per_block;
    const int i0 = blockIdx.x * R;
    const int nrows = min(R, A.n - i0);
    const bool have_bonds = kSinter && A.bptr != nullptr;

    // ---- stage: bulk asynchronous copies of every contiguous range the CTA reads ---------------------------
    const double* gx = A.x + 3 * (size_t)i0;
    const double* gv = A.v + 3 * (size_t)i0;
    const double* gw = A.w + 3 * (size_t)i0;
    const double* gth = A.th + 3 * (size_t)i0;
    const uint32_t* grow = A.row_ptr + i0;
    const uint32_t* gbond = have_bonds ? A.bptr + i0 : nullptr;
    if (tid == 0) {
        mbar_init(&sm.bar[0], 1);
        mbar_init(&sm.bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        const uint32_t nb3 = 24u * (uint32_t)nrows, nbp = 4u * (uint32_t)(nrows + 1);
        uint32_t total = window_bytes(gx, nb3) + window_bytes(gv, nb3) + window_bytes(gw, nb3) + window_bytes(grow, nbp);
        if constexpr (MODE == MODE_FUSED) total += window_bytes(gth, nb3);
        if (have_bonds) total += window_bytes(gbond, nbp);
        mbar_expect_tx(&sm.bar[0], total);
        bulk_load(sm.wx, gx, nb3, &sm.bar[0]);
        bulk_load(sm.wv, gv, nb3, &sm.bar[0]);
        bulk_load(sm.ww, gw, nb3, &sm.bar[0]);
        if constexpr (MODE == MODE_FUSED) bulk_load(sm.wth, gth, nb3, &sm.bar[0]);
        bulk_load(sm.wrow, grow, nbp, &sm.bar[0]);
        if (have_bonds) bulk_load(sm.wbond, gbond, nbp, &sm.bar[0]);
        // the ranges of pairs and bond records: four independent loads, one round trip, overlapped with the copies above
        const uint32_t rp0 = grow[0], rp1 = grow[nrows];
        const uint32_t bp0 = have_bonds ? gbond[0] : 0u, bp1 = have_bonds ? gbond[nrows] : 0u;
        const uint32_t np0 = min(rp1 - rp0, (uint32_t)TILE);           // pairs of the first tile
        const uint32_t nbr = (bp1 - bp0 <= (uint32_t)CCAP) ? bp1 - bp0 : 0u;  // staged bond records
        const int32_t* gcols = A.cols + rp0;
        const uint8_t* gtag = A.tag + rp0;
        const double* gconn = A.bconn + 15 * (size_t)bp0;
        const double* glen = A.blen + 5 * (size_t)bp0;
        const int32_t* grev = A.brev + bp0;
        uint32_t total1 = window_bytes(gcols, 4u * np0);
        if constexpr (kContact) total1 += window_bytes(gtag, np0);
        if (nbr) total1 += window_bytes(gconn, 120u * nbr) + window_bytes(glen, 40u * nbr) + window_bytes(grev, 4u * nbr);
        mbar_expect_tx(&sm.bar[1], total1);
        bulk_load(sm.wcols, gcols, 4u * np0, &sm.bar[1]);
        if constexpr (kContact) bulk_load(sm.wtag, gtag, np0, &sm.bar[1]);
        if (nbr) {
            bulk_load(sm.wconn, gconn, 120u * nbr, &sm.bar[1]);
            bulk_load(sm.wlen, glen, 40u * nbr, &sm.bar[1]);
            bulk_load(sm.wrev, grev, 4u * nbr, &sm.bar[1]);
        }
        sm.cnt[0] = sm.cnt[1] = 0;
    }
    for (int idx = tid; idx < 6 * nrows; idx += TB) sm.acc[idx] = 0.0;
    __syncthreads();  // the mbarriers are initialised
    mbar_wait(&sm.bar[0], 0);

    // payloads inside the windows
    const double* sx = sm.wx + (window_lead(gx) >> 3);
    const double* sv = sm.wv + (window_lead(gv) >> 3);
    const double* sw = sm.ww + (window_lead(gw) >> 3);
    const double* sth = sm.wth + (window_lead(gth) >> 3);
    const uint32_t* sptr = sm.wrow + (window_lead(grow) >> 2);
    const uint32_t* sbptr = sm.wbond + (window_lead(gbond) >> 2);  // valid only when have_bonds
    const uint32_t p_begin = sptr[0], p_end = sptr[nrows];
    const uint32_t b0 = have_bonds ? sbptr[0] : 0u;
    const uint32_t nbond = have_bonds ? sbptr[nrows] - b0 : 0u;
    const bool conn_staged = nbond > 0 && nbond <= (uint32_t)CCAP;
    const int32_t* scols = sm.wcols + (window_lead(A.cols + p_begin) >> 2);
    const uint8_t* stag = sm.wtag + window_lead(A.tag + p_begin);
    const double* sconn = sm.wconn + (window_lead(A.bconn + 15 * (size_t)b0) >> 3);
    const double* slen = sm.wlen + (window_lead(A.blen + 5 * (size_t)b0) >> 3);
    const int32_t* srev = sm.wrev + (window_lead(A.brev + b0) >> 2);

    if (tid < nrows) {
        // local-row map of the first tile, and of this particle's staged bond records (connector stage)
        const uint32_t lo = sptr[tid] - p_begin, hi = min(sptr[tid + 1] - p_begin, (uint32_t)TILE);
        for (uint32_t q = lo; q < hi; q++) sm.srow[q] = (uint8_t)tid;
        if (MODE == MODE_FUSED && conn_staged)
            for (uint32_t b = sbptr[tid]; b < sbptr[tid + 1]; b++) sm.brow[b - b0] = (uint8_t)tid;
    }
    mbar_wait(&sm.bar[1], 0);
    __syncthreads();  // srow / brow

    auto particle = [&](int r, int which) -> d3 { return load3((which == 0 ? sx : which == 1 ? sv : sw) + 3 * r); };
    auto neighbour = [&](int j, int which, const double* gsrc) -> d3 {
        const unsigned rj = (unsigned)(j - i0);
        return rj < (unsigned)nrows ? load3((which == 0 ? sx : which == 1 ? sv : sw) + 3 * rj) : load3(gsrc + 3 * (size_t)j);
    };

    for (uint32_t tile_start = p_begin; tile_start < p_end; tile_start += TILE) {
        const int cntp = (int)min((uint32_t)TILE, p_end - tile_start);
        const bool first = tile_start == p_begin;  // columns and tags of the first tile are staged

        // ---- phase A: light work for every pair, classify ------------------------------------------------
        // All of a thread's PPT pairs are loaded before any is computed (PPT gathers in flight per thread), and
        // the straight-line geometry + Hamaker arithmetic of the PPT pairs is independent (instruction-level
        // parallelism for the FP64 pipe); classification, which branches, comes last.
        int jv[PPT], tv[PPT];
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            const int q = tid + k * TB;
            jv[k] = -1;
            tv[k] = kTagNone;
            if (q < cntp) {
                jv[k] = first ? scols[q] : A.cols[tile_start + q];
                if constexpr (kContact) tv[k] = first ? stag[q] : A.tag[tile_start + q];
            }
        }
        d3 xjv[PPT];
#pragma unroll
        for (int k = 0; k < PPT; k++) xjv[k] = jv[k] >= 0 ? neighbour(jv[k], 0, A.x) : zero3();
        d3 hv[PPT];
        double ov[PPT];  // contact overlap (exact mode: the reference's value; fast mode: 2r - len, refined below)
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            const int q = tid + k * TB;
            const int r = q < cntp ? sm.srow[q] : 0;
            const d3 xi = particle(r, 0);
            hv[k] = zero3();
            if constexpr (FAST) {
                const PairGeom g = pair_geometry_fast(xi, xjv[k]);
                ov[k] = A.fc.two_r - g.len;
                if constexpr (kHamaker) hv[k] = hamaker_accel_fast(A.fc, g);
            } else {
                const PairGeom g = pair_geometry(xi, xjv[k]);
                ov[k] = contact_overlap(g, 2.0 * A.mp.r);
                if constexpr (kHamaker) hv[k] = hamaker_accel(A.mp.hamaker_A, A.mp.hamaker_h0, A.mp.r, A.mp.mass, g, A.mp.hamaker_kind, A.mp.hamaker_cutoff);
            }
        }
#pragma unroll
        for (int k = 0; k < PPT; k++) {
            const int q = tid + k * TB;
            int kind = KIND_LIGHT;
            if (q < cntp) {
                if constexpr (kContact) {
                    const int t = tv[k];
                    if (kSinter && t >= kTagBond && t != kTagNone) {
                        kind = KIND_BOND;
                    } else {
                        double overlap = ov[k];
                        if constexpr (FAST) {  // inside a rounding band of zero: decide on the exact expression
                            if (fabs(overlap) < A.fc.band)
                                overlap = contact_overlap(pair_geometry(particle(sm.srow[q], 0), xjv[k]), A.fc.two_r);
                        }
                        if (overlap <= 0) {
                            if (t != kTagNone) {  // reset_springs (contact_force.h:68-71)
                                A.sgid[(size_t)(i0 + sm.srow[q]) * A.S + t] = -1;
                                A.tag[tile_start + q] = (uint8_t)kTagNone;
                            }
                        } else {
                            kind = KIND_CONTACT;
                        }
                    }
                }
                sm.tile[0 * TS + q] = hv[k].x;
                sm.tile[1 * TS + q] = hv[k].y;
                sm.tile[2 * TS + q] = hv[k].z;
                sm.tile[3 * TS + q] = 0.0;
                sm.tile[4 * TS + q] = 0.0;
                sm.tile[5 * TS + q] = 0.0;
            }
            if constexpr (kContact) {
                const unsigned mb = __ballot_sync(0xffffffffu, kind == KIND_BOND);
                const unsigned mc = __ballot_sync(0xffffffffu, kind == KIND_CONTACT);
                if (mb | mc) {
                    int base_b = 0, base_c = 0;
                    if (lane == 0) {
                        if (mb) base_b = atomicAdd(&sm.cnt[0], __popc(mb));  // list position only; results unaffected
                        if (mc) base_c = atomicAdd(&sm.cnt[1], __popc(mc));
                    }
                    base_b = __shfl_sync(0xffffffffu, base_b, 0);
                    base_c = __shfl_sync(0xffffffffu, base_c, 0);
                    const unsigned lt = (1u << lane) - 1u;
                    if (kind == KIND_BOND) sm.heavy[base_b + __popc(mb & lt)] = (uint16_t)q;
                    if (kind == KIND_CONTACT) sm.heavy[TILE - 1 - (base_c + __popc(mc & lt))] = (uint16_t)q;
                }
            }
        }
        __syncthreads();

        // ---- phase B: bonded and overlapping pairs, densely packed -----------------------------------------
        if constexpr (kContact) {
            const int nB = sm.cnt[0], nC = sm.cnt[1];
            for (int e = tid; e < nB + nC; e += TB) {
                const bool is_bond = e < nB;
                const int q = is_bond ? sm.heavy[e] : sm.heavy[TILE - 1 - (e - nB)];
                const uint32_t p = tile_start + q;
                const int r = sm.srow[q];
                const int j = first ? scols[q] : A.cols[p];
                Accel c;
                if (kSinter && is_bond) {
                    const int t = first ? stag[q] : A.tag[p];
                    const uint32_t b = sbptr[r] + (uint32_t)(t - kTagBond);
                    const double *ci, *cj, *len;
                    if (conn_staged) {  // own record from shared memory; the mirror too when this CTA owns the partner
                        const uint32_t bl = b - b0, ml = (uint32_t)srev[bl] - b0;
                        ci = sconn + 15 * bl;
                        len = slen + 5 * bl;
                        cj = ml < nbond ? sconn + 15 * ml : A.bconn + (size_t)srev[bl] * 15;
                    } else {
                        ci = A.bconn + (size_t)b * 15;
                        cj = A.bconn + (size_t)A.brev[b] * 15;
                        len = A.blen + (size_t)b * 5;
                    }
                    const d3 xi = particle(r, 0), xj = neighbour(j, 0, A.x);
                    d3 force = zero3(), torque = zero3();
                    if constexpr (FAST) {
                        if (A.mp.gamma_d != 0.0) {
                            const d3 vi = particle(r, 1), wi = particle(r, 2);
                            const d3 vj = neighbour(j, 1, A.v), wj = neighbour(j, 2, A.w);
#pragma unroll BU
                            for (int k = 0; k < 5; k++)
                                bonded_spring_fast<true>(A.mp.k_d, A.mp.gamma_d, len[k], load3(ci + 3 * k), load3(cj + 3 * k), xi,
                                                         xj, vi, vj, wi, wj, force, torque);
                        } else {  // undamped bridges (the reference's tests): velocities are not needed at all
#pragma unroll BU
                            for (int k = 0; k < 5; k++)
                                bonded_spring_fast<false>(A.mp.k_d, 0.0, len[k], load3(ci + 3 * k), load3(cj + 3 * k), xi, xj,
                                                          zero3(), zero3(), zero3(), zero3(), force, torque);
                        }
                        c = Accel{force * A.fc.inv_mass, torque * A.fc.inv_inertia};
                    } else {
                        const d3 vi = particle(r, 1), wi = particle(r, 2);
                        const d3 vj = neighbour(j, 1, A.v), wj = neighbour(j, 2, A.w);
#pragma unroll 1
                        for (int k = 0; k < 5; k++)
                            bonded_spring(A.mp.k_d, A.mp.gamma_d, len[k], load3(ci + 3 * k), load3(cj + 3 * k), xi, xj, vi, vj,
                                          wi, wj, force, torque);
                        c = Accel{force / A.mp.mass, torque / A.mp.inertia};
                    }
                } else {
                    double out6[6];
                    contact_heavy(A, i0 + r, j, p, out6);
                    c = Accel{load3(out6), load3(out6 + 3)};
                }
                if constexpr (kHamaker) {  // pair value = (contact | sinter)(i,j) + hamaker(i,j)
                    sm.tile[0 * TS + q] = c.a.x + sm.tile[0 * TS + q];
                    sm.tile[1 * TS + q] = c.a.y + sm.tile[1 * TS + q];
                    sm.tile[2 * TS + q] = c.a.z + sm.tile[2 * TS + q];
                } else {
                    sm.tile[0 * TS + q] = c.a.x;
                    sm.tile[1 * TS + q] = c.a.y;
                    sm.tile[2 * TS + q] = c.a.z;
                }
                sm.tile[3 * TS + q] = c.alpha.x;
                sm.tile[4 * TS + q] = c.alpha.y;
                sm.tile[5 * TS + q] = c.alpha.z;
            }
            __syncthreads();
        }

        // ---- phase C: ordered accumulation (ascending j), one thread per (particle, component) ----------------
        for (int wk = tid; wk < 6 * nrows; wk += TB) {
            const int r = wk / 6, c = wk - 6 * r;
            const uint32_t lo = max(sptr[r], tile_start) - tile_start;
            const uint32_t hi = min(sptr[r + 1], tile_start + (uint32_t)cntp) - tile_start;
            if (lo < hi && sptr[r + 1] > tile_start) {
                const double* t = sm.tile + c * TS;
                double a = sm.acc[wk];
                for (uint32_t q = lo; q < hi; q++) a += t[q];
                sm.acc[wk] = a;
            }
        }
        if (tile_start + TILE < p_end) {  // local-row map of the next tile (srow is not read in phase C)
            const uint32_t nxt = tile_start + TILE;
            if (tid < nrows) {
                const uint32_t r0 = sptr[tid], r1 = sptr[tid + 1];
                if (r1 > nxt && r0 < nxt + TILE) {
                    const uint32_t lo = max(r0, nxt) - nxt, hi = min(r1 - nxt, (uint32_t)TILE);
                    for (uint32_t q = lo; q < hi; q++) sm.srow[q] = (uint8_t)tid;
                }
            }
        }
        if (tid == 0) sm.cnt[0] = sm.cnt[1] = 0;
        __syncthreads();
    }

    // ---- finalize: unary forces, then output through shared memory (the tile area is free now) --------------
    double* out = sm.tile;  // [0,3R) [3R,6R) ... blocks of 3 * RMAX doubles
    constexpr int OB = 3 * RMAX;
    if (tid < nrows) {
        const d3 xi = particle(tid, 0), vi = particle(tid, 1), wi = particle(tid, 2);
        d3 acc_a = load3(sm.acc + 6 * tid), acc_al = load3(sm.acc + 6 * tid + 3);
        if constexpr (FACETS) {
            const Accel u = unary_forces(A, i0 + tid, xi, vi, wi);
            acc_a += u.a;
            acc_al += u.alpha;
        }
        if constexpr (MODE == MODE_FORCE) {
            store3(out + 0 * OB + 3 * tid, acc_a);
            store3(out + 1 * OB + 3 * tid, acc_al);
        } else {
            const double dt = A.dt;
            const d3 v = vi + acc_a * dt;    // increment_v(n, a * dt)
            const d3 dx = v * dt;            // increment_x(n, v * dt), v already updated
            const d3 x = xi + dx;
            const d3 w = wi + acc_al * dt;   // increment_omega
            const d3 dth = w * dt;           // increment_theta, omega already updated
            store3(out + 0 * OB + 3 * tid, x);
            store3(out + 1 * OB + 3 * tid, v);
            store3(out + 2 * OB + 3 * tid, w);
            store3(out + 3 * OB + 3 * tid, dth);
            store3(out + 4 * OB + 3 * tid, dx);
        }
    }
    __syncthreads();
    for (int idx = tid; idx < 3 * nrows; idx += TB) {
        const size_t gidx = 3 * (size_t)i0 + idx;
        if constexpr (MODE == MODE_FORCE) {
            A.a[gidx] = out[0 * OB + idx];
            A.al[gidx] = out[1 * OB + idx];
        } else {
            A.x_next[gidx] = out[0 * OB + idx];
            A.v_next[gidx] = out[1 * OB + idx];
            A.w_next[gidx] = out[2 * OB + idx];
            A.th[gidx] = sth[idx] + out[3 * OB + idx];
        }
    }
    if constexpr (MODE == MODE_FUSED && kSinter) {
        if (have_bonds) {
            // connectors of the CTA's particles: c += dx ; c += dtheta x (c - x_new)   (sinter_bridge.h:36-41,67-74)
            const int ncon = (int)nbond * 5;
            for (int cq = tid; cq < ncon; cq += TB) {
                const int bl = cq / 5;
                const int r = conn_staged ? (int)sm.brow[bl] : find_row(sbptr, nrows, b0 + (uint32_t)bl);
                const size_t off = (size_t)b0 * 15 + 3 * (size_t)cq;
                d3 ck = (conn_staged ? load3(sconn + 3 * cq) : load3(A.bconn + off)) + load3(out + 4 * OB + 3 * r);
                const d3 arm = ck - load3(out + 0 * OB + 3 * r);
                ck = ck + cross(load3(out + 3 * OB + 3 * r), arm);
                store3(A.bconn_next + off, ck);
            }
        }
    }
}

// ======================================================================================================================
// step_rows_kernel -- the relaxed-arithmetic (LG_MATH_FAST) fused step for SHORT rows (aggregates: ~10 neighbors).
//
// Same inputs, same outputs and the same staging as step_tiled_kernel, but the work is cut the other way round:
//   pass 1  one thread per BOND RECORD of the CTA's particles: the five bridge springs (full lanes, no compaction);
//           the (force / m, torque / I) of record b waits in shared memory for the row pass
//   pass 2  one thread per PARTICLE walks its row in ascending j with the accumulators in registers: geometry +
//           Hamaker per pair, the bonded value looked up, Luding contact (rare for aggregates) evaluated in place
//   pass 3  integrate, coalesced stores through the staged windows, connector update
// There is no pair tile, no row map, no ballot compaction and no ordered shared-memory reduction: ~2.4x fewer
// instructions per particle than the tile kernel (profiles/README.md r01f).  Accumulation order: ascending j within
// the pairs the hot loop evaluates, then ascending j within the pairs it only marked (contacts, rounding band,
// unstaged bridges; rows_slow_pairs) -- a fixed order (deterministic), but not the reference's single ascending-j sweep
// whenever a marked pair exists; the pair value is formed as (sinter | contact) + hamaker as in the reference.  The
// arithmetic inside the laws is the relaxed one.  MODE_FORCE writes a, alpha and stops (lg_get_field(LG_A), first-step
// evaluations); MODE_FUSED integrates.  Rows of a dense packing (~80 neighbors, several contacts per
// particle) would serialise 80 iterations per thread and diverge on the contact law, so the host launches this kernel
// only when the mean row length is small (launch_step_tiled: rows_kernel_suits).
#ifndef LG_ROWS_TB
#define LG_ROWS_TB 128
#endif
#ifndef LG_ROWS_MINB
#define LG_ROWS_MINB 4
#endif
constexpr int RT = LG_ROWS_TB;   // threads = particles per CTA
constexpr int RPC = 12 * RT;     // staged columns / tags per CTA (pairs beyond are read from global memory)
constexpr int RBC = 2 * RT + 32; // bond records per CTA whose results are kept in shared memory
constexpr int RH = RT;           // x is staged for RH particles before and after the CTA's own: neighbours of an
                                 // aggregate are its own ~150 consecutive particles, so nearly every x_j is a shared-memory hit
static_assert(RT % 32 == 0 && RPC % 16 == 0 && RBC % 16 == 0, "window sizes");

struct alignas(16) RowsSmem {
    double wx[3 * (RT + 2 * RH) + 4];  // window: x of the CTA's particles and of RH particles either side
                              // windows: v, omega, theta of the CTA's particles; pass 3 reuses all four for the outputs
    double wv[3 * RT + 4];
    double ww[3 * RT + 4];
    double wth[3 * RT + 4];
    double res[6 * RBC];      // pass 1 results per bond record; pass 3: dx, dtheta of every particle
    uint32_t wrow[RT + 12];   // window: row_ptr[i0 .. i0 + nrows]
    uint32_t wbond[RT + 12];  // window: bptr[i0 .. i0 + nrows]
    int32_t wcols[RPC + 8];   // window: columns of the CTA's first RPC pairs
    uint8_t wtag[RPC + 32];   // window: their tags (snapshot at kernel entry)
    uint8_t brow[RBC];        // local row of every staged bond record
    unsigned long long bar[2];
};

template <int MASK, int MODE, bool FACETS>
__global__ void __launch_bounds__(RT, LG_ROWS_MINB) step_rows_kernel(const __grid_constant__ ForceArgs A) {
    constexpr bool kContact = (MASK & (MODEL_CONTACT | MODEL_SINTER)) != 0;
    constexpr bool kSinter = (MASK & MODEL_SINTER) != 0;
    constexpr bool kHamaker = (MASK & MODEL_HAMAKER) != 0;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    RowsSmem& sm = *reinterpret_cast<RowsSmem*>(smem_raw);

    const int tid = threadIdx.x;
    const int i0 = blockIdx.x * RT;
    const int nrows = min(RT, A.n - i0);
    const bool have_bonds = kSinter && A.bptr != nullptr;

    // ---- stage (as step_tiled_kernel): state, pointer slices | columns and tags of the CTA's rows ------------
    const int xlo = max(i0 - RH, 0), xcnt = min(A.n, i0 + RT + RH) - xlo;  // staged x range
    const double* gx = A.x + 3 * (size_t)xlo;
    const double* gv = A.v + 3 * (size_t)i0;
    const double* gw = A.w + 3 * (size_t)i0;
    const double* gth = A.th + 3 * (size_t)i0;
    const uint32_t* grow = A.row_ptr + i0;
    const uint32_t* gbond = have_bonds ? A.bptr + i0 : nullptr;
    if (tid == 0) {
        mbar_init(&sm.bar[0], 1);
        mbar_init(&sm.bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        const uint32_t nb3 = 24u * (uint32_t)nrows, nbp = 4u * (uint32_t)(nrows + 1), nbx = 24u * (uint32_t)xcnt;
        uint32_t total = window_bytes(gx, nbx) + window_bytes(gv, nb3) + window_bytes(gw, nb3) + window_bytes(gth, nb3) +
                         window_bytes(grow, nbp);
        if (have_bonds) total += window_bytes(gbond, nbp);
        mbar_expect_tx(&sm.bar[0], total);
        bulk_load(sm.wx, gx, nbx, &sm.bar[0]);
        bulk_load(sm.wv, gv, nb3, &sm.bar[0]);
        bulk_load(sm.ww, gw, nb3, &sm.bar[0]);
        bulk_load(sm.wth, gth, nb3, &sm.bar[0]);
        bulk_load(sm.wrow, grow, nbp, &sm.bar[0]);
        if (have_bonds) bulk_load(sm.wbond, gbond, nbp, &sm.bar[0]);
        const uint32_t rp0 = grow[0], rp1 = grow[nrows];
        const uint32_t np0 = min(rp1 - rp0, (uint32_t)RPC);
        const int32_t* gcols = A.cols + rp0;
        const uint8_t* gtag = A.tag + rp0;
        uint32_t total1 = window_bytes(gcols, 4u * np0);
        if constexpr (kContact) total1 += window_bytes(gtag, np0);
        mbar_expect_tx(&sm.bar[1], total1);
        bulk_load(sm.wcols, gcols, 4u * np0, &sm.bar[1]);
        if constexpr (kContact) bulk_load(sm.wtag, gtag, np0, &sm.bar[1]);
    }
    // The first bond record of this thread (record b0 + tid) does not depend on anything staged: its indices, both
    // connector sets and the rest lengths are requested NOW, so that their L2 round trips (indices, then connectors)
    // run underneath the bulk copies instead of after them.
    double e_ci[15], e_cj[15], e_len[5];
    int e_j = 0, e_j1 = 0, e_rev1 = -1;  // e_rev1 >= 0: the indices of the thread's second record (b0 + tid + RT) are loaded too
    bool e_have = false;
    if constexpr (kSinter) {
        if (have_bonds) {
            const uint32_t eb0 = gbond[0], enb = gbond[nrows] - eb0;
            if (enb <= (uint32_t)RBC && (uint32_t)tid < enb) {
                e_have = true;
                const uint32_t b = eb0 + (uint32_t)tid;
                e_j = A.bj[b];
                const double* ci = A.bconn + (size_t)b * 15;
                const double* cj = A.bconn + (size_t)A.brev[b] * 15;
                const double* len = A.blen + (size_t)b * 5;
#pragma unroll
                for (int k = 0; k < 15; k++) {
                    e_ci[k] = ci[k];
                    e_cj[k] = cj[k];
                }
#pragma unroll
                for (int k = 0; k < 5; k++) e_len[k] = len[k];
                if ((uint32_t)tid + (uint32_t)RT < enb) {
                    e_j1 = A.bj[b + RT];
                    e_rev1 = A.brev[b + RT];
                }
            }
        }
    }
    __syncthreads();  // the mbarriers are initialised
    mbar_wait(&sm.bar[0], 0);

    double* sxw = sm.wx + (window_lead(gx) >> 3);  // x of particle xlo + k at sxw[3 k]
    double* sx = sxw + 3 * (i0 - xlo);             // the CTA's own particles
    double* sv = sm.wv + (window_lead(gv) >> 3);
    double* sw = sm.ww + (window_lead(gw) >> 3);
    double* sth = sm.wth + (window_lead(gth) >> 3);
    const uint32_t* sptr = sm.wrow + (window_lead(grow) >> 2);
    const uint32_t* sbptr = sm.wbond + (window_lead(gbond) >> 2);  // valid only when have_bonds
    const uint32_t p_begin = sptr[0];
    const uint32_t b0 = have_bonds ? sbptr[0] : 0u;
    const uint32_t nbond = have_bonds ? sbptr[nrows] - b0 : 0u;
    const bool bonds_staged = nbond > 0 && nbond <= (uint32_t)RBC;
    const int32_t* scols = sm.wcols + (window_lead(A.cols + p_begin) >> 2);
    const uint8_t* stag = sm.wtag + window_lead(A.tag + p_begin);

    auto neighbour = [&](int j, const double* s, const double* gsrc) -> d3 {  // v, omega: own particles staged
        const unsigned rj = (unsigned)(j - i0);
        return rj < (unsigned)nrows ? load3(s + 3 * rj) : load3(gsrc + 3 * (size_t)j);
    };
    auto neighbour_x = [&](int j) -> d3 {
        const unsigned rj = (unsigned)(j - xlo);
        return rj < (unsigned)xcnt ? load3(sxw + 3 * rj) : load3(A.x + 3 * (size_t)j);
    };

    // ---- pass 1: one thread per bond record ---------------------------------------------------------------
    if constexpr (kSinter) {
        if (bonds_staged) {
            if (tid < nrows)
                for (uint32_t b = sbptr[tid]; b < sbptr[tid + 1]; b++) sm.brow[b - b0] = (uint8_t)tid;
            __syncthreads();
            if (e_have) {  // record b0 + tid, loaded early (e_have implies bonds_staged and tid < nbond)
                const int r = sm.brow[tid];
                const d3 xi = load3(sx + 3 * r), xj = neighbour_x(e_j);
                d3 fa, ta;
                if (A.mp.gamma_d != 0.0)
                    bridge_fast<true>(A, e_ci, e_cj, e_len, xi, xj, load3(sv + 3 * r), neighbour(e_j, sv, A.v), load3(sw + 3 * r),
                                      neighbour(e_j, sw, A.w), fa, ta);
                else
                    bridge_fast<false>(A, e_ci, e_cj, e_len, xi, xj, zero3(), zero3(), zero3(), zero3(), fa, ta);
                store3(sm.res + 6 * tid, fa);
                store3(sm.res + 6 * tid + 3, ta);
            }
            for (uint32_t bl = tid + RT; bl < nbond; bl += RT) {
                const uint32_t b = b0 + bl;
                const int r = sm.brow[bl];
                const bool early = e_rev1 >= 0 && bl == (uint32_t)tid + (uint32_t)RT;
                const int j = early ? e_j1 : A.bj[b];
                const double* ci = A.bconn + (size_t)b * 15;
                const double* cj = A.bconn + (size_t)(early ? e_rev1 : A.brev[b]) * 15;
                const double* len = A.blen + (size_t)b * 5;
                const d3 xi = load3(sx + 3 * r), xj = neighbour_x(j);
                d3 fa, ta;
                if (A.mp.gamma_d != 0.0)
                    bridge_fast<true>(A, ci, cj, len, xi, xj, load3(sv + 3 * r), neighbour(j, sv, A.v), load3(sw + 3 * r),
                                      neighbour(j, sw, A.w), fa, ta);
                else
                    bridge_fast<false>(A, ci, cj, len, xi, xj, zero3(), zero3(), zero3(), zero3(), fa, ta);
                store3(sm.res + 6 * bl, fa);
                store3(sm.res + 6 * bl + 3, ta);
            }
        }
    }
    mbar_wait(&sm.bar[1], 0);
    __syncthreads();  // pass 1 results

    // ---- pass 2: one thread per particle, its row in ascending j ---------------------------------------------
    d3 xn = zero3(), vn = zero3(), wn = zero3(), dth = zero3(), dxv = zero3();
    if (tid < nrows) {
        const d3 xi = load3(sx + 3 * tid), vi = load3(sv + 3 * tid), wi = load3(sw + 3 * tid);
        d3 acc_a = zero3(), acc_al = zero3();
        const uint32_t pb = sptr[tid], pe = sptr[tid + 1];
        // Pairs that need an out-of-line law (Luding contact; a bridge whose record is not staged) are only MARKED in
        // the hot loop and evaluated afterwards: a call inside the loop would force its live values into local memory.
        unsigned long long slow = 0ull;  // bit k: pair pb + k (k < 64) is evaluated by the second loop
        for (uint32_t p = pb; p < pe; p++) {
            const uint32_t ps = p - p_begin;
            const int j = ps < (uint32_t)RPC ? scols[ps] : A.cols[p];
            const d3 xj = neighbour_x(j);
            const PairGeom g = pair_geometry_fast(xi, xj);
            d3 pa = zero3();
            if constexpr (kHamaker) pa = hamaker_accel_fast(A.fc, g);
            if constexpr (kContact) {
                const int t = ps < (uint32_t)RPC ? stag[ps] : A.tag[p];
                if (kSinter && t >= kTagBond && t != kTagNone) {
                    if (bonds_staged) {
                        const uint32_t bl = sbptr[tid] + (uint32_t)(t - kTagBond) - b0;
                        const d3 sa = load3(sm.res + 6 * bl);
                        pa = kHamaker ? sa + pa : sa;  // pair value = sinter(i,j) + hamaker(i,j)
                        acc_al += load3(sm.res + 6 * bl + 3);
                    } else {
                        if (p - pb < 64u) slow |= 1ull << (p - pb);
                        continue;
                    }
                } else {
                    const double overlap = A.fc.two_r - g.len;
                    if (overlap <= -A.fc.band) {  // clearly apart
                        if (t != kTagNone) {  // reset_springs (contact_force.h:68-71)
                            A.sgid[(size_t)(i0 + tid) * A.S + t] = -1;
                            A.tag[p] = (uint8_t)kTagNone;
                        }
                    } else {  // overlapping, or inside the rounding band of zero (decided on the exact expression below)
                        if (p - pb < 64u) slow |= 1ull << (p - pb);
                        continue;
                    }
                }
            }
            acc_a += pa;
        }
        if constexpr (kContact) {
            if (slow != 0ull || pe - pb > 64u) {  // rare: out of line, so that the hot loop keeps its registers
                double out6[6];
                rows_slow_pairs<kSinter, kHamaker>(A, sxw, xlo, xcnt, tid + (i0 - xlo), pb, pe, pb, 1u, slow, bonds_staged, have_bonds ? sbptr[tid] : 0u, out6);
                acc_a += load3(out6);
                acc_al += load3(out6 + 3);
            }
        }
        if constexpr (FACETS) {
            const Accel u = unary_forces(A, i0 + tid, xi, vi, wi);
            acc_a += u.a;
            acc_al += u.alpha;
        }
        if constexpr (MODE == MODE_FORCE) {  // force only: a, alpha leave through the (now free) v / omega windows
            vn = acc_a;
            wn = acc_al;
        } else {
            const double dt = A.dt;
            vn = vi + acc_a * dt;   // increment_v(n, a * dt)
            dxv = vn * dt;          // increment_x(n, v * dt), v already updated
            xn = xi + dxv;
            wn = wi + acc_al * dt;  // increment_omega
            dth = wn * dt;          // increment_theta, omega already updated
        }
    }
    __syncthreads();  // every read of the staged state and of the pass-1 results is done

    if constexpr (MODE == MODE_FORCE) {
        if (tid < nrows) {
            store3(sv + 3 * tid, vn);
            store3(sw + 3 * tid, wn);
        }
        __syncthreads();
        for (int idx = tid; idx < 3 * nrows; idx += RT) {
            const size_t gidx = 3 * (size_t)i0 + idx;
            A.a[gidx] = sv[idx];
            A.al[gidx] = sw[idx];
        }
        return;
    }

    // ---- pass 3: outputs through the windows (coalesced stores), connector update -----------------------------
    if (tid < nrows) {
        const d3 thn = load3(sth + 3 * tid) + dth;
        store3(sx + 3 * tid, xn);
        store3(sv + 3 * tid, vn);
        store3(sw + 3 * tid, wn);
        store3(sth + 3 * tid, thn);
        store3(sm.res + 6 * tid, dxv);
        store3(sm.res + 6 * tid + 3, dth);
    }
    __syncthreads();
    for (int idx = tid; idx < 3 * nrows; idx += RT) {
        const size_t gidx = 3 * (size_t)i0 + idx;
        A.x_next[gidx] = sx[idx];
        A.v_next[gidx] = sv[idx];
        A.w_next[gidx] = sw[idx];
        A.th[gidx] = sth[idx];
    }
    if constexpr (kSinter) {
        if (have_bonds) {
            // connectors of the CTA's particles: c += dx ; c += dtheta x (c - x_new)   (sinter_bridge.h:36-41,67-74)
            const int ncon = (int)nbond * 5;
            for (int cq = tid; cq < ncon; cq += RT) {
                const int bl = cq / 5;
                const int r = bonds_staged ? (int)sm.brow[bl] : find_row(sbptr, nrows, b0 + (uint32_t)bl);
                const size_t off = (size_t)b0 * 15 + 3 * (size_t)cq;
                d3 ck = load3(A.bconn + off) + load3(sm.res + 6 * r);
                const d3 arm = ck - load3(sx + 3 * r);
                ck = ck + cross(load3(sm.res + 6 * r + 3), arm);
                store3(A.bconn_next + off, ck);
            }
        }
    }
}

template <int MASK, int MODE, bool FACETS>
void launch_rows_one(const ForceArgs& args, cudaStream_t st) {
    static bool configured = false;  // per instantiation
    if (!configured) {
        cudaFuncSetAttribute(step_rows_kernel<MASK, MODE, FACETS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RowsSmem));
        configured = true;
    }
    step_rows_kernel<MASK, MODE, FACETS><<<(args.n + RT - 1) / RT, RT, sizeof(RowsSmem), st>>>(args);
    g_launch_count++;
}

template <int MASK, int MODE>
void launch_rows_facets(const ForceArgs& args, cudaStream_t st) {
    if (args.n_facets > 0) launch_rows_one<MASK, MODE, true>(args, st); else launch_rows_one<MASK, MODE, false>(args, st);
}

template <int MODE>
void launch_rows(const ForceArgs& args, int mask, cudaStream_t st) {
    switch (mask & (MODEL_CONTACT | MODEL_HAMAKER | MODEL_SINTER)) {
        case MODEL_CONTACT: launch_rows_facets<MODEL_CONTACT, MODE>(args, st); break;
        case MODEL_CONTACT | MODEL_HAMAKER: launch_rows_facets<MODEL_CONTACT | MODEL_HAMAKER, MODE>(args, st); break;
        case MODEL_SINTER: launch_rows_facets<MODEL_SINTER, MODE>(args, st); break;
        case MODEL_SINTER | MODEL_HAMAKER: launch_rows_facets<MODEL_SINTER | MODEL_HAMAKER, MODE>(args, st); break;
        case MODEL_HAMAKER: launch_rows_facets<MODEL_HAMAKER, MODE>(args, st); break;
        default: launch_rows_facets<0, MODE>(args, st); break;
    }
}

template <int MASK, int MODE, bool FACETS, bool FAST>
void launch_one(const ForceArgs& args, cudaStream_t st) {
    static bool configured = false;  // per instantiation
    if (!configured) {
        cudaFuncSetAttribute(step_tiled_kernel<MASK, MODE, FACETS, FAST>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)sizeof(Smem));
        configured = true;
    }
    const int R = args.rows_per_block;
    const int blocks = (args.n + R - 1) / R;
    step_tiled_kernel<MASK, MODE, FACETS, FAST><<<blocks, TB, sizeof(Smem), st>>>(args);
    g_launch_count++;
}

template <int MASK, int MODE>
void launch_facets(const ForceArgs& args, cudaStream_t st) {
    if (args.n_facets > 0) {
        if (args.fast) launch_one<MASK, MODE, true, true>(args, st); else launch_one<MASK, MODE, true, false>(args, st);
    } else {
        if (args.fast) launch_one<MASK, MODE, false, true>(args, st); else launch_one<MASK, MODE, false, false>(args, st);
    }
}

template <int MODE>
void launch_mode(const ForceArgs& args, int mask, cudaStream_t st) {
    switch (mask & (MODEL_CONTACT | MODEL_HAMAKER | MODEL_SINTER)) {
        case MODEL_CONTACT: launch_facets<MODEL_CONTACT, MODE>(args, st); break;
        case MODEL_CONTACT | MODEL_HAMAKER: launch_facets<MODEL_CONTACT | MODEL_HAMAKER, MODE>(args, st); break;
        case MODEL_SINTER: launch_facets<MODEL_SINTER, MODE>(args, st); break;
        case MODEL_SINTER | MODEL_HAMAKER: launch_facets<MODEL_SINTER | MODEL_HAMAKER, MODE>(args, st); break;
        case MODEL_HAMAKER: launch_facets<MODEL_HAMAKER, MODE>(args, st); break;
        default: launch_facets<0, MODE>(args, st); break;
    }
}

}  // namespace

int tiled_max_rows_per_block() { return RMAX; }
int tiled_tile_pairs() { return TILE; }

// fused == false: a, alpha <- forces at (x, v, omega)                    [MODE_FORCE]
// fused == true : additionally integrates into x_next, v_next, w_next, th, bconn_next  [MODE_FUSED]
void launch_step_tiled(const ForceArgs& args, int mask, bool fused, cudaStream_t st) {
    if (args.n <= 0) return;
    if (args.fast && args.use_rows) {  // force-only and fused launches of the SAME kernel: what lg_get_field(LG_A) returns in
                                        // fast mode is what the benched kernel computes
        if (args.use_rows == 2) launch_rows2(args, mask, fused, st);  // persistent pipelined organisation (kernels_rows.cu)
        else if (fused) launch_rows<MODE_FUSED>(args, mask, st);
        else launch_rows<MODE_FORCE>(args, mask, st);
    } else if (fused)
        launch_mode<MODE_FUSED>(args, mask, st);
    else
        launch_mode<MODE_FORCE>(args, mask, st);
}

}  // namespace lg
