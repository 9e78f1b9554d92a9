// kernels_rows.cu -- step_rows2_kernel: the PERSISTENT, software-pipelined organisation of the short-row fused step
// (relaxed arithmetic, LG_MATH_FAST; aggregates with ~10 neighbors and ~2 bond records per particle).
//
// Same mathematics, same inputs and outputs as step_rows_kernel (kernels_force_tiled.cu); what changes is how the
// bytes move:
//   * a grid of (SMs x resident CTAs) blocks loops over tiles of RT2 consecutive particles;
//   * every contiguous range of a tile -- x (with a halo of RH2 particles either side), v, omega, theta, row / bond
//     pointer slices, columns, tags, and now also the tile's OWN bond records: connectors (120 B each), rest lengths,
//     partner and mirror indices -- arrives by cp.async.bulk (UBLKCP) against one mbarrier; the copies of tile t + 1 are
//     issued by one thread the moment tile t's results have left the stage, and at that time they are L2 hits, because
//     the same thread asked for them with cp.async.bulk.prefetch.L2 when tile t started.  Two CTAs share an SM and
//     cover each other's copy phases.  (A two-stage ring with a dedicated producer warp per CTA hid the copies
//     completely but left 8 compute warps per SM: 477 us per launch against 310 us; profiles/README.md r02b.)
//   * pass 1 (one thread per bond record) reads its connectors from shared memory, and the mirror record too when the
//     partner belongs to the tile (the common case inside an aggregate);
//   * pass 2 walks each particle's row with LPP cooperating lanes (partial sums combined by shuffles, fixed order);
//   * pass 3 integrates and moves the connectors IN the staged copy, and the results leave by bulk stores
//     (cp.async.bulk.global.shared::cta) straight from the stage.
// Replaces (like the other step kernels) LT/rotational_system/rotational_binary_system_neighbors_omp.h:78-98,
// LG/granular_system/binary_force_container.h:20-32, LG/contact_force/contact_force.h:57-147,
// LG/hamaker_force/hamaker_force.h:28-43, LG/sinter_bridge/sinter_bridge.h:24-94,214-259 and
// LT/rotational_integrator/rotational_velocity_verlet_half.h:63-73.
#include <cstdlib>

#include "step_common.cuh"

namespace lg {

namespace {

#ifndef LG_R2_RT
#define LG_R2_RT 96    // default shape: particles per tile (96 x 2 lanes, 3 CTAs/SM measured best on cfg 4: profiles/README.md r02f, r02j)
#endif
#ifndef LG_R2_LPP
#define LG_R2_LPP 2    // default shape: lanes per particle in pass 2
#endif
// resident CTAs per SM the launch bound asks for: 5 for CTAs of 96 threads, 3 up to 192 threads, else 2
#ifndef LG_R2_MINB
#define LG_R2_MINB(threads) ((threads) <= 96 ? 5 : (threads) <= 192 ? 3 : 2)
#endif

// Tile shape: RT particles per tile, LPP cooperating lanes per particle in the pair loop (RT * LPP threads per CTA).
//   < 96, 2>  rows of ~10 neighbors (aggregates)          staged x: the tile +- 128 particles, 12 pairs per particle
//   < 64, 4>  rows of ~20-45 (fractal clusters, loose packings)     +- 192                  48
//   < 32, 8>  rows of ~50-100 (dense packings)                      +- 256                  96
//   < 48, 2>  shape < 96, 2> on half-size tiles (same lanes per particle = same summation order): small per-rank systems
//             of slab runs, where 96-particle tiles leave the last wave of CTAs a quarter full
template <int RT_, int LPP_>
struct Shape {
    static constexpr int RT = RT_, LPP = LPP_;
    static constexpr int NT = RT * LPP;                                   // compute threads
    static constexpr int RH = LPP == 2 ? 128 : LPP == 4 ? 192 : 256;      // halo of the staged x window (particles either side)
    static constexpr int RPC = (LPP == 2 ? 12 : LPP == 4 ? 48 : 96) * RT; // staged columns / tags per tile
    static constexpr int RBC = 2 * RT + 32;                               // staged bond records per tile
    static_assert(RT % 16 == 0 && (LPP == 1 || LPP == 2 || LPP == 4 || LPP == 8) && RPC % 16 == 0 && RBC % 16 == 0, "tile shape");

    struct alignas(16) Stage {
        double wx[3 * (RT + 2 * RH) + 4];
        double wv[3 * RT + 4];
        double ww[3 * RT + 4];
        double wth[3 * RT + 4];
        double wconn[15 * RBC + 4];
        double wlen[5 * RBC + 4];
        uint32_t wrow[RT + 12];
        uint32_t wbond[RT + 12];
        int32_t wcols[RPC + 8];
        int32_t wbj[RBC + 8];
        int32_t wrev[RBC + 8];
        uint16_t wwork[RBC + 16];
        uint8_t wtag[RPC + 32];
    };
    struct alignas(16) Smem {
        Stage st;
        double res[6 * RBC];     // pass 1 results per bond record; pass 3: dx, dtheta of every particle
        uint8_t brow[RBC];       // local row of every staged bond record
        unsigned long long full;
        uint32_t nwork;           // bridge work items of the staged tile
        uint32_t h0, h1;          // halo items of the staged tile (fused launches of slab runs)
    };
};

__device__ __forceinline__ void compute_barrier() { __syncthreads(); }

// Bulk store shared -> global of [ssrc, ssrc + nbytes) to gdst; both addresses are congruent mod 16 (0 or 8) and nbytes is
// a multiple of 8: the 16-byte aligned middle goes by cp.async.bulk, a leading / trailing double by an ordinary store.
__device__ __forceinline__ void bulk_store(double* gdst, const double* ssrc, uint32_t nbytes) {
    if (nbytes == 0) return;
    if ((uintptr_t)gdst & 15u) {
        *gdst = *ssrc;
        gdst++;
        ssrc++;
        nbytes -= 8u;
    }
    const uint32_t mid = nbytes & ~15u;
    if (mid)
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(ssrc)), "r"(mid) : "memory");
    if (nbytes & 8u) gdst[mid >> 3] = ssrc[mid >> 3];
}

// global -> L2 prefetch of the 16-byte aligned window around [begin, begin + nbytes)
__device__ __forceinline__ void l2_prefetch(const void* begin, uint32_t nbytes) {
    const uint32_t bytes = window_bytes(begin, nbytes);
    if (bytes == 0) return;
    const void* src = (const void*)((uintptr_t)begin & ~(uintptr_t)15u);
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}

template <int MASK, int MODE, bool FACETS, int RT2, int LPP>
__global__ void __launch_bounds__(RT2 * LPP, LG_R2_MINB(RT2 * LPP)) step_rows2_kernel(const __grid_constant__ ForceArgs A) {
    using SH = Shape<RT2, LPP>;
    using Stage2 = typename SH::Stage;
    using Rows2Smem = typename SH::Smem;
    constexpr int NT2 = SH::NT, RH2 = SH::RH, RPC2 = SH::RPC, RBC2 = SH::RBC;
    constexpr bool kContact = (MASK & (MODEL_CONTACT | MODEL_SINTER)) != 0;
    constexpr bool kSinter = (MASK & MODEL_SINTER) != 0;
    constexpr bool kHamaker = (MASK & MODEL_HAMAKER) != 0;

    extern __shared__ __align__(16) unsigned char smem_raw[];
    Rows2Smem& sm = *reinterpret_cast<Rows2Smem*>(smem_raw);

    const int tid = threadIdx.x;
    const int n_tiles = (A.n + RT2 - 1) / RT2;
    const bool have_bonds = kSinter && A.bptr != nullptr;

    // pair / bond-record ranges of tile t: four independent loads (requested a tile ahead, consumed when its copies are issued)
    auto load_desc = [&](int t, uint32_t d[7]) {
        const int i0 = t * RT2, nr = min(RT2, A.n - i0);
        d[0] = A.row_ptr[i0];
        d[1] = A.row_ptr[i0 + nr];
        d[2] = have_bonds ? A.bptr[i0] : 0u;
        d[3] = have_bonds ? A.bptr[i0 + nr] : 0u;
        d[4] = have_bonds ? A.bwork_cnt[t] : 0u;  // with the others: no dependent load when the copies are issued
        d[5] = d[6] = 0u;
        if constexpr (MODE == MODE_FUSED) {
            if (A.hf.tile_ptr != nullptr) {
                d[5] = A.hf.tile_ptr[t];
                d[6] = A.hf.tile_ptr[t + 1];
            }
        }
    };
    // bulk copies of every contiguous range tile t reads, completion counted on sm.full (one thread)
    auto issue_loads = [&](int tile, const uint32_t d[7]) {
        Stage2& S = sm.st;
        const int i0 = tile * RT2, nrows = min(RT2, A.n - i0);
        const int xlo = max(i0 - RH2, 0), xcnt = min(A.n, i0 + RT2 + RH2) - xlo;
        const double* gx = A.x + 3 * (size_t)xlo;
        const double* gv = A.v + 3 * (size_t)i0;
        const double* gw = A.w + 3 * (size_t)i0;
        const double* gth = A.th + 3 * (size_t)i0;
        const uint32_t* grow = A.row_ptr + i0;
        const uint32_t* gbond = have_bonds ? A.bptr + i0 : nullptr;
        const uint32_t nb3 = 24u * (uint32_t)nrows, nbp = 4u * (uint32_t)(nrows + 1), nbx = 24u * (uint32_t)xcnt;
        const uint32_t np0 = min(d[1] - d[0], (uint32_t)RPC2);
        const uint32_t nbr = (d[3] - d[2] <= (uint32_t)RBC2) ? d[3] - d[2] : 0u;
        const int32_t* gcols = A.cols + d[0];
        const uint8_t* gtag = A.tag + d[0];
        const double* gconn = A.bconn + 15 * (size_t)d[2];
        const double* glen = A.blen + 5 * (size_t)d[2];
        const int32_t* gbj = A.bj + d[2];
        const int32_t* grev = A.brev + d[2];
        const uint16_t* gwork = A.bwork + d[2];
        uint32_t total = window_bytes(gx, nbx) + window_bytes(gv, nb3) + window_bytes(gw, nb3) + window_bytes(grow, nbp) +
                         window_bytes(gcols, 4u * np0);
        if constexpr (MODE == MODE_FUSED) total += window_bytes(gth, nb3);
        if (have_bonds) total += window_bytes(gbond, nbp);
        if constexpr (kContact) total += window_bytes(gtag, np0);
        if (nbr)
            total += window_bytes(gconn, 120u * nbr) + window_bytes(glen, 40u * nbr) + window_bytes(gbj, 4u * nbr) +
                     window_bytes(grev, 4u * nbr) + window_bytes(gwork, 2u * nbr);
        sm.h0 = d[5];
        sm.h1 = d[6];
        sm.nwork = nbr ? d[4] : 0u;  // published to the other threads by the barrier's release / acquire
        mbar_expect_tx(&sm.full, total);
        if (nbr) {
            bulk_load(S.wconn, gconn, 120u * nbr, &sm.full);
            bulk_load(S.wlen, glen, 40u * nbr, &sm.full);
            bulk_load(S.wbj, gbj, 4u * nbr, &sm.full);
            bulk_load(S.wrev, grev, 4u * nbr, &sm.full);
            bulk_load(S.wwork, gwork, 2u * nbr, &sm.full);
        }
        bulk_load(S.wx, gx, nbx, &sm.full);
        bulk_load(S.wv, gv, nb3, &sm.full);
        bulk_load(S.ww, gw, nb3, &sm.full);
        if constexpr (MODE == MODE_FUSED) bulk_load(S.wth, gth, nb3, &sm.full);
        bulk_load(S.wrow, grow, nbp, &sm.full);
        if (have_bonds) bulk_load(S.wbond, gbond, nbp, &sm.full);
        bulk_load(S.wcols, gcols, 4u * np0, &sm.full);
        if constexpr (kContact) bulk_load(S.wtag, gtag, np0, &sm.full);
    };
    // the big ranges of tile t pulled into L2 (no shared memory needed): by the time the stage is free its copies are L2 hits
    auto prefetch_tile = [&](int t, const uint32_t d[7]) {
        const int j0 = t * RT2, nr = min(RT2, A.n - j0);
        l2_prefetch(A.v + 3 * (size_t)j0, 24u * (uint32_t)nr);
        l2_prefetch(A.w + 3 * (size_t)j0, 24u * (uint32_t)nr);
        l2_prefetch(A.x + 3 * (size_t)j0, 24u * (uint32_t)nr);
        if constexpr (MODE == MODE_FUSED) l2_prefetch(A.th + 3 * (size_t)j0, 24u * (uint32_t)nr);
        l2_prefetch(A.cols + d[0], 4u * (d[1] - d[0]));
        if (have_bonds && d[3] > d[2]) {
            l2_prefetch(A.bconn + 15 * (size_t)d[2], 120u * (d[3] - d[2]));
            l2_prefetch(A.blen + 5 * (size_t)d[2], 40u * (d[3] - d[2]));
        }
    };

    // Order in which the grid walks the tiles: ascending, or -- fused launches of slab runs -- the tiles that hold halo items
    // first, so that their peer stores, fences and the publication of the exchange run underneath the interior tiles instead
    // of at the kernel's end.
    auto tile_of = [&](int slot) -> int {
        if constexpr (MODE == MODE_FUSED) {
            if (A.hf.tile_order != nullptr) return (int)A.hf.tile_order[slot];
        }
        return slot;
    };
    uint32_t dn[7] = {0u, 0u, 0u, 0u, 0u, 0u, 0u};  // thread 0: ranges of the NEXT tile of this CTA
    if (tid == 0) {
        mbar_init(&sm.full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        uint32_t d0[7];
        const int t0 = tile_of((int)blockIdx.x);
        load_desc(t0, d0);
        issue_loads(t0, d0);
        if ((int)(blockIdx.x + gridDim.x) < n_tiles) load_desc(tile_of((int)(blockIdx.x + gridDim.x)), dn);
    }
    __syncthreads();

    const int r = tid / LPP, lp = tid % LPP;  // particle of the tile and lane within it (pass 2)
    int k = 0;
    for (int slot = blockIdx.x; slot < n_tiles; slot += gridDim.x, k++) {
        Stage2& S = sm.st;
        const int tile = tile_of(slot);
        const int i0 = tile * RT2, nrows = min(RT2, A.n - i0);
        const int xlo = max(i0 - RH2, 0), xcnt = min(A.n, i0 + RT2 + RH2) - xlo;
        const int nslot = slot + (int)gridDim.x;
        const int nxt = nslot < n_tiles ? tile_of(nslot) : n_tiles;  // n_tiles: none
        mbar_wait(&sm.full, (uint32_t)(k & 1));

        double* sxw = S.wx + (window_lead(A.x + 3 * (size_t)xlo) >> 3);  // x of particle xlo + q at sxw[3 q]
        double* sx = sxw + 3 * (i0 - xlo);
        double* sv = S.wv + (window_lead(A.v + 3 * (size_t)i0) >> 3);
        double* sw = S.ww + (window_lead(A.w + 3 * (size_t)i0) >> 3);
        double* sth = S.wth + (window_lead(A.th + 3 * (size_t)i0) >> 3);
        const uint32_t* sptr = S.wrow + (window_lead(A.row_ptr + i0) >> 2);
        const uint32_t* sbptr = S.wbond + (have_bonds ? (window_lead(A.bptr + i0) >> 2) : 0u);
        const uint32_t p_begin = sptr[0];
        const uint32_t b0 = have_bonds ? sbptr[0] : 0u;
        const uint32_t nbond = have_bonds ? sbptr[nrows] - b0 : 0u;
        const bool bonds_staged = nbond > 0 && nbond <= (uint32_t)RBC2;
        const int32_t* scols = S.wcols + (window_lead(A.cols + p_begin) >> 2);
        const uint8_t* stag = S.wtag + window_lead(A.tag + p_begin);
        double* sconn = S.wconn + (window_lead(A.bconn + 15 * (size_t)b0) >> 3);
        const double* slen = S.wlen + (window_lead(A.blen + 5 * (size_t)b0) >> 3);
        const int32_t* sbj = S.wbj + (window_lead(A.bj + b0) >> 2);
        const int32_t* srev = S.wrev + (window_lead(A.brev + b0) >> 2);
        const uint16_t* swork = S.wwork + (window_lead(A.bwork + b0) >> 1);

        auto neighbour = [&](int j, const double* sp, const double* gsrc) -> d3 {  // v, omega: own particles staged
            const unsigned rj = (unsigned)(j - i0);
            return rj < (unsigned)nrows ? load3(sp + 3 * rj) : load3(gsrc + 3 * (size_t)j);
        };
        auto neighbour_x = [&](int j) -> d3 {
            const unsigned rj = (unsigned)(j - xlo);
            return rj < (unsigned)xcnt ? load3(sxw + 3 * rj) : load3(A.x + 3 * (size_t)j);
        };

        // ---- pass 1: one thread per bridge work item, operands from the stage.  An item is a bond record (i, j) of the
        //      tile; when its mirror (j, i) is in the tile as well (97 % of the records of config 4) the item carries bit 15
        //      and the thread produces BOTH records' values from one evaluation of the five springs -----------------------
        if constexpr (kSinter) {
            if (bonds_staged) {
                if (tid < nrows)
                    for (uint32_t b = sbptr[tid]; b < sbptr[tid + 1]; b++) sm.brow[b - b0] = (uint8_t)tid;
                compute_barrier();
                const uint32_t nwork = sm.nwork;
                const bool damped = A.mp.gamma_d != 0.0;
                for (uint32_t wq = tid; wq < nwork; wq += NT2) {
                    const uint32_t item = swork[wq];
                    const uint32_t bl = item & 0x7fffu;
                    const bool paired = (item & 0x8000u) != 0u;
                    const int rr = sm.brow[bl];
                    const int j = sbj[bl];
                    const uint32_t ml = (uint32_t)srev[bl] - b0;
                    const double* ci = sconn + 15 * bl;
                    const double* cj = paired ? sconn + 15 * ml : A.bconn + (size_t)srev[bl] * 15;
                    const double* len = slen + 5 * bl;
                    const d3 xi = load3(sx + 3 * rr), xj = neighbour_x(j);
                    d3 vi = zero3(), vj = zero3(), wi = zero3(), wj = zero3();
                    if (damped) {
                        vi = load3(sv + 3 * rr);
                        vj = neighbour(j, sv, A.v);
                        wi = load3(sw + 3 * rr);
                        wj = neighbour(j, sw, A.w);
                    }
                    d3 force = zero3(), ti = zero3(), tj = zero3();
                    if (damped) {
#pragma unroll RBU
                        for (int q = 0; q < 5; q++)
                            bonded_spring_pair_fast<true>(A.mp.k_d, A.mp.gamma_d, len[q], load3(ci + 3 * q), load3(cj + 3 * q), xi, xj,
                                                          vi, vj, wi, wj, force, ti, tj);
                    } else {
#pragma unroll RBU
                        for (int q = 0; q < 5; q++)
                            bonded_spring_pair_fast<false>(A.mp.k_d, 0.0, len[q], load3(ci + 3 * q), load3(cj + 3 * q), xi, xj, vi, vj,
                                                           wi, wj, force, ti, tj);
                    }
                    const d3 fa = force * A.fc.inv_mass;
                    store3(sm.res + 6 * bl, fa);
                    store3(sm.res + 6 * bl + 3, ti * A.fc.inv_inertia);
                    if (paired) {
                        store3(sm.res + 6 * ml, -fa);
                        store3(sm.res + 6 * ml + 3, tj * A.fc.inv_inertia);
                    }
                }
            }
        }
        compute_barrier();  // pass 1 results
        if (tid == 0 && nxt < n_tiles) prefetch_tile(nxt, dn);  // its ranges were requested a tile ago: no wait here

        // ---- pass 2: LPP lanes per particle walk its row (lane l: pairs pb + l, pb + l + LPP, ...) ---------------------
        d3 acc_a = zero3(), acc_al = zero3();
        unsigned long long slow = 0ull;  // bit q: pair pb + q (q < 64) is evaluated out of line afterwards
        const bool live = r < nrows;
        const d3 xi = live ? load3(sx + 3 * r) : zero3();
        uint32_t pb = 0u, pe = 0u;
        if (live) {
            pb = sptr[r];
            pe = sptr[r + 1];
            for (uint32_t p = pb + (uint32_t)lp; p < pe; p += LPP) {
                const uint32_t ps = p - p_begin;
                const int j = ps < (uint32_t)RPC2 ? scols[ps] : A.cols[p];
                const d3 xj = neighbour_x(j);
                const PairGeom g = pair_geometry_fast(xi, xj);
                d3 pa = zero3();
                if constexpr (kHamaker) pa = hamaker_accel_fast(A.fc, g);
                if constexpr (kContact) {
                    const int t = ps < (uint32_t)RPC2 ? stag[ps] : A.tag[p];
                    if (kSinter && t >= kTagBond && t != kTagNone) {
                        if (bonds_staged) {
                            const uint32_t bl = sbptr[r] + (uint32_t)(t - kTagBond) - b0;
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
                                A.sgid[(size_t)(i0 + r) * A.S + t] = -1;
                                A.tag[p] = (uint8_t)kTagNone;
                            }
                        } else {  // overlapping, or inside the rounding band of zero (decided on the exact expression)
                            if (p - pb < 64u) slow |= 1ull << (p - pb);
                            continue;
                        }
                    }
                }
                acc_a += pa;
            }
        }
        if constexpr (kContact) {
            // rare: out of line, so that the hot loop keeps its registers.  Every lane evaluates the pairs IT marked (and its
            // share of the pairs beyond the 64-bit mask), so the contacts of a dense row spread over the LPP lanes.
            if (live && (slow != 0ull || pe - pb > 64u)) {
                double out6[6];
                rows_slow_pairs<kSinter, kHamaker>(A, sxw, xlo, xcnt, r + (i0 - xlo), pb, pe, pb + (uint32_t)lp, (uint32_t)LPP, slow,
                                                   bonds_staged, have_bonds ? sbptr[r] : 0u, out6);
                acc_a += load3(out6);
                acc_al += load3(out6 + 3);
            }
        }
        if constexpr (LPP > 1) {  // fixed-shape combination of the lanes' partial sums (all 32 lanes take part)
#pragma unroll
            for (int off = 1; off < LPP; off <<= 1) {
                acc_a.x += __shfl_xor_sync(0xffffffffu, acc_a.x, off);
                acc_a.y += __shfl_xor_sync(0xffffffffu, acc_a.y, off);
                acc_a.z += __shfl_xor_sync(0xffffffffu, acc_a.z, off);
                acc_al.x += __shfl_xor_sync(0xffffffffu, acc_al.x, off);
                acc_al.y += __shfl_xor_sync(0xffffffffu, acc_al.y, off);
                acc_al.z += __shfl_xor_sync(0xffffffffu, acc_al.z, off);
            }
        }
        d3 xn = zero3(), vn = zero3(), wn = zero3(), dth = zero3(), dxv = zero3();
        if (live && lp == 0) {
            const d3 vi = load3(sv + 3 * r), wi = load3(sw + 3 * r);
            if constexpr (FACETS) {
                const Accel u = unary_forces(A, i0 + r, xi, vi, wi);
                acc_a += u.a;
                acc_al += u.alpha;
            }
            if constexpr (MODE == MODE_FORCE) {
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
        compute_barrier();  // every read of the staged state and of the pass-1 results is done

        // ---- pass 3: results into the stage, connectors moved in place, bulk stores ----------------------------------------
        if (live && lp == 0) {
            store3(sv + 3 * r, vn);
            store3(sw + 3 * r, wn);
            if constexpr (MODE == MODE_FUSED) {
                store3(sx + 3 * r, xn);
                store3(sth + 3 * r, load3(sth + 3 * r) + dth);
                store3(sm.res + 6 * r, dxv);
                store3(sm.res + 6 * r + 3, dth);
            }
        }
        if constexpr (MODE == MODE_FUSED && kSinter) {
            if (have_bonds) {
                compute_barrier();
                // connectors of the tile's particles: c += dx ; c += dtheta x (c - x_new)   (sinter_bridge.h:36-41,67-74)
                const int ncon = (int)nbond * 5;
                if (bonds_staged) {
                    for (int cq = tid; cq < ncon; cq += NT2) {
                        const int rr = sm.brow[cq / 5];
                        d3 ck = load3(sconn + 3 * cq) + load3(sm.res + 6 * rr);
                        const d3 arm = ck - load3(sx + 3 * rr);
                        ck = ck + cross(load3(sm.res + 6 * rr + 3), arm);
                        store3(sconn + 3 * cq, ck);
                    }
                } else {  // more bond records than the stage holds: straight from / to global memory
                    for (int cq = tid; cq < ncon; cq += NT2) {
                        const int rr = find_row(sbptr, nrows, b0 + (uint32_t)(cq / 5));
                        const size_t off = (size_t)b0 * 15 + 3 * (size_t)cq;
                        d3 ck = load3(A.bconn + off) + load3(sm.res + 6 * rr);
                        const d3 arm = ck - load3(sx + 3 * rr);
                        ck = ck + cross(load3(sm.res + 6 * rr + 3), arm);
                        store3(A.bconn_next + off, ck);
                    }
                }
            }
        }
        bool halo_tile = false;  // this tile has stored halo items into peer memory
        if constexpr (MODE == MODE_FUSED) {
            if (A.hf.tile_ptr != nullptr) {  // slab runs: this tile's halo items go straight into the neighbours' receive buffers
                const uint32_t h0 = sm.h0, h1 = sm.h1;  // loaded with the tile's descriptors a tile ahead
                if (h1 > h0) {  // uniform over the CTA
                    halo_tile = true;
                    compute_barrier();  // the staged results of every thread (and connectors written to global) are visible
                    // 16 consecutive threads per item, one double each: an item's 72 / 120 bytes leave as one or two
                    // contiguous NVLink writes instead of 9 / 15 scattered ones
                    const uint32_t n_slots = (h1 - h0) * 16u;
                    for (uint32_t e = (uint32_t)tid; e < n_slots; e += NT2) {
                        const uint2 it = A.hf.items[h0 + (e >> 4)];
                        const int c = (int)(e & 15u);
                        const int peer = (int)(it.x >> 24), idx = (int)(it.x & 0x7fffffu);
                        if (it.x & 0x800000u) {  // bond record: 5 connectors
                            if (c < 15) {
                                const double* src = bonds_staged ? sconn + 15 * idx : A.bconn_next + 15 * ((size_t)b0 + (size_t)idx);
                                A.hf.remote[peer][A.hf.rec_off[peer] + 15 * (size_t)it.y + c] = src[c];
                            }
                        } else if (c < 9) {  // particle: x, v, omega
                            const double* src = c < 3 ? sx : (c < 6 ? sv : sw);
                            A.hf.remote[peer][9 * (size_t)it.y + c] = src[3 * idx + (c % 3)];
                        }
                    }
                    __threadfence_system();  // this thread's peer stores are visible system-wide before the tile is counted
                }
            }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // this thread's shared-memory writes -> async proxy
        compute_barrier();
        if (tid == 0) {
            if (halo_tile) {  // count the tile; the launch's last tile with items publishes the exchange to every peer
                const unsigned int prev = atomicAdd(A.hf.done, 1u);
                if (prev == A.hf.expected - 1u) {
                    *A.hf.done = 0u;  // next exchange (stream-ordered after this kernel)
                    __threadfence_system();
                    for (int p = 0; p < A.hf.n_peers; p++)
                        asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(A.hf.remote_flag[p]), "l"(A.hf.seq) : "memory");
                }
            }
            const uint32_t nb3 = 24u * (uint32_t)nrows;
            const size_t g0 = 3 * (size_t)i0;
            if constexpr (MODE == MODE_FORCE) {
                bulk_store(A.a + g0, sv, nb3);
                bulk_store(A.al + g0, sw, nb3);
            } else {
                bulk_store(A.x_next + g0, sx, nb3);
                bulk_store(A.v_next + g0, sv, nb3);
                bulk_store(A.w_next + g0, sw, nb3);
                bulk_store(A.th + g0, sth, nb3);
                if (kSinter && bonds_staged) bulk_store(A.bconn_next + 15 * (size_t)b0, sconn, 120u * nbond);
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // the stage may be overwritten
            if (nxt < n_tiles) {
                issue_loads(nxt, dn);
                if (nslot + (int)gridDim.x < n_tiles) load_desc(tile_of(nslot + (int)gridDim.x), dn);
            }
        }
    }
    if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // all stores have landed before the CTA retires
}

template <int MASK, int MODE, bool FACETS, int RT2, int LPP>
void launch_rows2_one(const ForceArgs& args, cudaStream_t st) {
    using SH = Shape<RT2, LPP>;
    static int grid_cap = 0;  // per instantiation
    if (grid_cap == 0) {
        const void* fn = (const void*)step_rows2_kernel<MASK, MODE, FACETS, RT2, LPP>;
        cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(typename SH::Smem));
        int dev = 0, sms = 0, occ = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, fn, SH::NT, sizeof(typename SH::Smem));
        grid_cap = sms * (occ > 0 ? occ : 1);
    }
    const int n_tiles = (args.n + RT2 - 1) / RT2;
    step_rows2_kernel<MASK, MODE, FACETS, RT2, LPP><<<min(n_tiles, grid_cap), SH::NT, sizeof(typename SH::Smem), st>>>(args);
    g_launch_count++;
}

template <int MASK, int MODE, bool FACETS>
void launch_rows2_shape(const ForceArgs& args, cudaStream_t st) {
    switch (args.rows_shape) {
        case 1: launch_rows2_one<MASK, MODE, FACETS, 64, 4>(args, st); break;
        case 2: launch_rows2_one<MASK, MODE, FACETS, 32, 8>(args, st); break;
        case 3: launch_rows2_one<MASK, MODE, FACETS, 48, 2>(args, st); break;  // shape 0's arithmetic on half-size tiles
        default: launch_rows2_one<MASK, MODE, FACETS, LG_R2_RT, LG_R2_LPP>(args, st); break;
    }
}

template <int MASK, int MODE>
void launch_rows2_facets(const ForceArgs& args, cudaStream_t st) {
    if (args.n_facets > 0) launch_rows2_shape<MASK, MODE, true>(args, st); else launch_rows2_shape<MASK, MODE, false>(args, st);
}

template <int MODE>
void launch_rows2_mode(const ForceArgs& args, int mask, cudaStream_t st) {
    switch (mask & (MODEL_CONTACT | MODEL_HAMAKER | MODEL_SINTER)) {
        case MODEL_CONTACT: launch_rows2_facets<MODEL_CONTACT, MODE>(args, st); break;
        case MODEL_CONTACT | MODEL_HAMAKER: launch_rows2_facets<MODEL_CONTACT | MODEL_HAMAKER, MODE>(args, st); break;
        case MODEL_SINTER: launch_rows2_facets<MODEL_SINTER, MODE>(args, st); break;
        case MODEL_SINTER | MODEL_HAMAKER: launch_rows2_facets<MODEL_SINTER | MODEL_HAMAKER, MODE>(args, st); break;
        case MODEL_HAMAKER: launch_rows2_facets<MODEL_HAMAKER, MODE>(args, st); break;
        default: launch_rows2_facets<0, MODE>(args, st); break;
    }
}

}  // namespace

// tile shape the launch will use for a mean row length (api.cu builds the bridge work lists for its tile size)
int rows2_shape_for(double k_mean, int n, bool bonds) {
    static const int forced = [] { const char* e = std::getenv("LGB200_ROWS_SHAPE"); return e ? std::atoi(e) : -1; }();
    if (forced >= 0 && forced <= 2) return forced;
    // measured (profiles/README.md r02g): <96,2> for rows of ~10, <64,4> for loose packings (rows of ~35, no bridges),
    // <32,8> for dense rows, for bridged clusters with long rows (fewer tiles overflow the staged records) and for systems
    // too small to fill the SMs with larger tiles
    if (k_mean <= 16.0) return n >= 40000 ? 0 : 2;
    if (k_mean <= 48.0) return (bonds || n < 40000) ? 2 : 1;
    return 2;
}
int rows2_tile_particles(int shape) { return shape == 1 ? 64 : shape == 2 ? 32 : shape == 3 ? 48 : LG_R2_RT; }
int rows2_tile_records(int shape) { return 2 * rows2_tile_particles(shape) + 32; }

void launch_rows2(const ForceArgs& args, int mask, bool fused, cudaStream_t st) {
    if (fused) launch_rows2_mode<MODE_FUSED>(args, mask, st); else launch_rows2_mode<MODE_FORCE>(args, mask, st);
}

}  // namespace lg
