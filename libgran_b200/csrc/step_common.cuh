// step_common.cuh -- device helpers shared by the fused step kernels (kernels_force_tiled.cu: step_tiled_kernel,
// step_rows_kernel; kernels_rows.cu: step_rows2_kernel): mbarrier / bulk-copy wrappers, the out-of-line Luding contact,
// the unary (facet) fold, the bridge evaluation and the marked-pair loop of the row kernels.
// Everything sits in an anonymous namespace: each translation unit gets its own copy.
#pragma once
#include "system.h"

namespace lg {

namespace {

enum : int { MODE_FORCE = 0, MODE_FUSED = 1 };

#ifndef LG_ROWS_BU
#define LG_ROWS_BU 5
#endif
constexpr int RBU = LG_ROWS_BU;  // unroll factor of the five bridge springs in the row kernels

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    uint32_t done;
    do {
        asm volatile(
            "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
            : "=r"(done)
            : "r"(a), "r"(parity)
            : "memory");
    } while (!done);
}
// 16-byte aligned window around [begin, begin + nbytes): returns the bytes to copy (0 for an empty range)
__device__ __forceinline__ uint32_t window_bytes(const void* begin, uint32_t nbytes) {
    if (nbytes == 0) return 0u;
    const uint32_t lead = (uint32_t)((uintptr_t)begin & 15u);
    return (lead + nbytes + 15u) & ~15u;
}
__device__ __forceinline__ uint32_t window_lead(const void* begin) { return (uint32_t)((uintptr_t)begin & 15u); }
// one bulk copy global -> shared of the window around [begin, begin + nbytes), completion counted on `bar`
__device__ __forceinline__ void bulk_load(void* dst, const void* begin, uint32_t nbytes, unsigned long long* bar) {
    const uint32_t bytes = window_bytes(begin, nbytes);
    if (bytes == 0) return;
    const void* src = (const void*)((uintptr_t)begin & ~(uintptr_t)15u);
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
                 "l"(src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ Springs ld_springs(const double* p) {
    Springs s;
    s.xi[0] = load3(p);
    s.xi[1] = load3(p + 3);
    s.xi[2] = load3(p + 6);
    return s;
}
__device__ __forceinline__ void st_springs(double* p, const Springs& s) {
    store3(p, s.xi[0]);
    store3(p + 3, s.xi[1]);
    store3(p + 6, s.xi[2]);
}

// largest r in [0, nrows) with ptr[r] <= p (ptr[nrows] > p guaranteed by the caller)
__device__ __forceinline__ int find_row(const uint32_t* ptr, int nrows, uint32_t p) {
    int lo = 0, hi = nrows;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (ptr[mid] <= p) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// unary forces: right fold over facets, see kernels_force.cu:unary_accel
__device__ __forceinline__ Accel unary_forces(const ForceArgs& A, int i, d3 xi, d3 vi, d3 wi) {
    Accel total = zero_accel();
    for (int f = A.n_facets - 1; f >= 0; f--) {
        const Facet& F = A.facets[f];
        const d3 xf = facet_closest_point(F, xi);
        const PairGeom g = pair_geometry(xi, xf);
        Accel c = zero_accel();
        if (F.has_contact) {
            const double overlap = contact_overlap(g, A.mp.r);
            double* sp = A.wspr[f] + (size_t)i * 9;
            const uint8_t act = A.wact[f][i];
            if (overlap <= 0) {
                if (act) {
                    Springs z;
                    z.xi[0] = z.xi[1] = z.xi[2] = zero3();
                    st_springs(sp, z);
                    A.wact[f][i] = 0;
                }
            } else {
                Springs s;
                if (act) s = ld_springs(sp); else s.xi[0] = s.xi[1] = s.xi[2] = zero3();
                c = wall_contact_force(F.sp, A.mp.r, A.mp.mass, A.mp.inertia, g, overlap, vi, F.v, wi, s);
                st_springs(sp, s);
                if (!act) A.wact[f][i] = 1;
            }
        }
        if (F.has_hamaker) {
            const Accel h = Accel{wall_hamaker_accel(F.A, F.h0, A.mp.r, A.mp.mass, g), zero3()};
            c = F.has_contact ? c + h : h;
        }
        total = (f == A.n_facets - 1) ? c : c + total;
    }
    return total;
}

// The Luding contact law with its history springs for ONE overlapping ordered pair (phase B).  Kept out of line
// on purpose: it needs ~110 registers and runs for a small minority of pairs, so the kernel's register budget
// (and with it the number of resident warps) is set by the Hamaker / bridge paths instead; everything it needs
// is re-read from global memory (L1/L2 hits).  Always the reference's exact arithmetic, in both math modes: a relaxed
// version (measured: cfg 5d 5.5 -> 4.1 ms) cannot stay within any tolerance of the reference, because the torsion spring
// accumulates ALONG the contact normal (v_o is parallel to n, contact_force.h:84) and is then projected into the tangent plane
// and rescaled to its old length (:115-122) -- the direction that survives is rounding residue, reproducible only bit for bit.
__device__ __noinline__ void contact_heavy(const ForceArgs& A, int i, int j, uint32_t p, double* out6) {
    const d3 xi = load3(A.x + 3 * (size_t)i), vi = load3(A.v + 3 * (size_t)i), wi = load3(A.w + 3 * (size_t)i);
    const d3 xj = load3(A.x + 3 * (size_t)j), vj = load3(A.v + 3 * (size_t)j), wj = load3(A.w + 3 * (size_t)j);
    const int t = A.tag[p];
    const PairGeom g = pair_geometry(xi, xj);
    const double overlap = contact_overlap(g, 2.0 * A.mp.r);
    int slot = t;
    Springs s;
    if (t == kTagNone) {  // first overlapping evaluation: claim a free slot of particle i
        slot = -1;
        for (int k = 0; k < A.S; k++)
            if (atomicCAS(&A.sgid[(size_t)i * A.S + k], -1, j) == -1) {  // bookkeeping only
                slot = k;
                break;
            }
        s.xi[0] = s.xi[1] = s.xi[2] = zero3();
        if (slot >= 0) A.tag[p] = (uint8_t)slot; else atomicOr(A.err, ERR_SLOT_OVERFLOW);
    } else {
        s = ld_springs(A.spr + ((size_t)i * A.S + slot) * 9);
    }
    const Accel c = contact_force(A.mp.cp, A.mp.r, A.mp.mass, A.mp.inertia, g, overlap, vi, vj, wi, wj, s);
    if (slot >= 0) st_springs(A.spr + ((size_t)i * A.S + slot) * 9, s);
    store3(out6, c.a);
    store3(out6 + 3, c.alpha);
}

template <bool DAMPED>
__device__ __forceinline__ void bridge_fast(const ForceArgs& A, const double* ci, const double* cj, const double* len, d3 xi, d3 xj,
                                            d3 vi, d3 vj, d3 wi, d3 wj, d3& fa, d3& ta) {
    d3 force = zero3(), torque = zero3();
#pragma unroll RBU
    for (int k = 0; k < 5; k++)
        bonded_spring_fast<DAMPED>(A.mp.k_d, A.mp.gamma_d, len[k], load3(ci + 3 * k), load3(cj + 3 * k), xi, xj, vi, vj, wi, wj,
                                   force, torque);
    fa = force * A.fc.inv_mass;
    ta = torque * A.fc.inv_inertia;
}

// bonded value of record b (particle i = i0 + r) evaluated from global memory: used when the CTA holds more bond
// records than fit in shared memory
__device__ __noinline__ void bridge_from_global(const ForceArgs& A, int i, int j, uint32_t b, double* out6) {
    const d3 xi = load3(A.x + 3 * (size_t)i), xj = load3(A.x + 3 * (size_t)j);
    const double* ci = A.bconn + (size_t)b * 15;
    const double* cj = A.bconn + (size_t)A.brev[b] * 15;
    const double* len = A.blen + (size_t)b * 5;
    d3 fa, ta;
    if (A.mp.gamma_d != 0.0)
        bridge_fast<true>(A, ci, cj, len, xi, xj, load3(A.v + 3 * (size_t)i), load3(A.v + 3 * (size_t)j),
                          load3(A.w + 3 * (size_t)i), load3(A.w + 3 * (size_t)j), fa, ta);
    else
        bridge_fast<false>(A, ci, cj, len, xi, xj, zero3(), zero3(), zero3(), zero3(), fa, ta);
    store3(out6, fa);
    store3(out6 + 3, ta);
}

// The pairs of particle i0 + r that the hot loop of step_rows_kernel only marked (bit k of `slow`: pair pb + k), and,
// for rows longer than 64, every later pair that it skipped: bridges whose record is not staged, overlapping pairs
// (Luding contact) and pairs inside the rounding band of zero overlap (decided on the reference's exact expression).
// Returns the sum of their pair values (law + Hamaker) in ascending j.
template <bool kSinter, bool kHamaker>
__device__ __noinline__ void rows_slow_pairs(const ForceArgs& A, const double* sx, int i0, int nrows, int r,  // sx: staged x of particles i0 .. i0 + nrows; r: this particle's index in it
                                             uint32_t pb, uint32_t pe, uint32_t p_first, uint32_t stride,      // this caller's share of the row: p_first, p_first + stride, ... (< pe)
                                             unsigned long long slow, bool bonds_staged, uint32_t bond_base, double* out6) {
    const d3 xi = load3(sx + 3 * r);
    d3 acc_a = zero3(), acc_al = zero3();
    for (uint32_t p = p_first; p < pe; p += stride) {
        const uint32_t k = p - pb;
        if (k < 64u && !((slow >> k) & 1ull)) continue;
        const int j = A.cols[p];
        const int t = A.tag[p];
        const unsigned rj = (unsigned)(j - i0);
        const d3 xj = rj < (unsigned)nrows ? load3(sx + 3 * rj) : load3(A.x + 3 * (size_t)j);
        const bool bond = kSinter && t >= kTagBond && t != kTagNone;
        if (k >= 64u && bond && bonds_staged) continue;  // the hot loop has added it
        const PairGeom gf = pair_geometry_fast(xi, xj);
        d3 ha = zero3();
        if constexpr (kHamaker) ha = hamaker_accel_fast(A.fc, gf);
        double v6[6];
        if (bond) {
            bridge_from_global(A, i0 + r, j, bond_base + (uint32_t)(t - kTagBond), v6);
        } else {
            double overlap = A.fc.two_r - gf.len;
            if (k >= 64u && overlap <= -A.fc.band) continue;  // the hot loop has handled it
            if (fabs(overlap) < A.fc.band) overlap = contact_overlap(pair_geometry(xi, xj), A.fc.two_r);
            if (overlap <= 0) {  // inside the band but not overlapping: Hamaker only, springs reset
                if (t != kTagNone) {
                    A.sgid[(size_t)(i0 + r) * A.S + t] = -1;
                    A.tag[p] = (uint8_t)kTagNone;
                }
                acc_a += ha;
                continue;
            }
            contact_heavy(A, i0 + r, j, p, v6);
        }
        acc_a += kHamaker ? load3(v6) + ha : load3(v6);
        acc_al += load3(v6 + 3);
    }
    store3(out6, acc_a);
    store3(out6 + 3, acc_al);
}

}  // namespace

}  // namespace lg
