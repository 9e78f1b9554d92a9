// kernels_force.cu -- the pair-force evaluation (sm_100a, FP64 CUDA cores; nothing here is a dense
// contraction, so no tensor cores).
//
// Replaces the hot loop of LT/rotational_system/rotational_binary_system_neighbors_omp.h:78-98 (and
// the all-pairs twin rotational_binary_system_omp.h:69-90) together with the functor fold of
// LG/granular_system/binary_force_container.h:20-32 and the unary call :84-89:
//   a[i] = 0; for j in row(i) ascending: a[i] += f1(i,j) + f2(i,j); then a[i] += unary(i)
// Every ordered pair is evaluated by the owner of i and writes only i's accumulators and i's slice of
// the history tables, so there are no atomics and the result is deterministic; the accumulation
// order is the reference's (ascending j), so a, alpha are bit-identical to it.
//
// Variant 0 ("simple"): one thread per particle walks its row.  Kept as the validation kernel.
// Variant 1 ("tiled") : pair-parallel evaluation into shared memory + ordered per-row reduction
//                       (kernels_force_tiled.cu).
#include "system.h"

namespace lg {

// ---- history-table helpers -------------------------------------------------------------------------
__device__ __forceinline__ Springs load_springs(const double* p) {
    Springs s;
    s.xi[0] = load3(p);
    s.xi[1] = load3(p + 3);
    s.xi[2] = load3(p + 6);
    return s;
}
__device__ __forceinline__ void store_springs(double* p, const Springs& s) {
    store3(p, s.xi[0]);
    store3(p + 3, s.xi[1]);
    store3(p + 6, s.xi[2]);
}
__device__ __forceinline__ Springs zero_springs() {
    Springs s;
    s.xi[0] = s.xi[1] = s.xi[2] = zero3();
    return s;
}

// Bonded force on i from bond record b of i (sinter_bridge.h:235-259).
__device__ __forceinline__ Accel bonded_accel(const ForceArgs& A, uint32_t b, d3 xi, d3 xj, d3 vi, d3 vj, d3 wi, d3 wj) {
    const double* ci = A.bconn + (size_t)b * 15;
    const double* cj = A.bconn + (size_t)A.brev[b] * 15;
    const double* len = A.blen + (size_t)b * 5;
    d3 force = zero3(), torque = zero3();
#pragma unroll
    for (int k = 0; k < 5; k++)
        bonded_spring(A.mp.k_d, A.mp.gamma_d, len[k], load3(ci + 3 * k), load3(cj + 3 * k), xi, xj, vi, vj, wi, wj, force,
                      torque);
    return Accel{force / A.mp.mass, torque / A.mp.inertia};
}

// Unary forces on i: right fold over the facets in insertion order,
// f_1 + (f_2 + (... + f_F)), each f = wall contact + wall Hamaker (triangular_facet.h:46-49).
__device__ __forceinline__ Accel unary_accel(const ForceArgs& A, int i, d3 xi, d3 vi, d3 wi) {
    Accel total = zero_accel();
    for (int f = A.n_facets - 1; f >= 0; f--) {
        const Facet& F = A.facets[f];
        d3 xf = facet_closest_point(F, xi);
        PairGeom g = pair_geometry(xi, xf);
        Accel c = zero_accel();
        if (F.has_contact) {
            double overlap = contact_overlap(g, A.mp.r);
            double* sp = A.wspr[f] + (size_t)i * 9;
            uint8_t act = A.wact[f][i];
            if (overlap <= 0) {
                if (act) {  // reset_springs (surface_contact_force.h:98-100)
                    store_springs(sp, zero_springs());
                    A.wact[f][i] = 0;
                }
            } else {
                Springs s = act ? load_springs(sp) : zero_springs();
                c = wall_contact_force(F.sp, A.mp.r, A.mp.mass, A.mp.inertia, g, overlap, vi, F.v, wi, s);
                store_springs(sp, s);
                if (!act) A.wact[f][i] = 1;
            }
        }
        if (F.has_hamaker) {
            Accel h = Accel{wall_hamaker_accel(F.A, F.h0, A.mp.r, A.mp.mass, g), zero3()};
            c = F.has_contact ? c + h : h;
        }
        total = (f == A.n_facets - 1) ? c : c + total;
    }
    return total;
}

template <int MASK>
__global__ void __launch_bounds__(128) force_simple_kernel(const __grid_constant__ ForceArgs A) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A.n) return;
    constexpr bool kContact = (MASK & (MODEL_CONTACT | MODEL_SINTER)) != 0;
    constexpr bool kSinter = (MASK & MODEL_SINTER) != 0;
    constexpr bool kHamaker = (MASK & MODEL_HAMAKER) != 0;

    const d3 xi = load3(A.x + 3 * (size_t)i);
    const d3 vi = load3(A.v + 3 * (size_t)i);
    const d3 wi = load3(A.w + 3 * (size_t)i);
    d3 acc_a = zero3(), acc_al = zero3();

    const uint32_t p0 = A.row_ptr[i], p1 = A.row_ptr[i + 1];
    for (uint32_t p = p0; p < p1; p++) {
        const int j = A.cols[p];
        const d3 xj = load3(A.x + 3 * (size_t)j);
        const PairGeom g = pair_geometry(xi, xj);
        Accel c = zero_accel();
        if constexpr (kContact) {
            const int t = A.tag[p];
            if (kSinter && t >= kTagBond && t != kTagNone) {
                const d3 vj = load3(A.v + 3 * (size_t)j), wj = load3(A.w + 3 * (size_t)j);
                const uint32_t b = A.bptr[i] + (uint32_t)(t - kTagBond);
                if (A.mp.alt_bonds) {  // alt_sinter_functor: the record's own three springs, no connectors
                    double* sp = A.bspr + (size_t)b * 9;
                    Springs s = load_springs(sp);
                    c = alt_bonded_force(A.mp.alt, A.mp.alt_dt, A.mp.r, A.mp.mass, A.mp.inertia, g, vi, vj, wi, wj, s);
                    store_springs(sp, s);
                } else {
                    c = bonded_accel(A, b, xi, xj, vi, vj, wi, wj);
                }
            } else {
                const double overlap = contact_overlap(g, 2.0 * A.mp.r);
                if (overlap <= 0) {
                    if (t != kTagNone) {  // reset_springs (contact_force.h:68-71): drop the record
                        A.sgid[(size_t)i * A.S + t] = -1;
                        A.tag[p] = (uint8_t)kTagNone;
                    }
                } else {
                    int slot = t;
                    Springs s;
                    if (t == kTagNone) {  // first overlapping evaluation of this ordered pair
                        slot = -1;
                        for (int q = 0; q < A.S; q++)
                            if (A.sgid[(size_t)i * A.S + q] < 0) {
                                slot = q;
                                break;
                            }
                        s = zero_springs();
                        if (slot >= 0) {
                            A.sgid[(size_t)i * A.S + slot] = j;
                            A.tag[p] = (uint8_t)slot;
                        } else {
                            atomicOr(A.err, ERR_SLOT_OVERFLOW);
                        }
                    } else {
                        s = load_springs(A.spr + ((size_t)i * A.S + slot) * 9);
                    }
                    const d3 vj = load3(A.v + 3 * (size_t)j), wj = load3(A.w + 3 * (size_t)j);
                    c = contact_force(A.mp.cp, A.mp.r, A.mp.mass, A.mp.inertia, g, overlap, vi, vj, wi, wj, s);
                    if (slot >= 0) store_springs(A.spr + ((size_t)i * A.S + slot) * 9, s);
                }
            }
        }
        if constexpr (kHamaker) {
            const d3 h = hamaker_accel(A.mp.hamaker_A, A.mp.hamaker_h0, A.mp.r, A.mp.mass, g, A.mp.hamaker_kind, A.mp.hamaker_cutoff);
            if constexpr (kContact)
                c.a = c.a + h;  // (contact|sinter)(i,j) + hamaker(i,j): two-term fold, torque + 0
            else
                c.a = h;
        }
        acc_a += c.a;
        acc_al += c.alpha;
    }
    if (A.n_facets > 0) {
        const Accel u = unary_accel(A, i, xi, vi, wi);
        acc_a += u.a;
        acc_al += u.alpha;
    }
    if (A.central_on) {  // central_attraction_functor (Custom-unary-force-model.md:99-100): a = k_a / mass * (x_i - x0), alpha = 0
        acc_a += (A.central_k / A.mp.mass) * (xi - A.central_x0);
        acc_al += zero3();
    }
    store3(A.a + 3 * (size_t)i, acc_a);
    store3(A.al + 3 * (size_t)i, acc_al);
}

template <int MASK>
static void launch_simple(const ForceArgs& args, cudaStream_t st) {
    const int threads = 128;
    const int blocks = (args.n + threads - 1) / threads;
    force_simple_kernel<MASK><<<blocks, threads, 0, st>>>(args);
    g_launch_count++;
}

void launch_force(const ForceArgs& args, int mask, int variant, cudaStream_t st) {
    if (args.n <= 0) return;
    if (variant >= 1) {
        launch_step_tiled(args, mask, false, st);
        return;
    }
    switch (mask & (MODEL_CONTACT | MODEL_HAMAKER | MODEL_SINTER)) {
        case MODEL_CONTACT: launch_simple<MODEL_CONTACT>(args, st); break;
        case MODEL_CONTACT | MODEL_HAMAKER: launch_simple<MODEL_CONTACT | MODEL_HAMAKER>(args, st); break;
        case MODEL_SINTER: launch_simple<MODEL_SINTER>(args, st); break;
        case MODEL_SINTER | MODEL_HAMAKER: launch_simple<MODEL_SINTER | MODEL_HAMAKER>(args, st); break;
        case MODEL_HAMAKER: launch_simple<MODEL_HAMAKER>(args, st); break;
        default: launch_simple<0>(args, st); break;  // unary forces only
    }
}

}  // namespace lg
