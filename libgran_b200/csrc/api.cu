// api.cu -- the extern "C" layer of include/lgb200/lgb200.h: owns the device buffers of one granular
// system and sequences the kernels.  Host-side counterpart of
//   LT/rotational_system/rotational_system.h:41-130               (state ownership, getters)
//   LT/rotational_integrator/rotational_velocity_verlet_half.h:43-76 (do_step schedule, first-step kick)
//   LT/rotational_system/rotational_binary_system_neighbors_omp.h:102-116 (update_neighbor_list)
//   LG/sinter_bridge/sinter_bridge.h:148-211                      (bond construction, host, init only)
// No CPU fallback exists: every compute entry point launches CUDA kernels or fails.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <new>
#include <string>
#include <vector>

#include "../../include/lgb200/lgb200.h"
#include "system.h"
#include "bond_build.cuh"

namespace lg {
thread_local int64_t g_launch_count = 0;
}

namespace {

thread_local std::string g_error;

int fail(int code, const std::string& msg) {
    g_error = msg;
    return code;
}

}  // namespace
namespace lg {
int fail_message(int code, const char* msg) { return fail(code, msg); }  // for the other translation units
}
namespace {

#define LG_CUDA(expr)                                                                                   \
    do {                                                                                                \
        cudaError_t e_ = (expr);                                                                        \
        if (e_ != cudaSuccess)                                                                          \
            return fail(e_ == cudaErrorMemoryAllocation ? LG_ENOMEM : LG_ECUDA,                         \
                        std::string(#expr) + ": " + cudaGetErrorString(e_));                            \
    } while (0)

#define LG_TRY(expr)              \
    do {                          \
        int rc_ = (expr);         \
        if (rc_ != LG_OK) return rc_; \
    } while (0)

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;  // elements
    ~DevBuf() { release(); }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    // grow-only; contents are NOT preserved
    cudaError_t reserve(size_t elems, double headroom = 1.0) {
        if (elems <= cap) return cudaSuccess;
        release();
        size_t want = (size_t)((double)elems * headroom) + 16;
        cudaError_t e = cudaMalloc(&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
};

// double-buffered array: kernels of step k read `p` and write `next`; the host swaps after the launch
struct PingPong {
    DevBuf<double> b[2];
    int cur = 0;
    double* p = nullptr;
    double* next = nullptr;
    cudaError_t reserve(size_t elems) {
        cudaError_t e = b[0].reserve(elems);
        if (e != cudaSuccess) return e;
        e = b[1].reserve(elems);
        p = b[cur].p;
        next = b[cur ^ 1].p;
        return e;
    }
    void swap() {
        cur ^= 1;
        p = b[cur].p;
        next = b[cur ^ 1].p;
    }
};

}  // namespace

using namespace lg;

struct lg_system {
    int device = 0;
    cudaStream_t stream = nullptr;
    int64_t n = 0;
    ModelParams mp{};
    int mask = 0;  // binary models
    int variant = 1;
    int math_mode = 0;  // 0 = bit-exact (reference operand order, IEEE division), 1 = relaxed (lg_set_math_mode)
    int fixed_particle = -1;  // fixed_particle_step_handler: increment_v of this particle is dropped (lg_set_fixed_particle)
    int integrator = 0;       // 0 = rotational_velocity_verlet_half, 1 = rotational_forward_euler (lg_set_integrator)
    bool simple_only = false; // a SURVEY 8f variant is active (alt_sinter, forward Euler, fixed particle, central attraction):
                              // the system runs the reference's schedule literally (kernel variant 0, exact arithmetic)
    bool central_on = false;  // central_attraction_functor (lg_set_central_attraction)
    double central_k = 0.0;
    double central_x0[3] = {0.0, 0.0, 0.0};
    DevBuf<double> bspr;      // alt_sinter_functor: three springs (9 doubles) per ordered bond record
    // Spatial renumbering (lg_set_reorder): internal index k holds the caller's particle perm[k]; inv is the inverse.
    // Fixed by the first lg_set_state; every entry point that takes or returns particle indices / per-particle arrays maps.
    int reorder_mode = 0;
    bool reordered = false;
    DevBuf<int32_t> d_perm, d_inv;
    std::vector<int32_t> h_perm, h_inv;
    DevBuf<double> io_scratch;  // 3n doubles: caller-order staging of lg_get_field / lg_set_state

    PingPong x, v, w;         // read by other particles' pair evaluations -> double-buffered
    DevBuf<double> a, th, al;  // private to their particle -> single
    bool state_set = false;
    bool kicked = false;
    bool half_kicked = false;  // lg_first_half_kick ran: the first do_step resumes at its second evaluation
    DevBuf<uint8_t> owned;     // slab decomposition: 1 = owned, 0 = ghost (rows stay empty); unused when !has_owned
    bool has_owned = false;
    // Lazy evaluation (variant 1): the reference ends do_step with a force evaluation at the new state
    // (rotational_velocity_verlet_half.h:75); the fused kernel performs that evaluation at the START of the
    // next step instead.  `pending` means a, alpha (and the contact springs) have not yet been advanced to
    // the current x, v, omega; flush() materialises them whenever something could observe the difference.
    bool pending = false;
    std::vector<Facet> pending_facets;  // facet state as of the do_step call that left the evaluation pending

    // neighbor rows
    bool use_nl = false;
    double r_verlet = 0.0;
    bool allpairs_built = false;
    DevBuf<uint32_t> row_ptr;
    DevBuf<int32_t> cols;
    DevBuf<uint8_t> tag;
    int64_t n_pairs = 0;
    // Rows rebuilt while a closing evaluation was still pending: that evaluation belongs to the OLD rows, so the new
    // ones wait here and are installed (swap + retag) right after it has run -- the step after a rebuild then stays
    // ONE fused launch instead of force-only + integrate (rebuild(), evaluate()).
    DevBuf<uint32_t> row_ptr_new;
    DevBuf<int32_t> cols_new;
    DevBuf<uint8_t> tag_new;
    int64_t n_pairs_new = 0;
    bool rows_staged = false;

    // contact slots
    int S = 16;
    DevBuf<double> spr;
    DevBuf<int32_t> sgid;
    bool slots_ready = false;

    // bonds
    int64_t n_bonds = 0;  // ordered
    DevBuf<uint32_t> bptr;
    DevBuf<int32_t> bj, brev;
    PingPong bconn;
    DevBuf<double> blen;
    std::vector<uint32_t> h_bptr;
    std::vector<int32_t> h_bj;
    // step_rows2_kernel: per tile of rows2_tile_particles() particles, the bridges to evaluate (build_bridge_work)
    DevBuf<uint16_t> bwork;
    DevBuf<uint32_t> bwork_cnt;
    int bwork_shape = -1;            // tile shape the lists were built for (-1: none)
    // slab runs (lg_run_slab): the halo whose stores the fused step kernel performs itself (kernels_rows.cu, HaloFusedArgs)
    struct lg_halo* active_halo = nullptr;
    bool halo_stored_by_step = false;  // the last step launch has stored and published the halo: no push kernel
    // tile shape of step_rows2_kernel: heuristic (deterministic) or, after lg_set_autotune(1), measured on the first fused launches
    int autotune = 0;
    int forced_shape = -1;           // lg_set_rows_shape: the caller pins the shape (slab runs: one shape on every rank)
    int tuned_shape = -1;
    double tuned_k = 0.0;
    int tune_launch = 0;             // 0 .. 3 * kTuneReps - 1 while measuring
    float tune_ms[3] = {0.f, 0.f, 0.f};
    cudaEvent_t tune_ev[2] = {nullptr, nullptr};
    std::vector<int32_t> h_brev;     // host copies the lists are rebuilt from when the shape changes
    std::vector<uint8_t> h_pairable; // record and its mirror carry the same rest lengths

    // facets
    std::vector<Facet> facets;
    std::vector<std::unique_ptr<DevBuf<double>>> wspr;
    std::vector<std::unique_ptr<DevBuf<uint8_t>>> wact;

    // cell-list scratch
    DevBuf<double> bbox_scratch, bbox_out;
    DevBuf<int32_t> cell_of, order;
    DevBuf<int32_t> row_hits;  // rebuild scratch: pairs found by the count pass, [entry][particle]
    DevBuf<uint32_t> cell_start, cell_cursor, scan_scratch;
    int64_t last_cells = 0;
    // The cell grid of the previous rebuild is reused for kGridReuse rebuilds: cell_coords clamps, so a grid that no
    // longer covers every particle still yields the exact rows (the cell list is only a candidate filter, cell edge
    // >= r_verlet), and the bounding-box kernels plus their host read-back drop out of most rebuilds.
    NeighborGrid grid{};
    bool grid_valid = false;
    int grid_age = 0;
    // Steady-state rebuilds make no host round trip: the bounding box that refreshes the grid and
    // the pair count that sizes the column buffers are read back asynchronously and consumed at a later rebuild; a
    // device-side guard (launch_pair_guard) keeps an unexpectedly large list from overrunning the buffers.
    cudaEvent_t ev_bbox = nullptr, ev_pairs = nullptr;
    bool bbox_pending = false, pairs_pending = false;
    int64_t last_total = -1;          // exact pair count of the most recent rebuild whose read-back has been consumed
    DevBuf<uint32_t> pair_total;      // [0] pairs of the last guarded rebuild, [1] overflow flag

    // diagnostics scratch
    DevBuf<double> diag_partials, diag_out;
    DevBuf<unsigned long long> counter;

    DevBuf<int> err;
    double* h_pinned = nullptr;  // 32 doubles of pinned scratch

    int64_t steps = 0, rebuilds = 0;
    int64_t launches0 = 0;
    const char* last_kernel = "";  // lg_last_step_kernel

    // optional per-kernel timing (lg_set_profiling)
    bool profiling = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_events;
    size_t prof_used = 0;
    double prof_ms = 0.0;
    int64_t prof_count = 0;

    // Releases what the DevBuf members do not own; runs on every exit path of lg_create as well (error returns
    // destroy the half-built system through its unique_ptr).  The caller makes `device` current first.
    ~lg_system() {
        if (stream) {
            cudaStreamSynchronize(stream);
            cudaStreamDestroy(stream);
        }
        if (h_pinned) cudaFreeHost(h_pinned);
        if (ev_bbox) cudaEventDestroy(ev_bbox);
        if (ev_pairs) cudaEventDestroy(ev_pairs);
        for (auto& ev : tune_ev)
            if (ev) cudaEventDestroy(ev);
        for (auto& ev : prof_events) {
            cudaEventDestroy(ev.first);
            cudaEventDestroy(ev.second);
        }
    }
};

namespace {

int check_async(lg_system* s) {
    int h = 0;
    LG_CUDA(cudaMemcpyAsync(&h, s->err.p, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
    LG_CUDA(cudaStreamSynchronize(s->stream));
    if (h & ERR_SLOT_OVERFLOW)
        return fail(LG_ECAPACITY, "a particle has more simultaneous contacts than lg_set_max_contacts allows");
    if (h & ERR_BOND_LOOKUP) return fail(LG_ECAPACITY, "bond table lookup failed");
    if (h & ERR_PAIR_CAPACITY)
        return fail(LG_ECAPACITY, "the neighbor list grew by more than 15 % within one rebuild period and did not fit its buffers; "
                                  "rows were emptied (call lg_update_neighbor_list again after lg_synchronize)");
    if (h & ERR_HALO_TIMEOUT) return fail(LG_ECUDA, "halo exchange: a neighbouring rank did not publish its message in time");
    return LG_OK;
}

int ensure_slots(lg_system* s) {
    if (s->slots_ready) return LG_OK;
    const size_t n = (size_t)s->n;
    LG_CUDA(s->spr.reserve(n * s->S * 9));
    LG_CUDA(s->sgid.reserve(n * s->S));
    LG_CUDA(cudaMemsetAsync(s->sgid.p, 0xff, n * s->S * sizeof(int32_t), s->stream));
    s->slots_ready = true;
    return LG_OK;
}

int ensure_pair_capacity(lg_system* s, int64_t pairs) {
    if ((size_t)pairs > s->cols.cap) {
        LG_CUDA(s->cols.reserve((size_t)pairs, 1.25));
        LG_CUDA(s->tag.reserve(s->cols.cap));
    }
    return LG_OK;
}

int retag(lg_system* s) {
    LG_TRY(ensure_slots(s));
    launch_retag((int)s->n, s->row_ptr.p, s->cols.p, s->tag.p, s->sgid.p, s->spr.p, s->S,
                 s->n_bonds > 0 ? s->bptr.p : nullptr, s->bj.p, s->err.p, s->stream);
    LG_CUDA(cudaGetLastError());
    return LG_OK;
}

template <typename T>
void swap_buf(DevBuf<T>& a, DevBuf<T>& b) {
    std::swap(a.p, b.p);
    std::swap(a.cap, b.cap);
}

int install_staged_rows(lg_system* s) {
    swap_buf(s->row_ptr, s->row_ptr_new);
    swap_buf(s->cols, s->cols_new);
    swap_buf(s->tag, s->tag_new);
    s->n_pairs = s->n_pairs_new;
    s->rows_staged = false;
    return retag(s);
}

int build_allpairs(lg_system* s) {
    const int64_t n = s->n;
    if (n > 65536) return fail(LG_ECAPACITY, "all-pairs loop supports n <= 65536; call lg_enable_neighbor_list");
    const int64_t pairs = n * (n - 1);
    LG_TRY(ensure_pair_capacity(s, pairs));
    if (n == 1) {
        LG_CUDA(cudaMemsetAsync(s->row_ptr.p, 0, 2 * sizeof(uint32_t), s->stream));
    } else {
        launch_allpairs_fill((int)n, s->row_ptr.p, s->cols.p, s->stream);
        LG_CUDA(cudaGetLastError());
    }
    s->n_pairs = pairs;
    s->allpairs_built = true;
    return retag(s);
}

int build_bridge_work(lg_system* s, int shape);  // below: work lists of step_rows2_kernel's bridge pass
int fill_halo_fused(lg_system* s, ForceArgs& A);  // below: halo stores fused into step_rows2_kernel

int drain_profile(lg_system* s) {
    LG_CUDA(cudaStreamSynchronize(s->stream));
    for (size_t k = 0; k < s->prof_used; k++) {
        float ms = 0.f;
        LG_CUDA(cudaEventElapsedTime(&ms, s->prof_events[k].first, s->prof_events[k].second));
        s->prof_ms += (double)ms;
        s->prof_count++;
    }
    s->prof_used = 0;
    return LG_OK;
}

int rows_per_block(const lg_system* s) {
    // aim at one (almost full) tile of pairs per CTA
    const double k_mean = s->n > 0 ? (double)s->n_pairs / (double)s->n : 1.0;
    static const double fill = [] {  // tuning knob: fraction of the tile the average CTA fills
        const char* e = std::getenv("LGB200_TILE_FILL");
        const double f = e ? std::atof(e) : 0.0;
        return (f > 0.05 && f <= 1.0) ? f : 0.977;
    }();
    int r = (int)std::lround(fill * tiled_tile_pairs() / std::max(k_mean, 1.0));
    r = std::max(8, std::min(r, tiled_max_rows_per_block()));
    return r;
}

// One launch of the pair-force kernel at the current state.
//   fused_dt < 0 : force only, a / alpha <- result (CLEAN afterwards)
//   fused_dt >= 0: variant 1 only; force at the current state, then integrate with fused_dt into the next-step
//                  buffers and swap them
constexpr int kTuneReps = 3;

int evaluate(lg_system* s, const std::vector<Facet>& facets, double fused_dt = -1.0) {
    ForceArgs A{};
    bool tuning = false;
    A.n = (int)s->n;
    A.x = s->x.p; A.v = s->v.p; A.w = s->w.p; A.a = s->a.p; A.al = s->al.p;
    A.row_ptr = s->row_ptr.p; A.cols = s->cols.p; A.tag = s->tag.p;
    A.spr = s->spr.p; A.sgid = s->sgid.p; A.S = s->S;
    A.bptr = s->n_bonds > 0 ? s->bptr.p : nullptr;
    A.bj = s->bj.p; A.brev = s->brev.p; A.bconn = s->bconn.p; A.blen = s->blen.p;
    A.bwork = s->bwork.p; A.bwork_cnt = s->bwork_cnt.p;
    A.bspr = s->bspr.p;
    A.fixed = s->fixed_particle;
    A.central_on = s->central_on ? 1 : 0;
    A.central_k = s->central_k;
    A.central_x0 = mk3(s->central_x0[0], s->central_x0[1], s->central_x0[2]);
    A.mp = s->mp;
    A.n_facets = (int)facets.size();
    for (int f = 0; f < A.n_facets; f++) {
        A.facets[f] = facets[(size_t)f];
        A.wspr[f] = s->wspr[(size_t)f]->p;
        A.wact[f] = s->wact[(size_t)f]->p;
    }
    A.err = s->err.p;
    A.rows_per_block = rows_per_block(s);
    A.fast = s->math_mode;
    {   // short rows (aggregates) -> row kernel; long rows (packings) -> tile kernel.  LGB200_ROWS_KERNEL=0 disables it.
        static const int rows_enabled = [] { const char* e = std::getenv("LGB200_ROWS_KERNEL"); return e ? std::atoi(e) : 2; }();  // 2 = step_rows2_kernel, 1 = step_rows_kernel, 0 = tile kernel
        const double k_mean = s->n > 0 ? (double)s->n_pairs / (double)s->n : 0.0;
        static const double rows_kmax = [] { const char* e = std::getenv("LGB200_ROWS_KMAX"); return e ? std::atof(e) : 1.0e9; }();  // mean row length above which the tile kernel takes over
        A.use_rows = (s->math_mode && (s->variant == 2 || (rows_enabled && k_mean <= (rows_enabled == 1 ? 24.0 : rows_kmax)))) ? (rows_enabled == 1 ? 1 : 2) : 0;  // variant 2: row kernel whatever the row length
        A.rows_shape = s->forced_shape >= 0 ? s->forced_shape : rows2_shape_for(k_mean, (int)s->n, s->n_bonds > 0);
        if (A.use_rows == 2 && s->autotune && s->forced_shape < 0 && std::getenv("LGB200_ROWS_SHAPE") == nullptr) {
            if (s->tuned_shape >= 0 && std::fabs(k_mean - s->tuned_k) <= 0.25 * s->tuned_k) {
                A.rows_shape = s->tuned_shape;
            } else if (fused_dt >= 0.0) {  // measure: kTuneReps fused launches per shape, the first of each discarded
                if (s->tuned_shape >= 0) { s->tuned_shape = -1; s->tune_launch = 0; }
                A.rows_shape = s->tune_launch / kTuneReps;
                tuning = true;
            }
        }
        if (A.use_rows == 2 && s->n_bonds > 0 && s->bwork_shape != A.rows_shape) {
            LG_TRY(build_bridge_work(s, A.rows_shape));
            A.bwork = s->bwork.p; A.bwork_cnt = s->bwork_cnt.p;
        }
    }
    A.fc = fast_consts(s->mp.hamaker_A, s->mp.hamaker_h0, s->mp.r, s->mp.mass, s->mp.inertia, s->mp.hamaker_kind, s->mp.hamaker_cutoff);
    const bool fused = fused_dt >= 0.0;
    A.dt = fused ? fused_dt : 0.0;
    s->halo_stored_by_step = false;
    if (fused && A.fast && A.use_rows == 2 && s->active_halo != nullptr) LG_TRY(fill_halo_fused(s, A));
    if (tuning) {
        if (!s->tune_ev[0]) {
            LG_CUDA(cudaEventCreate(&s->tune_ev[0]));
            LG_CUDA(cudaEventCreate(&s->tune_ev[1]));
        }
        LG_CUDA(cudaEventRecord(s->tune_ev[0], s->stream));
    }
    A.x_next = s->x.next; A.v_next = s->v.next; A.w_next = s->w.next; A.th = s->th.p; A.bconn_next = s->bconn.next;
    if (s->profiling) {
        if (s->prof_used == s->prof_events.size()) {
            if (s->prof_used >= 4096) LG_TRY(drain_profile(s));
            if (s->prof_used == s->prof_events.size()) {
                cudaEvent_t e0, e1;
                LG_CUDA(cudaEventCreate(&e0));
                LG_CUDA(cudaEventCreate(&e1));
                s->prof_events.emplace_back(e0, e1);
            }
        }
        LG_CUDA(cudaEventRecord(s->prof_events[s->prof_used].first, s->stream));
    }
    if (fused)
        launch_step_tiled(A, s->mask, true, s->stream);
    else
        launch_force(A, s->mask, s->variant, s->stream);
    s->last_kernel = s->variant == 0 ? "force_simple_kernel" : (A.fast && A.use_rows) ? (A.use_rows == 2 ? "step_rows2_kernel" : "step_rows_kernel") : "step_tiled_kernel";
    if (s->profiling) LG_CUDA(cudaEventRecord(s->prof_events[s->prof_used++].second, s->stream));
    LG_CUDA(cudaGetLastError());
    if (tuning) {
        LG_CUDA(cudaEventRecord(s->tune_ev[1], s->stream));
        LG_CUDA(cudaEventSynchronize(s->tune_ev[1]));
        float ms = 0.f;
        LG_CUDA(cudaEventElapsedTime(&ms, s->tune_ev[0], s->tune_ev[1]));
        const int shape = s->tune_launch / kTuneReps, rep_idx = s->tune_launch % kTuneReps;
        if (rep_idx == 1 || (rep_idx > 1 && ms < s->tune_ms[shape])) s->tune_ms[shape] = ms;
        if (++s->tune_launch == 3 * kTuneReps) {
            int best = 0;
            for (int k = 1; k < 3; k++)
                if (s->tune_ms[k] < s->tune_ms[best]) best = k;
            s->tuned_shape = best;
            s->tuned_k = s->n > 0 ? (double)s->n_pairs / (double)s->n : 0.0;
            s->tune_launch = 0;
        }
    }
    if (fused) {
        s->x.swap();
        s->v.swap();
        s->w.swap();
        if (s->n_bonds > 0) s->bconn.swap();
    }
    if (s->rows_staged) LG_TRY(install_staged_rows(s));  // this was the evaluation the old rows were kept for
    return LG_OK;
}

// Materialise the evaluation a previous do_step left pending (a, alpha, contact springs at the current state).
int flush(lg_system* s) {
    if (!s->pending) return LG_OK;
    LG_TRY(evaluate(s, s->pending_facets));
    s->pending = false;
    return LG_OK;
}

int prepare(lg_system* s) {
    if (!s->state_set) return fail(LG_EINVAL, "lg_set_state has not been called");
    if (s->mask == 0 && s->facets.empty()) return fail(LG_EINVAL, "no force model installed");
    if ((s->mask & MODEL_CONTACT) && (s->mask & MODEL_SINTER))
        return fail(LG_EINVAL, "contact and sinter models are mutually exclusive (sinter embeds contact)");
    LG_TRY(ensure_slots(s));
    if (!s->use_nl && !s->allpairs_built) LG_TRY(build_allpairs(s));
    return LG_OK;
}

int integrate_in_place(lg_system* s, double dt) {
    IntegrateArgs I{};
    I.n = (int)s->n;
    I.x = s->x.p; I.v = s->v.p; I.th = s->th.p; I.w = s->w.p; I.a = s->a.p; I.al = s->al.p;
    I.bptr = (s->n_bonds > 0 && !s->mp.alt_bonds) ? s->bptr.p : nullptr;  // alt_sinter bonds carry no connectors
    I.bconn = s->bconn.p;
    I.n_bonds = s->n_bonds;
    I.dt = dt;
    I.scheme = s->integrator;
    I.fixed = s->fixed_particle;
    launch_integrate(I, s->stream);
    LG_CUDA(cudaGetLastError());
    return LG_OK;
}

int step(lg_system* s, double dt) {
    if (s->integrator == 1) {
        // rotational_forward_euler.h:44-62: update_acceleration(), then the two-term Taylor update; no half kick, and
        // a / alpha stay those of the state BEFORE the step, as get_a() of the reference shows them
        LG_TRY(evaluate(s, s->facets));
        LG_TRY(integrate_in_place(s, dt));
        s->steps++;
        return LG_OK;
    }
    if (!s->kicked) {
        // first call: evaluate, half kick back, evaluate again (rotational_velocity_verlet_half.h:45-60)
        if (!s->half_kicked) {
            LG_TRY(evaluate(s, s->facets));
            launch_half_kick_back((int)s->n, s->v.p, s->w.p, s->a.p, s->al.p, dt, s->fixed_particle, s->stream);
            LG_CUDA(cudaGetLastError());
        }
        LG_TRY(evaluate(s, s->facets));
        s->kicked = true;
        s->half_kicked = false;
        s->pending = false;
    }
    if (s->variant == 0) {
        // validation path: the reference's schedule literally -- integrate, then evaluate
        LG_TRY(integrate_in_place(s, dt));
        LG_TRY(evaluate(s, s->facets));
    } else if (!s->pending) {
        // a, alpha are current: integrate with them; the closing evaluation is left pending
        LG_TRY(integrate_in_place(s, dt));
        s->pending = true;
    } else {
        // steady state: ONE kernel = (pending evaluation of the previous step) + (this step's integration)
        LG_TRY(evaluate(s, s->pending_facets, dt));
    }
    if (s->pending) s->pending_facets = s->facets;  // the closing evaluation sees the facets as they are now
    s->steps++;
    return LG_OK;
}

// The exact pair count of the most recent rebuild, if its asynchronous read-back is still outstanding.
int sync_pairs(lg_system* s) {
    if (!s->pairs_pending) return LG_OK;
    LG_CUDA(cudaEventSynchronize(s->ev_pairs));
    const int64_t v = (int64_t)reinterpret_cast<const uint32_t*>(s->h_pinned + 16)[0];
    (s->rows_staged ? s->n_pairs_new : s->n_pairs) = v;
    s->last_total = v;
    s->pairs_pending = false;
    return LG_OK;
}

int grid_from_bbox(lg_system* s, const double* bb) {
    const int n = (int)s->n;
    NeighborGrid g{};
    double cell = s->r_verlet * (1.0 + 1e-6);
    const double max_cells = 16.0 * (double)n + 65536.0;
    for (;;) {
        double total = 1.0;
        for (int k = 0; k < 3; k++) {
            double ext = bb[3 + k] - bb[k];
            if (!(ext >= 0.0) || !std::isfinite(ext)) return fail(LG_EINVAL, "non-finite particle positions");
            double d = std::floor(ext / cell) + 1.0;
            g.dim[k] = d > 2.0e9 ? 2000000000 : (int)d;
            total *= d;
        }
        if (total <= max_cells) {
            g.n_cells = (int64_t)g.dim[0] * g.dim[1] * g.dim[2];
            break;
        }
        cell *= 1.26;
    }
    g.cell = cell;
    for (int k = 0; k < 3; k++) g.lo[k] = bb[k];
    s->grid = g;
    s->grid_valid = true;
    s->grid_age = 0;
    return LG_OK;
}

int rebuild(lg_system* s) {
    if (!s->use_nl) return LG_OK;  // granular_system (all pairs) has no update_neighbor_list
    if (!s->state_set) return fail(LG_EINVAL, "lg_set_state has not been called");
    LG_TRY(ensure_slots(s));
    LG_TRY(sync_pairs(s));  // the previous rebuild's count: long since arrived in the steady state
    // The pending evaluation belongs to the OLD rows (it closed the previous do_step).  In the steady state of the
    // fused variant it stays pending: the new rows are built next to the old ones (they depend on x only, which
    // the evaluation does not change) and are installed by evaluate() once the evaluation has run.
    const bool defer = s->pending && s->variant >= 1 && s->rebuilds > 0 && s->n_pairs > 0;
    if (!defer) LG_TRY(flush(s));
    const int n = (int)s->n;
    if (s->variant >= 1) {  // the second set of row buffers is allocated up front: no cudaMalloc in the steady state
        LG_CUDA(s->row_ptr_new.reserve((size_t)n + 1));
        if (s->cols.cap > s->cols_new.cap) {
            LG_CUDA(s->cols_new.reserve(s->cols.cap - 16));
            LG_CUDA(s->tag_new.reserve(s->cols_new.cap));
        }
    }
    DevBuf<uint32_t>& row_ptr = defer ? s->row_ptr_new : s->row_ptr;
    DevBuf<int32_t>& cols = defer ? s->cols_new : s->cols;
    DevBuf<uint8_t>& tag = defer ? s->tag_new : s->tag;
    cudaStream_t st = s->stream;
    if (!s->ev_bbox) {
        LG_CUDA(cudaEventCreateWithFlags(&s->ev_bbox, cudaEventDisableTiming));
        LG_CUDA(cudaEventCreateWithFlags(&s->ev_pairs, cudaEventDisableTiming));
        LG_CUDA(s->pair_total.reserve(2));
    }
    // 1. cell grid.  cell_coords clamps, so a grid that no longer covers every particle still yields the exact rows (the
    //    cell list is only a candidate filter, cell edge >= r_verlet): the grid is refreshed from a bounding box that
    //    was measured at an EARLIER rebuild and has arrived in pinned memory since -- no wait.  Only the first rebuild
    //    after new positions were set measures and waits.
    constexpr int kGridRefresh = 4;
    double* h_bbox = s->h_pinned + 20;  // its own pinned slots: the copy may land at any later time
    if (!s->grid_valid) {
        if (s->bbox_pending) {  // a measurement of positions that have been replaced since: discard
            LG_CUDA(cudaEventSynchronize(s->ev_bbox));
            s->bbox_pending = false;
        }
        launch_bbox(n, s->x.p, s->bbox_scratch.p, s->bbox_out.p, st);
        LG_CUDA(cudaGetLastError());
        LG_CUDA(cudaMemcpyAsync(h_bbox, s->bbox_out.p, 6 * sizeof(double), cudaMemcpyDeviceToHost, st));
        LG_CUDA(cudaStreamSynchronize(st));
        LG_TRY(grid_from_bbox(s, h_bbox));
    } else if (s->bbox_pending) {
        if (cudaEventQuery(s->ev_bbox) == cudaSuccess) {
            s->bbox_pending = false;
            LG_TRY(grid_from_bbox(s, h_bbox));
        }
    } else if (++s->grid_age >= kGridRefresh) {
        launch_bbox(n, s->x.p, s->bbox_scratch.p, s->bbox_out.p, st);
        LG_CUDA(cudaGetLastError());
        LG_CUDA(cudaMemcpyAsync(h_bbox, s->bbox_out.p, 6 * sizeof(double), cudaMemcpyDeviceToHost, st));
        LG_CUDA(cudaEventRecord(s->ev_bbox, st));
        s->bbox_pending = true;
        s->grid_age = 0;
    }
    const NeighborGrid g = s->grid;
    s->last_cells = g.n_cells;
    // 2. histogram, scan, scatter
    const size_t nc1 = (size_t)g.n_cells + 1;
    LG_CUDA(s->cell_start.reserve(nc1, 1.25));
    LG_CUDA(s->cell_cursor.reserve(nc1, 1.25));
    LG_CUDA(s->scan_scratch.reserve(scan_scratch_elems(std::max(nc1, (size_t)n + 1)), 1.25));
    LG_CUDA(cudaMemsetAsync(s->cell_start.p, 0, nc1 * sizeof(uint32_t), st));
    LG_CUDA(cudaMemsetAsync(s->cell_cursor.p, 0, nc1 * sizeof(uint32_t), st));
    launch_cell_count(n, s->x.p, g, s->cell_of.p, s->cell_start.p, st);
    launch_exclusive_scan_u32(s->cell_start.p, nc1, s->scan_scratch.p, st);
    launch_cell_fill(n, s->cell_of.p, s->cell_start.p, s->cell_cursor.p, s->order.p, st);
    // 3. rows: count, scan, fill + sort
    const uint8_t* owned = s->has_owned ? s->owned.p : nullptr;
    // the pairs the count pass finds travel to the fill pass in a hit list (no second sweep of the cells)
    LG_CUDA(s->row_hits.reserve((size_t)n * (size_t)neighbor_hit_list_entries()));
    launch_row_count(n, s->x.p, g, s->cell_start.p, s->order.p, s->r_verlet, row_ptr.p, s->row_hits.p, owned, st);
    LG_CUDA(cudaMemsetAsync(row_ptr.p + n, 0, sizeof(uint32_t), st));
    launch_exclusive_scan_u32(row_ptr.p, (size_t)n + 1, s->scan_scratch.p, st);
    LG_CUDA(cudaGetLastError());
    // The pair count sizes the column buffers.  With 15 % of headroom over the previous rebuild's count the buffers are
    // kept, the count is read back asynchronously (consumed by the next rebuild or by whoever asks for it) and a
    // device-side guard covers the impossible-in-practice case of a list that grew faster; otherwise wait for it.
    uint32_t* h_total = reinterpret_cast<uint32_t*>(s->h_pinned + 16);
    const bool async_ok = defer && s->rebuilds >= 2 && s->last_total > 0 &&
                          (double)cols.cap >= 1.15 * (double)s->last_total + 1024.0;
    int64_t total_pairs = 0;
    if (async_ok) {
        launch_pair_guard(n, row_ptr.p, (uint32_t)std::min<size_t>(cols.cap, 0xffffffffu), s->pair_total.p, s->err.p, st);
        LG_CUDA(cudaMemcpyAsync(h_total, s->pair_total.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        LG_CUDA(cudaEventRecord(s->ev_pairs, st));
        s->pairs_pending = true;
        total_pairs = s->last_total;  // estimate until the read-back is consumed (sync_pairs)
    } else {
        LG_CUDA(cudaMemcpyAsync(h_total, row_ptr.p + n, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
        LG_CUDA(cudaStreamSynchronize(st));
        total_pairs = (int64_t)h_total[0];
        s->last_total = total_pairs;
        if ((size_t)total_pairs > cols.cap || (double)cols.cap < 1.15 * (double)total_pairs + 1024.0) {
            LG_CUDA(cols.reserve((size_t)((double)total_pairs * 1.3) + 2048));
            LG_CUDA(tag.reserve(cols.cap));
        }
    }
    launch_row_fill(n, s->x.p, g, s->cell_start.p, s->order.p, s->r_verlet, row_ptr.p, cols.p, s->row_hits.p, owned, st);
    LG_CUDA(cudaGetLastError());
    s->rebuilds++;
    if (defer) {
        s->n_pairs_new = total_pairs;
        s->rows_staged = true;
        return LG_OK;
    }
    s->n_pairs = total_pairs;
    return retag(s);
}

// ---- host-side bond CSR construction ---------------------------------------------------------------
struct HostBond {
    int64_t i, j;
    double conn[15];
    double len[5];
};

}  // namespace

struct lg_bond_table {
    int64_t n = 0;
    int64_t bonds = 0;  // unordered
    std::vector<HostBond> recs;  // ordered records, two per bond
    bool alt = false;            // built by lg_alt_bond_table_create: no connectors, cycle-free
};

namespace {

// Work lists of step_rows2_kernel's bridge pass.  Inside a tile of T consecutive particles every bridge whose two ends
// both belong to the tile is evaluated ONCE, by the record with the lower index: the springs of (j, i) are those of (i, j)
// reversed, so force_j = -force_i exactly and only the torque arm differs (sinter_bridge.h:244-256).  Item = tile-relative
// record index, bit 15 set when the item also produces its mirror's value.  Records whose mirror lies in another tile, or
// whose two orientations carry different rest lengths (possible with explicit tables), stay one-sided.  Tiles with more
// records than the kernel stages get an empty list (the kernel evaluates them from global memory).
int build_bridge_work(lg_system* s, int shape) {
    const std::vector<uint32_t>& bptr = s->h_bptr;
    const std::vector<int32_t>& brev = s->h_brev;
    const int T = rows2_tile_particles(shape), cap = rows2_tile_records(shape);
    const size_t n = (size_t)s->n, nb = brev.size();
    const size_t n_tiles = (n + (size_t)T - 1) / (size_t)T;
    std::vector<uint16_t> work(nb + 16, 0);
    std::vector<uint32_t> cnt(n_tiles + 1, 0);
    for (size_t t = 0; t < n_tiles; t++) {
        const size_t i0 = t * (size_t)T, i1 = std::min(n, i0 + (size_t)T);
        const uint32_t b0 = bptr[i0], b1 = bptr[i1];
        if (b1 - b0 > (uint32_t)cap) continue;
        uint32_t w = 0;
        for (uint32_t b = b0; b < b1; b++) {
            const uint32_t m = (uint32_t)brev[b];
            const bool in_tile = m >= b0 && m < b1;
            if (in_tile && s->h_pairable[b]) {
                if (b < m) work[b0 + w++] = (uint16_t)((b - b0) | 0x8000u);
            } else {
                work[b0 + w++] = (uint16_t)(b - b0);
            }
        }
        cnt[t] = w;
    }
    LG_CUDA(s->bwork.reserve(nb + 16));
    LG_CUDA(s->bwork_cnt.reserve(n_tiles + 1));
    LG_CUDA(cudaMemcpyAsync(s->bwork.p, work.data(), (nb + 16) * sizeof(uint16_t), cudaMemcpyHostToDevice, s->stream));
    LG_CUDA(cudaMemcpyAsync(s->bwork_cnt.p, cnt.data(), (n_tiles + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, s->stream));
    LG_CUDA(cudaStreamSynchronize(s->stream));
    s->bwork_shape = shape;
    return LG_OK;
}

// host copies the work lists are derived from; the lists themselves are built for the tile shape of the first launch
void keep_bond_mirrors(lg_system* s, std::vector<int32_t>&& brev, const std::vector<double>& len) {
    const size_t nb = brev.size();
    s->h_pairable.assign(nb, 0);
    for (size_t b = 0; b < nb; b++) {
        const size_t m = (size_t)brev[b];
        s->h_pairable[b] = m < nb && std::memcmp(&len[5 * b], &len[5 * m], 5 * sizeof(double)) == 0;
    }
    s->h_brev = std::move(brev);
    s->bwork_shape = -1;
}

int install_bonds(lg_system* s, std::vector<HostBond>& recs) {
    if (s->reordered)  // the caller's indices -> internal numbering (connectors are absolute positions: untouched)
        for (auto& r : recs) {
            r.i = (int64_t)s->h_inv[(size_t)r.i];
            r.j = (int64_t)s->h_inv[(size_t)r.j];
        }
    // CSR by (i, j) ascending
    std::sort(recs.begin(), recs.end(), [](HostBond const& a, HostBond const& b) { return a.i != b.i ? a.i < b.i : a.j < b.j; });
    const size_t nb = recs.size();
    const size_t n = (size_t)s->n;
    std::vector<uint32_t> bptr(n + 1, 0);
    for (auto const& r : recs) bptr[(size_t)r.i + 1]++;
    for (size_t i = 0; i < n; i++) {
        if (bptr[i + 1] > (uint32_t)kMaxBondsPerParticle)
            return fail(LG_ECAPACITY, "more than 120 bonds on one particle");
        bptr[i + 1] += bptr[i];
    }
    std::vector<int32_t> bj(nb), brev(nb);
    std::vector<double> conn(nb * 15), len(nb * 5);
    for (size_t b = 0; b < nb; b++) {
        bj[b] = (int32_t)recs[b].j;
        std::memcpy(&conn[b * 15], recs[b].conn, sizeof(recs[b].conn));
        std::memcpy(&len[b * 5], recs[b].len, sizeof(recs[b].len));
        if (b > 0 && recs[b].i == recs[b - 1].i && recs[b].j == recs[b - 1].j) return fail(LG_EINVAL, "duplicate bond record");
    }
    for (size_t b = 0; b < nb; b++) {
        const size_t j = (size_t)recs[b].j;
        const int32_t want = (int32_t)recs[b].i;
        auto first = bj.begin() + bptr[j], last = bj.begin() + bptr[j + 1];
        auto it = std::lower_bound(first, last, want);
        if (it == last || *it != want) return fail(LG_EINVAL, "bond (i, j) given without its mirror (j, i)");
        brev[b] = (int32_t)(it - bj.begin());
    }
    LG_CUDA(s->bptr.reserve(n + 1));
    LG_CUDA(s->bj.reserve(nb));
    LG_CUDA(s->brev.reserve(nb));
    LG_CUDA(s->bconn.reserve(nb * 15 + 16));
    LG_CUDA(s->blen.reserve(nb * 5));
    LG_CUDA(cudaMemcpyAsync(s->bptr.p, bptr.data(), (n + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, s->stream));
    if (nb) {
        LG_CUDA(cudaMemcpyAsync(s->bj.p, bj.data(), nb * sizeof(int32_t), cudaMemcpyHostToDevice, s->stream));
        LG_CUDA(cudaMemcpyAsync(s->brev.p, brev.data(), nb * sizeof(int32_t), cudaMemcpyHostToDevice, s->stream));
        LG_CUDA(cudaMemcpyAsync(s->bconn.p, conn.data(), nb * 15 * sizeof(double), cudaMemcpyHostToDevice, s->stream));
        LG_CUDA(cudaMemcpyAsync(s->blen.p, len.data(), nb * 5 * sizeof(double), cudaMemcpyHostToDevice, s->stream));
    }
    LG_CUDA(cudaStreamSynchronize(s->stream));
    keep_bond_mirrors(s, std::move(brev), len);
    s->n_bonds = (int64_t)nb;
    s->h_bptr = std::move(bptr);
    s->h_bj = std::move(bj);
    // existing rows (if any) need their tags refreshed
    LG_TRY(flush(s));  // installs rows that were rebuilt behind a pending evaluation
    if (s->n_pairs > 0 || s->allpairs_built) LG_TRY(retag(s));
    return LG_OK;
}

// Host-side restatement of the bond scan of sinter_functor's constructor (sinter_bridge.h:148-211) on a uniform
// grid instead of the O(N^2) double loop; pairs are still visited in lexicographic (i < j) order so the
// generator is consumed exactly like the reference.  Needs no GPU.
// Uniform host grid over the positions: candidates(i) = particles j > i in the 27 cells around i, ascending, i.e. the
// partners the reference's i < j double loops (sinter_bridge.h:148-149, alt_sinter_bridge.h:65-66) could accept within `range`.
struct HostGrid {
    const std::vector<double>& hx;
    double lo[3], cell;
    int64_t dim[3];
    std::vector<int64_t> start, order;
    HostGrid(int64_t n, const std::vector<double>& hx_, double range) : hx(hx_) {
        double hi[3] = {-1e300, -1e300, -1e300};
        lo[0] = lo[1] = lo[2] = 1e300;
        for (int64_t i = 0; i < n; i++)
            for (int k = 0; k < 3; k++) {
                lo[k] = std::min(lo[k], hx[3 * (size_t)i + k]);
                hi[k] = std::max(hi[k], hx[3 * (size_t)i + k]);
            }
        cell = range * 1.0001;
        for (;;) {
            double total = 1.0;
            for (int k = 0; k < 3; k++) {
                double d = std::floor((hi[k] - lo[k]) / cell) + 1.0;
                dim[k] = (int64_t)d;
                total *= d;
            }
            if (total <= 16.0 * (double)n + 65536.0) break;
            cell *= 1.26;
        }
        const int64_t ncell = dim[0] * dim[1] * dim[2];
        start.assign((size_t)ncell + 1, 0);
        order.resize((size_t)n);
        std::vector<int64_t> cidx((size_t)n);
        for (int64_t i = 0; i < n; i++) {
            int64_t c[3];
            cell_of(i, c);
            cidx[(size_t)i] = (c[0] * dim[1] + c[1]) * dim[2] + c[2];
            start[(size_t)cidx[(size_t)i] + 1]++;
        }
        for (int64_t c = 0; c < ncell; c++) start[(size_t)c + 1] += start[(size_t)c];
        std::vector<int64_t> cur(start.begin(), start.end() - 1);
        for (int64_t i = 0; i < n; i++) order[(size_t)cur[(size_t)cidx[(size_t)i]]++] = i;  // ascending id inside a cell
    }
    void cell_of(int64_t i, int64_t c[3]) const {
        for (int k = 0; k < 3; k++) {
            int64_t t = (int64_t)std::floor((hx[3 * (size_t)i + k] - lo[k]) / cell);
            c[k] = std::min(std::max<int64_t>(t, 0), dim[k] - 1);
        }
    }
    void candidates(int64_t i, std::vector<int64_t>& cand) const {
        int64_t c[3];
        cell_of(i, c);
        cand.clear();
        for (int64_t dx = -1; dx <= 1; dx++)
            for (int64_t dy = -1; dy <= 1; dy++)
                for (int64_t dz = -1; dz <= 1; dz++) {
                    const int64_t cx = c[0] + dx, cy = c[1] + dy, cz = c[2] + dz;
                    if (cx < 0 || cy < 0 || cz < 0 || cx >= dim[0] || cy >= dim[1] || cz >= dim[2]) continue;
                    const int64_t cl = (cx * dim[1] + cy) * dim[2] + cz;
                    for (int64_t q = start[(size_t)cl]; q < start[(size_t)cl + 1]; q++)
                        if (order[(size_t)q] > i) cand.push_back(order[(size_t)q]);
                }
        std::sort(cand.begin(), cand.end());
    }
};

static void scan_bonds(int64_t n, const std::vector<double>& hx, double r, double crit, lg_unit_vector_fn gen, void* user,
                       std::vector<HostBond>& recs, int64_t& bonds) {
    auto X = [&](int64_t i) { return mk3(hx[3 * (size_t)i], hx[3 * (size_t)i + 1], hx[3 * (size_t)i + 2]); };
    // candidate partners j > i within 2r + crit, visited in lexicographic order so that the generator is consumed
    // exactly like the i < j double loop of sinter_bridge.h:148-149
    const HostGrid grid(n, hx, 2.0 * r + crit);
    std::vector<int64_t> cand;
    bonds = 0;
    for (int64_t i = 0; i < n - 1; i++) {
        grid.candidates(i, cand);
        const d3 part_i = X(i);
        for (int64_t j : cand) {
            const d3 part_j = X(j);
            if (!bridge_candidate(part_i, part_j, r, crit)) continue;  // sinter_bridge.h:155-157
            HostBond bi{}, bjr{};
            bi.i = i; bi.j = j; bjr.i = j; bjr.j = i;
            d3 tv;
            do {  // :171-175: draw again while the vector is collinear with the bond axis
                double t[3];
                gen(user, t);
                tv = mk3(t[0], t[1], t[2]);
            } while (!build_bridge(part_i, part_j, tv, r, bi.conn, bjr.conn));
            for (int k = 0; k < 5; k++) bi.len[k] = bjr.len[k] = (k < 4) ? 2.0 * r : 0.0;
            recs.push_back(bi);
            recs.push_back(bjr);
            bonds++;
        }
    }
}

// alt_sinter_functor's constructor (alt_sinter_bridge.h:56-89): bond when |dist - 2 r| < critical_separation (strict),
// i < j in lexicographic order; a bond that would close a cycle of bonds is skipped (:70-73 -- its subset relabelling,
// :83-87, is a "same connected component?" test, done here with a union-find).  Records carry no connectors.
static void scan_alt_bonds(int64_t n, const std::vector<double>& hx, double r, double crit, std::vector<HostBond>& recs,
                           int64_t& bonds) {
    auto X = [&](int64_t i) { return mk3(hx[3 * (size_t)i], hx[3 * (size_t)i + 1], hx[3 * (size_t)i + 2]); };
    const HostGrid grid(n, hx, 2.0 * r + crit);
    std::vector<int64_t> parent((size_t)n), cand;
    for (int64_t i = 0; i < n; i++) parent[(size_t)i] = i;
    auto find = [&](int64_t q) {
        int64_t root = q;
        while (parent[(size_t)root] != root) root = parent[(size_t)root];
        while (parent[(size_t)q] != root) {
            const int64_t nx = parent[(size_t)q];
            parent[(size_t)q] = root;
            q = nx;
        }
        return root;
    };
    bonds = 0;
    for (int64_t i = 0; i < n - 1; i++) {
        grid.candidates(i, cand);
        const d3 part_i = X(i);
        for (int64_t j : cand) {
            if (!(std::fabs(norm(part_i - X(j)) - 2.0 * r) < crit)) continue;  // :67
            const int64_t ri = find(i), rj = find(j);
            if (ri == rj) continue;  // "preventing neck insertion to avoid a cycle"
            parent[(size_t)rj] = ri;
            HostBond bi{}, bjr{};
            bi.i = i; bi.j = j; bjr.i = j; bjr.j = i;
            recs.push_back(bi);
            recs.push_back(bjr);
            bonds++;
        }
    }
}

void to_cp(const lg_contact_params* p, ContactParams& c) {
    c.k = p->k; c.gamma_n = p->gamma_n;
    c.k_t = p->k_t; c.gamma_t = p->gamma_t; c.mu_s = p->mu_s; c.phi_d = p->phi_d;
    c.k_r = p->k_r; c.gamma_r = p->gamma_r; c.mu_r = p->mu_r; c.phi_r = p->phi_r;
    c.k_o = p->k_o; c.gamma_o = p->gamma_o; c.mu_o = p->mu_o; c.phi_o = p->phi_o;
    c.dt = p->dt;
}

double* field_ptr(lg_system* s, int field) {
    switch (field) {
        case LG_X: return s->x.p;
        case LG_V: return s->v.p;
        case LG_A: return s->a.p;
        case LG_THETA: return s->th.p;
        case LG_OMEGA: return s->w.p;
        case LG_ALPHA: return s->al.p;
        default: return nullptr;
    }
}

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
        else prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};

}  // namespace

#define LG_ENTER(s)                                             \
    if ((s) == nullptr) return fail(LG_EINVAL, "null system"); \
    DeviceGuard guard_((s)->device)

extern "C" {

const char* lg_last_error(void) { return g_error.c_str(); }
int lg_abi_version(void) { return LGB200_ABI_VERSION; }

int lg_create(int64_t n, double r_part, double mass, double inertia, int device, lg_system** out) {
    if (out == nullptr) return fail(LG_EINVAL, "out is null");
    *out = nullptr;
    if (n <= 0 || n > 2000000000LL) return fail(LG_EINVAL, "n out of range");
    if (!(r_part > 0.0) || !(mass > 0.0) || !(inertia > 0.0)) return fail(LG_EINVAL, "r_part, mass, inertia must be positive");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count <= 0)
        return fail(LG_ENODEVICE, std::string("no CUDA device available (there is no CPU fallback): ") +
                                      (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
    if (device < 0) {
        LG_CUDA(cudaGetDevice(&device));
    } else if (device >= count) {
        return fail(LG_ENODEVICE, "device ordinal out of range");
    }
    DeviceGuard guard(device);  // declared first: still in force when an error return destroys the half-built system
    std::unique_ptr<lg_system> s(new (std::nothrow) lg_system());
    if (!s) return fail(LG_ENOMEM, "host allocation failed");
    s->device = device;
    s->n = n;
    s->mp.r = r_part;
    s->mp.mass = mass;
    s->mp.inertia = inertia;
    LG_CUDA(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    const size_t n3 = 3 * (size_t)n;
    LG_CUDA(s->x.reserve(n3));
    LG_CUDA(s->v.reserve(n3));
    LG_CUDA(s->a.reserve(n3));
    LG_CUDA(s->th.reserve(n3));
    LG_CUDA(s->w.reserve(n3));
    LG_CUDA(s->al.reserve(n3));
    LG_CUDA(s->row_ptr.reserve((size_t)n + 1));
    LG_CUDA(s->cols.reserve(16));
    LG_CUDA(s->tag.reserve(16));
    LG_CUDA(s->bptr.reserve((size_t)n + 1));
    LG_CUDA(s->bj.reserve(16));
    LG_CUDA(s->brev.reserve(16));
    LG_CUDA(s->bconn.reserve(16));
    LG_CUDA(s->blen.reserve(16));
    LG_CUDA(s->cell_of.reserve((size_t)n));
    LG_CUDA(s->order.reserve((size_t)n));
    LG_CUDA(s->bbox_scratch.reserve(6 * 1024));
    LG_CUDA(s->bbox_out.reserve(8));
    LG_CUDA(s->diag_partials.reserve(12 * 512));
    LG_CUDA(s->diag_out.reserve(16));
    LG_CUDA(s->counter.reserve(2));
    LG_CUDA(s->err.reserve(1));
    LG_CUDA(cudaMemsetAsync(s->err.p, 0, sizeof(int), s->stream));
    LG_CUDA(cudaMemsetAsync(s->row_ptr.p, 0, ((size_t)n + 1) * sizeof(uint32_t), s->stream));
    LG_CUDA(cudaMemsetAsync(s->bptr.p, 0, ((size_t)n + 1) * sizeof(uint32_t), s->stream));
    LG_CUDA(cudaMemsetAsync(s->a.p, 0, n3 * sizeof(double), s->stream));
    LG_CUDA(cudaMemsetAsync(s->al.p, 0, n3 * sizeof(double), s->stream));
    LG_CUDA(cudaMallocHost(&s->h_pinned, 32 * sizeof(double)));
    LG_CUDA(cudaStreamSynchronize(s->stream));
    s->launches0 = g_launch_count;
    *out = s.release();
    return LG_OK;
}

int lg_destroy(lg_system* s) {
    if (s == nullptr) return LG_OK;
    {
        DeviceGuard guard(s->device);
        delete s;  // ~lg_system releases the stream, pinned scratch and events; DevBuf destructors free device memory
    }
    return LG_OK;
}

// caller index -> internal index (identity unless the system was renumbered)
static inline int64_t to_internal(const lg_system* s, int64_t i) { return s->reordered ? (int64_t)s->h_inv[(size_t)i] : i; }
static inline int64_t to_caller(const lg_system* s, int64_t k) { return s->reordered ? (int64_t)s->h_perm[(size_t)k] : k; }

// Computes the Morton renumbering from the caller-order positions d_x (device) and keeps it on both sides.
static int compute_reorder(lg_system* s, const double* d_x) {
    const int n = (int)s->n;
    cudaStream_t st = s->stream;
    launch_bbox(n, d_x, s->bbox_scratch.p, s->bbox_out.p, st);
    double bb[6];
    LG_CUDA(cudaMemcpyAsync(bb, s->bbox_out.p, sizeof(bb), cudaMemcpyDeviceToHost, st));
    LG_CUDA(cudaStreamSynchronize(st));
    double ext = 0.0;
    for (int k = 0; k < 3; k++) {
        if (!std::isfinite(bb[k]) || !std::isfinite(bb[3 + k])) return fail(LG_EINVAL, "non-finite particle positions");
        ext = std::max(ext, bb[3 + k] - bb[k]);
    }
    const double cell = std::max(2.0 * s->mp.r, ext / 255.0);
    const size_t nc1 = morton_cells() + 1;
    DevBuf<uint32_t> key, count, cursor, scratch;
    DevBuf<int32_t> tmp;
    LG_CUDA(key.reserve((size_t)n));
    LG_CUDA(count.reserve(nc1));
    LG_CUDA(cursor.reserve(nc1));
    LG_CUDA(scratch.reserve(scan_scratch_elems(nc1)));
    LG_CUDA(tmp.reserve((size_t)n));
    LG_CUDA(s->d_perm.reserve((size_t)n));
    LG_CUDA(s->d_inv.reserve((size_t)n));
    launch_morton_order(n, d_x, bb, cell, key.p, count.p, cursor.p, scratch.p, tmp.p, s->d_perm.p, s->d_inv.p, st);
    LG_CUDA(cudaGetLastError());
    s->h_perm.resize((size_t)n);
    s->h_inv.resize((size_t)n);
    LG_CUDA(cudaMemcpyAsync(s->h_perm.data(), s->d_perm.p, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    LG_CUDA(cudaMemcpyAsync(s->h_inv.data(), s->d_inv.p, (size_t)n * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    LG_CUDA(cudaStreamSynchronize(st));
    s->reordered = true;
    return LG_OK;
}

int lg_set_reorder(lg_system* s, int mode) {
    LG_ENTER(s);
    if (mode != 0 && mode != 1) return fail(LG_EINVAL, "reorder mode must be 0 (caller order) or 1 (spatial)");
    if (s->state_set) return fail(LG_EINVAL, "lg_set_reorder must precede the first lg_set_state");
    s->reorder_mode = mode;
    return LG_OK;
}

int lg_set_state(lg_system* s, int64_t n_check, const double* x, const double* v, const double* theta, const double* omega) {
    LG_ENTER(s);
    if (n_check != s->n)
        return fail(LG_ESIZE, "size mismatch in arguments passed to lg_set_state (x0, v0, theta0, omega0 must hold n particles)");
    if (!x || !v || !theta || !omega) return fail(LG_EINVAL, "null state pointer");
    LG_TRY(sync_pairs(s));
    const size_t bytes = 3 * (size_t)s->n * sizeof(double);
    if (s->reorder_mode == 0) {
        LG_CUDA(cudaMemcpyAsync(s->x.p, x, bytes, cudaMemcpyHostToDevice, s->stream));
        LG_CUDA(cudaMemcpyAsync(s->v.p, v, bytes, cudaMemcpyHostToDevice, s->stream));
        LG_CUDA(cudaMemcpyAsync(s->th.p, theta, bytes, cudaMemcpyHostToDevice, s->stream));
        LG_CUDA(cudaMemcpyAsync(s->w.p, omega, bytes, cudaMemcpyHostToDevice, s->stream));
    } else {
        // caller order lands in buffers that hold nothing yet (the ping-pong alternates, a), then one gather per field
        LG_CUDA(cudaMemcpyAsync(s->x.next, x, bytes, cudaMemcpyHostToDevice, s->stream));
        LG_CUDA(cudaMemcpyAsync(s->v.next, v, bytes, cudaMemcpyHostToDevice, s->stream));
        LG_CUDA(cudaMemcpyAsync(s->a.p, theta, bytes, cudaMemcpyHostToDevice, s->stream));
        LG_CUDA(cudaMemcpyAsync(s->w.next, omega, bytes, cudaMemcpyHostToDevice, s->stream));
        if (!s->reordered) LG_TRY(compute_reorder(s, s->x.next));  // fixed by the first state: bond tables refer to it
        launch_gather3((int)s->n, s->d_perm.p, s->x.next, s->x.p, s->stream);
        launch_gather3((int)s->n, s->d_perm.p, s->v.next, s->v.p, s->stream);
        launch_gather3((int)s->n, s->d_perm.p, s->a.p, s->th.p, s->stream);
        launch_gather3((int)s->n, s->d_perm.p, s->w.next, s->w.p, s->stream);
        LG_CUDA(cudaGetLastError());
    }
    LG_CUDA(cudaMemsetAsync(s->a.p, 0, bytes, s->stream));
    LG_CUDA(cudaMemsetAsync(s->al.p, 0, bytes, s->stream));
    LG_CUDA(cudaStreamSynchronize(s->stream));  // the caller may reuse its buffers
    s->state_set = true;
    s->kicked = false;
    s->half_kicked = false;
    s->pending = false;
    s->rows_staged = false;
    s->grid_valid = false;  // new positions: the next rebuild measures the bounding box again
    // A new state is a new trajectory, i.e. what the reference gets from constructing a new system: no history springs
    // (contact_force.h:53 zero-initialised), an empty neighbor list until update_neighbor_list is called
    // (rotational_binary_system_neighbors_omp.h:66), step count 0 (lg_run's rebuild period restarts).  Bond tables stay.
    s->steps = 0;
    if (s->slots_ready) LG_CUDA(cudaMemsetAsync(s->sgid.p, 0xff, (size_t)s->n * s->S * sizeof(int32_t), s->stream));
    if (s->mp.alt_bonds && s->n_bonds > 0) LG_CUDA(cudaMemsetAsync(s->bspr.p, 0, (size_t)s->n_bonds * 9 * sizeof(double), s->stream));
    for (size_t f = 0; f < s->facets.size(); f++) {
        LG_CUDA(cudaMemsetAsync(s->wspr[f]->p, 0, (size_t)s->n * 9 * sizeof(double), s->stream));
        LG_CUDA(cudaMemsetAsync(s->wact[f]->p, 0, (size_t)s->n, s->stream));
    }
    if (s->use_nl) {
        s->n_pairs = 0;
        s->rebuilds = 0;
        s->last_total = -1;
        LG_CUDA(cudaMemsetAsync(s->row_ptr.p, 0, ((size_t)s->n + 1) * sizeof(uint32_t), s->stream));
    } else if (s->allpairs_built) {
        LG_TRY(retag(s));
    }
    return LG_OK;
}

int lg_get_field(lg_system* s, int field, double* out) {
    LG_ENTER(s);
    if (field == LG_A || field == LG_ALPHA) LG_TRY(flush(s));
    double* src = field_ptr(s, field);
    if (!src || !out) return fail(LG_EINVAL, "bad field id or null output");
    if (s->reordered) {  // back to the caller's particle order
        LG_CUDA(s->io_scratch.reserve(3 * (size_t)s->n));
        launch_gather3((int)s->n, s->d_inv.p, src, s->io_scratch.p, s->stream);
        LG_CUDA(cudaGetLastError());
        src = s->io_scratch.p;
    }
    LG_CUDA(cudaMemcpyAsync(out, src, 3 * (size_t)s->n * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    return check_async(s);
}

int lg_device_field(lg_system* s, int field, double** d_ptr) {
    LG_ENTER(s);
    if (field == LG_A || field == LG_ALPHA) LG_TRY(flush(s));
    double* src = field_ptr(s, field);
    if (!src || !d_ptr) return fail(LG_EINVAL, "bad field id or null output");
    if (s->reordered) return fail(LG_EINVAL, "device views are in the internal particle order: not available after lg_set_reorder");
    *d_ptr = src;
    return LG_OK;
}

int lg_set_contact(lg_system* s, const lg_contact_params* p) {
    LG_ENTER(s);
    if (!p) return fail(LG_EINVAL, "null params");
    if (s->mask & MODEL_SINTER) return fail(LG_EINVAL, "sinter model already installed (it embeds the contact model)");
    to_cp(p, s->mp.cp);
    s->mask |= MODEL_CONTACT;
    return LG_OK;
}

int lg_set_hamaker(lg_system* s, double A, double h0) {
    LG_ENTER(s);
    s->mp.hamaker_A = A;
    s->mp.hamaker_h0 = h0;
    s->mask |= MODEL_HAMAKER;
    return LG_OK;
}

/* ---- SURVEY.md section 8f variants: all of them run the reference's schedule literally (kernel variant 0) -------------- */
static int lock_simple(lg_system* s) {
    LG_TRY(flush(s));
    s->variant = 0;
    s->math_mode = 0;
    s->simple_only = true;
    return LG_OK;
}

int lg_set_hamaker_variant(lg_system* s, int kind, double r_cutoff) {
    LG_ENTER(s);
    if (kind < HAMAKER_STANDARD || kind > HAMAKER_SIMPLIFIED) return fail(LG_EINVAL, "unknown Hamaker variant");
    if (kind == HAMAKER_LIMITED_RANGE && !(r_cutoff > 0.0)) return fail(LG_EINVAL, "limited-range Hamaker needs r_cutoff > 0");
    LG_TRY(flush(s));
    s->mp.hamaker_kind = kind;
    s->mp.hamaker_cutoff = r_cutoff;
    return LG_OK;
}

int lg_set_integrator(lg_system* s, int integrator) {
    LG_ENTER(s);
    if (integrator != LG_INTEGRATOR_VERLET_HALF && integrator != LG_INTEGRATOR_FORWARD_EULER) return fail(LG_EINVAL, "unknown integrator");
    if (s->steps > 0 && integrator != s->integrator) return fail(LG_EINVAL, "the integrator is a constructor choice: set it before the first step");
    s->integrator = integrator;
    return integrator == LG_INTEGRATOR_FORWARD_EULER ? lock_simple(s) : LG_OK;
}

int lg_set_fixed_particle(lg_system* s, int64_t index) {
    LG_ENTER(s);
    if (index < -1 || index >= s->n) return fail(LG_EINVAL, "fixed particle index out of range");
    s->fixed_particle = index >= 0 ? (int)to_internal(s, index) : -1;
    return index >= 0 ? lock_simple(s) : LG_OK;
}

int lg_set_central_attraction(lg_system* s, double k_a, const double x0[3]) {
    LG_ENTER(s);
    if (!x0) return fail(LG_EINVAL, "null centre");
    if (!s->facets.empty()) return fail(LG_EINVAL, "central attraction together with facets is not supported");
    LG_TRY(lock_simple(s));
    s->central_on = true;
    s->central_k = k_a;
    for (int k = 0; k < 3; k++) s->central_x0[k] = x0[k];
    return LG_OK;
}

int lg_alt_bond_table_create(int64_t n, const double* x, double r_part, double crit, lg_bond_table** out) {
    if (!out) return fail(LG_EINVAL, "out is null");
    *out = nullptr;
    if (n <= 0 || !x || !(r_part > 0.0)) return fail(LG_EINVAL, "bad arguments to lg_alt_bond_table_create");
    std::unique_ptr<lg_bond_table> t(new (std::nothrow) lg_bond_table());
    if (!t) return fail(LG_ENOMEM, "host allocation failed");
    t->n = n;
    t->alt = true;
    std::vector<double> hx(x, x + 3 * (size_t)n);
    scan_alt_bonds(n, hx, r_part, crit, t->recs, t->bonds);
    *out = t.release();
    return LG_OK;
}

int lg_set_alt_bond_table(lg_system* s, const lg_contact_params* contact, const double bond_params[9], const lg_bond_table* t) {
    LG_ENTER(s);
    if (!contact || !bond_params || !t) return fail(LG_EINVAL, "null contact params, bond params or bond table");
    if (!t->alt) return fail(LG_EINVAL, "the table was built for sinter_functor (lg_bond_table_create)");
    if (t->n != s->n) return fail(LG_ESIZE, "bond table was built for a different particle count");
    if (s->mask & (MODEL_CONTACT | MODEL_SINTER)) return fail(LG_EINVAL, "a contact or sinter model is already installed; alt_sinter embeds contact");
    LG_TRY(lock_simple(s));
    std::vector<HostBond> recs = t->recs;
    to_cp(contact, s->mp.cp);
    for (int k = 0; k < 8; k++) s->mp.alt[k] = bond_params[k];
    s->mp.alt_dt = bond_params[8];
    s->mp.alt_bonds = 1;
    s->mp.k_d = 0.0;
    s->mp.gamma_d = 0.0;
    s->mask |= MODEL_SINTER;  // same row tags, same bond CSR; the bonded law is chosen by mp.alt_bonds
    LG_CUDA(s->bspr.reserve(recs.size() * 9 + 16));
    LG_CUDA(cudaMemsetAsync(s->bspr.p, 0, (recs.size() * 9 + 16) * sizeof(double), s->stream));
    return install_bonds(s, recs);
}

int lg_get_alt_bond(lg_system* s, int64_t i, int64_t j, int* bonded, double springs[9]) {
    LG_ENTER(s);
    if (!bonded || !springs || i < 0 || j < 0 || i >= s->n || j >= s->n) return fail(LG_EINVAL, "bad arguments");
    *bonded = 0;
    if (s->n_bonds == 0 || !s->mp.alt_bonds) return LG_OK;
    LG_TRY(flush(s));
    i = to_internal(s, i);
    j = to_internal(s, j);
    for (uint32_t b = s->h_bptr[(size_t)i]; b < s->h_bptr[(size_t)i + 1]; b++)
        if (s->h_bj[b] == (int32_t)j) {
            LG_CUDA(cudaMemcpyAsync(springs, s->bspr.p + (size_t)b * 9, 9 * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
            LG_CUDA(cudaStreamSynchronize(s->stream));
            *bonded = 1;
            break;
        }
    return LG_OK;
}

int lg_set_autotune(lg_system* s, int enabled) {
    LG_ENTER(s);
    s->autotune = enabled ? 1 : 0;
    if (!enabled) s->tuned_shape = -1;
    return LG_OK;
}

int lg_set_rows_shape(lg_system* s, int shape) {
    LG_ENTER(s);
    if (shape < -1 || shape > 3) return fail(LG_EINVAL, "shape must be -1 (rule / autotune), 0, 1, 2 or 3");
    s->forced_shape = shape;
    return LG_OK;
}

int lg_rows_shape_rule(double mean_row_length, int64_t n, int has_bonds) {
    return rows2_shape_for(mean_row_length, (int)std::min<int64_t>(n, 2000000000LL), has_bonds != 0);
}

int lg_get_rows_shape(lg_system* s, int* shape) {
    LG_ENTER(s);
    if (!shape) return fail(LG_EINVAL, "null output");
    if (s->forced_shape >= 0) {
        *shape = s->forced_shape;
        return LG_OK;
    }
    const double k_mean = s->n > 0 ? (double)s->n_pairs / (double)s->n : 0.0;
    *shape = (s->autotune && s->tuned_shape >= 0) ? s->tuned_shape : rows2_shape_for(k_mean, (int)s->n, s->n_bonds > 0);
    return LG_OK;
}

int lg_set_max_contacts(lg_system* s, int max_contacts) {
    LG_ENTER(s);
    if (s->slots_ready) return fail(LG_EINVAL, "lg_set_max_contacts must precede the first step / rebuild");
    if (max_contacts < 1 || max_contacts > kMaxSlots) return fail(LG_EINVAL, "max_contacts must be in 1..120");
    s->S = max_contacts;
    return LG_OK;
}

int lg_set_bonds(lg_system* s, const lg_contact_params* contact, double k_d, double gamma_d, int64_t n_ordered,
                 const int64_t* ij, const double* conn, const double* len) {
    LG_ENTER(s);
    if (!contact || n_ordered < 0 || (n_ordered > 0 && (!ij || !conn || !len))) return fail(LG_EINVAL, "bad bond arguments");
    if (s->mask & MODEL_CONTACT) return fail(LG_EINVAL, "contact model already installed; the sinter model embeds it");
    std::vector<HostBond> recs((size_t)n_ordered);
    for (int64_t b = 0; b < n_ordered; b++) {
        HostBond& r = recs[(size_t)b];
        r.i = ij[2 * b];
        r.j = ij[2 * b + 1];
        if (r.i < 0 || r.j < 0 || r.i >= s->n || r.j >= s->n || r.i == r.j) return fail(LG_EINVAL, "bond index out of range");
        std::memcpy(r.conn, conn + 15 * b, sizeof(r.conn));
        std::memcpy(r.len, len + 5 * b, sizeof(r.len));
    }
    to_cp(contact, s->mp.cp);
    s->mp.k_d = k_d;
    s->mp.gamma_d = gamma_d;
    s->mask |= MODEL_SINTER;
    return install_bonds(s, recs);
}

int lg_build_sinter_bonds(lg_system* s, const lg_contact_params* contact, double k_d, double gamma_d, double crit,
                          lg_unit_vector_fn gen, void* user, int64_t* n_bonds_out) {
    LG_ENTER(s);
    if (!contact || !gen) return fail(LG_EINVAL, "null contact params or generator");
    if (!s->state_set) return fail(LG_EINVAL, "lg_set_state must precede lg_build_sinter_bonds");
    if (s->mask & MODEL_CONTACT) return fail(LG_EINVAL, "contact model already installed; the sinter model embeds it");
    const int64_t n = s->n;
    std::vector<double> hx(3 * (size_t)n);
    const double* d_x = s->x.p;
    if (s->reordered) {  // the scan consumes `gen` in the CALLER's lexicographic pair order
        LG_CUDA(s->io_scratch.reserve(3 * (size_t)n));
        launch_gather3((int)n, s->d_inv.p, s->x.p, s->io_scratch.p, s->stream);
        d_x = s->io_scratch.p;
    }
    LG_CUDA(cudaMemcpyAsync(hx.data(), d_x, hx.size() * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    LG_CUDA(cudaStreamSynchronize(s->stream));
    std::vector<HostBond> recs;
    int64_t bonds = 0;
    scan_bonds(n, hx, s->mp.r, crit, gen, user, recs, bonds);
    to_cp(contact, s->mp.cp);
    s->mp.k_d = k_d;
    s->mp.gamma_d = gamma_d;
    s->mask |= MODEL_SINTER;
    if (n_bonds_out) *n_bonds_out = bonds;
    return install_bonds(s, recs);
}

int lg_bond_table_create(int64_t n, const double* x, double r_part, double crit, lg_unit_vector_fn gen, void* user,
                         lg_bond_table** out) {
    if (!out) return fail(LG_EINVAL, "out is null");
    *out = nullptr;
    if (n <= 0 || !x || !gen || !(r_part > 0.0)) return fail(LG_EINVAL, "bad arguments to lg_bond_table_create");
    std::unique_ptr<lg_bond_table> t(new (std::nothrow) lg_bond_table());
    if (!t) return fail(LG_ENOMEM, "host allocation failed");
    t->n = n;
    std::vector<double> hx(x, x + 3 * (size_t)n);
    scan_bonds(n, hx, r_part, crit, gen, user, t->recs, t->bonds);
    *out = t.release();
    return LG_OK;
}

// The same table built on the GPU (kernels_bonds.cu): candidates from the cell list, bond k in lexicographic pair order takes
// unit vector k.  Falls back to the sequential host scan when a vector is collinear with its bond axis (the reference would
// redraw and shift every later draw) -- the result is the reference's in either case.
int lg_bond_table_create_device(int64_t n, const double* x, double r_part, double crit, int64_t n_uv, const double* uv, int device,
                                lg_bond_table** out) {
    if (!out) return fail(LG_EINVAL, "out is null");
    *out = nullptr;
    if (n <= 0 || n > 2000000000LL || !x || !uv || n_uv < 0 || !(r_part > 0.0)) return fail(LG_EINVAL, "bad arguments to lg_bond_table_create_device");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0) return fail(LG_ENODEVICE, "no CUDA device available (use lg_bond_table_create)");
    if (device < 0) LG_CUDA(cudaGetDevice(&device));
    if (device >= count) return fail(LG_ENODEVICE, "device ordinal out of range");
    DeviceGuard guard(device);
    cudaStream_t st = nullptr;
    LG_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    struct StreamGuard { cudaStream_t s; ~StreamGuard() { cudaStreamDestroy(s); } } sg{st};
    const int ni = (int)n;
    const size_t n3 = 3 * (size_t)n;
    DevBuf<double> dx, duv, bbs, bbo, dconn, dlen;
    DevBuf<int32_t> cell_of, order, cols, dbj, dbrev;
    DevBuf<uint32_t> cell_start, cell_cursor, scratch, row_ptr, deg, up;
    DevBuf<int> flags;
    LG_CUDA(dx.reserve(n3));
    LG_CUDA(duv.reserve(3 * (size_t)std::max<int64_t>(n_uv, 1)));
    LG_CUDA(bbs.reserve(6 * 1024));
    LG_CUDA(bbo.reserve(8));
    LG_CUDA(flags.reserve(1));
    LG_CUDA(cudaMemcpyAsync(dx.p, x, n3 * sizeof(double), cudaMemcpyHostToDevice, st));
    if (n_uv) LG_CUDA(cudaMemcpyAsync(duv.p, uv, 3 * (size_t)n_uv * sizeof(double), cudaMemcpyHostToDevice, st));
    LG_CUDA(cudaMemsetAsync(flags.p, 0, sizeof(int), st));
    // candidate rows: every pair closer than a hair above 2 r + crit (the exact predicate is applied afterwards)
    const double range = (2.0 * r_part + crit) * (1.0 + 1e-9);
    launch_bbox(ni, dx.p, bbs.p, bbo.p, st);
    double bb[6];
    LG_CUDA(cudaMemcpyAsync(bb, bbo.p, sizeof(bb), cudaMemcpyDeviceToHost, st));
    LG_CUDA(cudaStreamSynchronize(st));
    NeighborGrid g{};
    double cell = range * (1.0 + 1e-6);
    for (;;) {
        double total = 1.0;
        for (int k = 0; k < 3; k++) {
            const double ext = bb[3 + k] - bb[k];
            if (!(ext >= 0.0) || !std::isfinite(ext)) return fail(LG_EINVAL, "non-finite particle positions");
            const double d = std::floor(ext / cell) + 1.0;
            g.dim[k] = d > 2.0e9 ? 2000000000 : (int)d;
            total *= d;
        }
        if (total <= 16.0 * (double)n + 65536.0) break;
        cell *= 1.26;
    }
    g.n_cells = (int64_t)g.dim[0] * g.dim[1] * g.dim[2];
    g.cell = cell;
    for (int k = 0; k < 3; k++) g.lo[k] = bb[k];
    const size_t nc1 = (size_t)g.n_cells + 1;
    LG_CUDA(cell_of.reserve((size_t)n));
    LG_CUDA(order.reserve((size_t)n));
    LG_CUDA(cell_start.reserve(nc1));
    LG_CUDA(cell_cursor.reserve(nc1));
    LG_CUDA(scratch.reserve(scan_scratch_elems(std::max(nc1, (size_t)n + 1))));
    LG_CUDA(row_ptr.reserve((size_t)n + 1));
    LG_CUDA(deg.reserve((size_t)n + 1));
    LG_CUDA(up.reserve((size_t)n + 1));
    LG_CUDA(cudaMemsetAsync(cell_start.p, 0, nc1 * sizeof(uint32_t), st));
    LG_CUDA(cudaMemsetAsync(cell_cursor.p, 0, nc1 * sizeof(uint32_t), st));
    launch_cell_count(ni, dx.p, g, cell_of.p, cell_start.p, st);
    launch_exclusive_scan_u32(cell_start.p, nc1, scratch.p, st);
    launch_cell_fill(ni, cell_of.p, cell_start.p, cell_cursor.p, order.p, st);
    launch_row_count(ni, dx.p, g, cell_start.p, order.p, range, row_ptr.p, nullptr, nullptr, st);
    LG_CUDA(cudaMemsetAsync(row_ptr.p + n, 0, sizeof(uint32_t), st));
    launch_exclusive_scan_u32(row_ptr.p, (size_t)n + 1, scratch.p, st);
    uint32_t n_cand = 0;
    LG_CUDA(cudaMemcpyAsync(&n_cand, row_ptr.p + n, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    LG_CUDA(cudaStreamSynchronize(st));
    LG_CUDA(cols.reserve((size_t)n_cand + 16));
    launch_row_fill(ni, dx.p, g, cell_start.p, order.p, range, row_ptr.p, cols.p, nullptr, nullptr, st);
    // bonds per particle, lexicographic ranks
    launch_bond_count(ni, dx.p, row_ptr.p, cols.p, r_part, crit, deg.p, up.p, st);
    LG_CUDA(cudaMemsetAsync(deg.p + n, 0, sizeof(uint32_t), st));
    LG_CUDA(cudaMemsetAsync(up.p + n, 0, sizeof(uint32_t), st));
    launch_exclusive_scan_u32(deg.p, (size_t)n + 1, scratch.p, st);
    launch_exclusive_scan_u32(up.p, (size_t)n + 1, scratch.p, st);
    uint32_t totals[2] = {0, 0};
    LG_CUDA(cudaMemcpyAsync(&totals[0], deg.p + n, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    LG_CUDA(cudaMemcpyAsync(&totals[1], up.p + n, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    LG_CUDA(cudaStreamSynchronize(st));
    const size_t nb = totals[0];
    if ((int64_t)totals[1] > n_uv) return fail(LG_EINVAL, "fewer unit vectors than bonds (" + std::to_string(totals[1]) + " bonds)");
    LG_CUDA(dbj.reserve(nb + 16));
    LG_CUDA(dbrev.reserve(nb + 16));
    LG_CUDA(dconn.reserve(nb * 15 + 16));
    LG_CUDA(dlen.reserve(nb * 5 + 16));
    launch_bond_fill(ni, dx.p, row_ptr.p, cols.p, r_part, crit, deg.p, up.p, duv.p, n_uv, dbj.p, dbrev.p, dconn.p, dlen.p, flags.p, st);
    LG_CUDA(cudaGetLastError());
    int h_flags = 0;
    std::vector<uint32_t> bptr((size_t)n + 1);
    std::vector<int32_t> bj(nb);
    std::vector<double> conn(nb * 15), len(nb * 5);
    LG_CUDA(cudaMemcpyAsync(&h_flags, flags.p, sizeof(int), cudaMemcpyDeviceToHost, st));
    LG_CUDA(cudaMemcpyAsync(bptr.data(), deg.p, ((size_t)n + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    if (nb) {
        LG_CUDA(cudaMemcpyAsync(bj.data(), dbj.p, nb * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        LG_CUDA(cudaMemcpyAsync(conn.data(), dconn.p, nb * 15 * sizeof(double), cudaMemcpyDeviceToHost, st));
        LG_CUDA(cudaMemcpyAsync(len.data(), dlen.p, nb * 5 * sizeof(double), cudaMemcpyDeviceToHost, st));
    }
    LG_CUDA(cudaStreamSynchronize(st));
    if (h_flags & 2) {  // a collinear draw: only the sequential scan knows how the reference's generator moves on
        struct Seq { const double* uv; int64_t n, k; bool short_; } seq{uv, n_uv, 0, false};
        auto draw = [](void* user, double o[3]) {
            Seq* q = static_cast<Seq*>(user);
            if (q->k >= q->n) { q->short_ = true; o[0] = 1.0; o[1] = o[2] = 0.0; return; }
            for (int c = 0; c < 3; c++) o[c] = q->uv[3 * q->k + c];
            q->k++;
        };
        LG_TRY(lg_bond_table_create(n, x, r_part, crit, draw, &seq, out));
        if (seq.short_) {
            lg_bond_table_destroy(*out);
            *out = nullptr;
            return fail(LG_EINVAL, "fewer unit vectors than draws (a collinear vector was drawn again)");
        }
        return LG_OK;
    }
    std::unique_ptr<lg_bond_table> t(new (std::nothrow) lg_bond_table());
    if (!t) return fail(LG_ENOMEM, "host allocation failed");
    t->n = n;
    t->bonds = (int64_t)totals[1];
    t->recs.resize(nb);
    for (int64_t i = 0; i < n; i++)
        for (uint32_t b = bptr[(size_t)i]; b < bptr[(size_t)i + 1]; b++) {
            HostBond& r = t->recs[b];
            r.i = i;
            r.j = bj[b];
            std::memcpy(r.conn, &conn[(size_t)b * 15], sizeof(r.conn));
            std::memcpy(r.len, &len[(size_t)b * 5], sizeof(r.len));
        }
    *out = t.release();
    return LG_OK;
}

int lg_bond_table_count(const lg_bond_table* t, int64_t* n_bonds) {
    if (!t || !n_bonds) return fail(LG_EINVAL, "null bond table");
    *n_bonds = t->bonds;
    return LG_OK;
}

int lg_bond_table_records(const lg_bond_table* t, int64_t* ij, double* conn, double* len) {
    if (!t || !ij || !conn || !len) return fail(LG_EINVAL, "null bond table or output");
    for (size_t b = 0; b < t->recs.size(); b++) {
        ij[2 * b] = t->recs[b].i;
        ij[2 * b + 1] = t->recs[b].j;
        std::memcpy(conn + 15 * b, t->recs[b].conn, sizeof(t->recs[b].conn));
        std::memcpy(len + 5 * b, t->recs[b].len, sizeof(t->recs[b].len));
    }
    return LG_OK;
}

int lg_bond_table_destroy(lg_bond_table* t) {
    delete t;
    return LG_OK;
}

int lg_set_bond_table(lg_system* s, const lg_contact_params* contact, double k_d, double gamma_d, const lg_bond_table* t) {
    LG_ENTER(s);
    if (!contact || !t) return fail(LG_EINVAL, "null contact params or bond table");
    if (t->n != s->n) return fail(LG_ESIZE, "bond table was built for a different particle count");
    if (s->mask & MODEL_CONTACT) return fail(LG_EINVAL, "contact model already installed; the sinter model embeds it");
    std::vector<HostBond> recs = t->recs;
    to_cp(contact, s->mp.cp);
    s->mp.k_d = k_d;
    s->mp.gamma_d = gamma_d;
    s->mask |= MODEL_SINTER;
    return install_bonds(s, recs);
}

int lg_add_facet(lg_system* s, const double vertices[9], const double v_facet[3], const lg_contact_params* sc, int has_hamaker,
                 double A, double h0, int* facet_id) {
    LG_ENTER(s);
    if (!vertices || !v_facet) return fail(LG_EINVAL, "null facet data");
    if (!sc && !has_hamaker) return fail(LG_EINVAL, "at least one surface force functor must be provided");
    if ((int)s->facets.size() >= kMaxFacets) return fail(LG_ECAPACITY, "too many facets (max 8)");
    Facet f{};
    for (int k = 0; k < 3; k++) f.vtx[k] = mk3(vertices[3 * k], vertices[3 * k + 1], vertices[3 * k + 2]);
    f.v = mk3(v_facet[0], v_facet[1], v_facet[2]);
    facet_init_geometry(f);
    f.has_contact = sc != nullptr;
    f.has_hamaker = has_hamaker != 0;
    if (sc) to_cp(sc, f.sp);
    f.A = A;
    f.h0 = h0;
    auto ws = std::make_unique<DevBuf<double>>();
    auto wa = std::make_unique<DevBuf<uint8_t>>();
    LG_CUDA(ws->reserve((size_t)s->n * 9));
    LG_CUDA(wa->reserve((size_t)s->n));
    LG_CUDA(cudaMemsetAsync(ws->p, 0, (size_t)s->n * 9 * sizeof(double), s->stream));
    LG_CUDA(cudaMemsetAsync(wa->p, 0, (size_t)s->n, s->stream));
    s->facets.push_back(f);
    s->wspr.push_back(std::move(ws));
    s->wact.push_back(std::move(wa));
    if (facet_id) *facet_id = (int)s->facets.size() - 1;
    return LG_OK;
}

int lg_facet_set_velocity(lg_system* s, int id, const double v[3]) {
    LG_ENTER(s);
    if (id < 0 || id >= (int)s->facets.size() || !v) return fail(LG_EINVAL, "bad facet id");
    s->facets[(size_t)id].v = mk3(v[0], v[1], v[2]);
    return LG_OK;
}

int lg_facet_set_vertices(lg_system* s, int id, const double vertices[9]) {
    LG_ENTER(s);
    if (id < 0 || id >= (int)s->facets.size() || !vertices) return fail(LG_EINVAL, "bad facet id");
    Facet& f = s->facets[(size_t)id];
    for (int k = 0; k < 3; k++) f.vtx[k] = mk3(vertices[3 * k], vertices[3 * k + 1], vertices[3 * k + 2]);
    return LG_OK;
}

int lg_facet_update_positions(lg_system* s, int id, double dt) {
    LG_ENTER(s);
    if (id < 0 || id >= (int)s->facets.size()) return fail(LG_EINVAL, "bad facet id");
    Facet& f = s->facets[(size_t)id];
    for (int k = 0; k < 3; k++) f.vtx[k] = f.vtx[k] + f.v * dt;  // triangular_facet.h:62-66
    return LG_OK;
}

int lg_facet_get_vertices(lg_system* s, int id, double out[9]) {
    LG_ENTER(s);
    if (id < 0 || id >= (int)s->facets.size() || !out) return fail(LG_EINVAL, "bad facet id");
    for (int k = 0; k < 3; k++) store3(out + 3 * k, s->facets[(size_t)id].vtx[k]);
    return LG_OK;
}

int lg_enable_neighbor_list(lg_system* s, double r_verlet) {
    LG_ENTER(s);
    if (!(r_verlet > 0.0)) return fail(LG_EINVAL, "r_verlet must be positive");
    // Contact springs are dropped when a pair leaves the list, which equals the reference's reset on a visited
    // non-overlapping evaluation only if every overlapping pair (distance < 2 r) is inside the list.
    if (r_verlet < 2.0 * s->mp.r) return fail(LG_EINVAL, "r_verlet must be at least 2 * r_part (contact range)");
    LG_TRY(sync_pairs(s));
    s->last_total = -1;
    s->use_nl = true;
    s->r_verlet = r_verlet;
    s->allpairs_built = false;
    s->n_pairs = 0;
    s->rows_staged = false;
    s->grid_valid = false;
    LG_CUDA(cudaMemsetAsync(s->row_ptr.p, 0, ((size_t)s->n + 1) * sizeof(uint32_t), s->stream));
    return LG_OK;
}

int lg_update_neighbor_list(lg_system* s) {
    LG_ENTER(s);
    return rebuild(s);
}

int lg_do_step(lg_system* s, double dt) {
    LG_ENTER(s);
    LG_TRY(prepare(s));
    return step(s, dt);
}

int lg_run(lg_system* s, double dt, int64_t n_steps, int64_t rebuild_every, int move_facets) {
    LG_ENTER(s);
    LG_TRY(prepare(s));
    for (int64_t k = 0; k < n_steps; k++) {
        if (s->use_nl && rebuild_every > 0 && s->steps % rebuild_every == 0) LG_TRY(rebuild(s));
        LG_TRY(step(s, dt));
        if (move_facets)
            for (auto& f : s->facets)
                for (int q = 0; q < 3; q++) f.vtx[q] = f.vtx[q] + f.v * dt;
    }
    return LG_OK;
}

int lg_reset_acceleration_buffers(lg_system* s) {
    LG_ENTER(s);
    LG_TRY(flush(s));
    const size_t bytes = 3 * (size_t)s->n * sizeof(double);
    LG_CUDA(cudaMemsetAsync(s->a.p, 0, bytes, s->stream));
    LG_CUDA(cudaMemsetAsync(s->al.p, 0, bytes, s->stream));
    return LG_OK;
}

int lg_synchronize(lg_system* s) {
    LG_ENTER(s);
    return check_async(s);
}

int lg_get_neighbor_count(lg_system* s, int64_t* n_pairs) {
    LG_ENTER(s);
    if (!n_pairs) return fail(LG_EINVAL, "null output");
    if (!s->use_nl && !s->allpairs_built) {
        *n_pairs = s->n * (s->n - 1);
        return LG_OK;
    }
    LG_TRY(flush(s));  // rows rebuilt behind a pending evaluation are installed by it
    LG_TRY(sync_pairs(s));
    *n_pairs = s->n_pairs;
    return LG_OK;
}

int lg_get_neighbor_rows(lg_system* s, int64_t* row_ptr, int64_t* cols) {
    LG_ENTER(s);
    if (!row_ptr || !cols) return fail(LG_EINVAL, "null output");
    if (!s->use_nl && !s->allpairs_built) {
        LG_TRY(ensure_slots(s));
        LG_TRY(build_allpairs(s));
    }
    LG_TRY(flush(s));  // rows rebuilt behind a pending evaluation are installed by it
    LG_TRY(sync_pairs(s));
    const size_t n = (size_t)s->n, p = (size_t)s->n_pairs;
    std::vector<uint32_t> hp(n + 1);
    std::vector<int32_t> hc(p);
    LG_CUDA(cudaMemcpyAsync(hp.data(), s->row_ptr.p, (n + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream));
    if (p) LG_CUDA(cudaMemcpyAsync(hc.data(), s->cols.p, p * sizeof(int32_t), cudaMemcpyDeviceToHost, s->stream));
    LG_CUDA(cudaStreamSynchronize(s->stream));
    if (!s->reordered) {
        for (size_t i = 0; i <= n; i++) row_ptr[i] = (int64_t)hp[i];
        for (size_t k = 0; k < p; k++) cols[k] = (int64_t)hc[k];
        return LG_OK;
    }
    // renumbered system: rows in the caller's order, columns in the caller's ids, ascending
    row_ptr[0] = 0;
    for (size_t i = 0; i < n; i++) {
        const size_t k = (size_t)s->h_inv[i];
        const size_t len = hp[k + 1] - hp[k];
        int64_t* dst = cols + row_ptr[i];
        for (size_t q = 0; q < len; q++) dst[q] = (int64_t)s->h_perm[(size_t)hc[hp[k] + q]];
        std::sort(dst, dst + len);
        row_ptr[i + 1] = row_ptr[i] + (int64_t)len;
    }
    return LG_OK;
}

int lg_get_contact_springs(lg_system* s, int64_t i, int64_t j, double out[9]) {
    LG_ENTER(s);
    if (!out || i < 0 || j < 0 || i >= s->n || j >= s->n) return fail(LG_EINVAL, "bad pair");
    std::fill(out, out + 9, 0.0);
    if (!s->slots_ready) return LG_OK;
    LG_TRY(flush(s));
    i = to_internal(s, i);
    j = to_internal(s, j);
    std::vector<int32_t> ids((size_t)s->S);
    LG_CUDA(cudaMemcpyAsync(ids.data(), s->sgid.p + (size_t)i * s->S, (size_t)s->S * sizeof(int32_t), cudaMemcpyDeviceToHost,
                            s->stream));
    LG_CUDA(cudaStreamSynchronize(s->stream));
    for (int q = 0; q < s->S; q++)
        if (ids[(size_t)q] == (int32_t)j) {
            LG_CUDA(cudaMemcpyAsync(out, s->spr.p + ((size_t)i * s->S + q) * 9, 9 * sizeof(double), cudaMemcpyDeviceToHost,
                                    s->stream));
            LG_CUDA(cudaStreamSynchronize(s->stream));
            break;
        }
    return LG_OK;
}

int lg_get_bond(lg_system* s, int64_t i, int64_t j, int* bonded, double conn[15], double len[5]) {
    LG_ENTER(s);
    if (!bonded || !conn || !len || i < 0 || j < 0 || i >= s->n || j >= s->n) return fail(LG_EINVAL, "bad arguments");
    *bonded = 0;
    if (s->n_bonds == 0) return LG_OK;
    i = to_internal(s, i);
    j = to_internal(s, j);
    for (uint32_t b = s->h_bptr[(size_t)i]; b < s->h_bptr[(size_t)i + 1]; b++)
        if (s->h_bj[b] == (int32_t)j) {
            LG_CUDA(cudaMemcpyAsync(conn, s->bconn.p + (size_t)b * 15, 15 * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
            LG_CUDA(cudaMemcpyAsync(len, s->blen.p + (size_t)b * 5, 5 * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
            LG_CUDA(cudaStreamSynchronize(s->stream));
            *bonded = 1;
            break;
        }
    return LG_OK;
}

int lg_get_counters(lg_system* s, int64_t out[8]) {
    LG_ENTER(s);
    if (!out) return fail(LG_EINVAL, "null output");
    unsigned long long contacts = 0;
    LG_TRY(flush(s));
    if (s->slots_ready) {
        launch_count_slots((int)s->n, s->sgid.p, s->S, s->counter.p, s->stream);
        LG_CUDA(cudaMemcpyAsync(&contacts, s->counter.p, sizeof(contacts), cudaMemcpyDeviceToHost, s->stream));
        LG_CUDA(cudaStreamSynchronize(s->stream));
    }
    int64_t pairs = 0;
    lg_get_neighbor_count(s, &pairs);
    out[0] = pairs;
    out[1] = (int64_t)contacts;
    out[2] = s->n_bonds;
    out[3] = s->steps;
    out[4] = s->rebuilds;
    out[5] = g_launch_count - s->launches0;
    out[6] = s->last_cells;
    out[7] = 0;
    return LG_OK;
}

int lg_diagnostics(lg_system* s, double out[9]) {
    LG_ENTER(s);
    if (!out) return fail(LG_EINVAL, "null output");
    if (!s->state_set) return fail(LG_EINVAL, "lg_set_state has not been called");
    const int n = (int)s->n;
    double* h = s->h_pinned;
    launch_diag_sums(n, s->x.p, s->v.p, s->w.p, s->mp.mass, s->mp.inertia, nullptr, s->diag_partials.p, s->diag_out.p, 0,
                     nullptr, s->stream);
    LG_CUDA(cudaMemcpyAsync(h, s->diag_out.p, 11 * sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    LG_CUDA(cudaStreamSynchronize(s->stream));
    double com[3] = {h[8] / (double)n, h[9] / (double)n, h[10] / (double)n};
    double sums[11];
    std::memcpy(sums, h, sizeof(sums));
    LG_CUDA(cudaMemcpyAsync(s->diag_out.p + 12, com, 3 * sizeof(double), cudaMemcpyHostToDevice, s->stream));
    launch_diag_sums(n, s->x.p, s->v.p, s->w.p, s->mp.mass, s->mp.inertia, s->diag_out.p + 12, s->diag_partials.p,
                     s->diag_out.p, 1, nullptr, s->stream);
    LG_CUDA(cudaMemcpyAsync(h, s->diag_out.p + 11, sizeof(double), cudaMemcpyDeviceToHost, s->stream));
    LG_CUDA(cudaStreamSynchronize(s->stream));
    out[0] = sums[0] + sums[1];
    out[1] = sums[0];
    out[2] = sums[1];
    out[3] = s->mp.mass * std::sqrt(sums[2] * sums[2] + sums[3] * sums[3] + sums[4] * sums[4]);
    out[4] = std::sqrt(sums[5] * sums[5] + sums[6] * sums[6] + sums[7] * sums[7]);
    out[5] = com[0];
    out[6] = com[1];
    out[7] = com[2];
    out[8] = std::sqrt(h[0] / (double)n);
    return LG_OK;
}

// ---- slab decomposition support ------------------------------------------------------
int lg_flush(lg_system* s) {
    LG_ENTER(s);
    if (!s->state_set) return LG_OK;
    return flush(s);
}

int lg_first_half_kick(lg_system* s, double dt) {
    LG_ENTER(s);
    LG_TRY(prepare(s));
    if (s->kicked || s->half_kicked) return LG_OK;
    LG_TRY(evaluate(s, s->facets));
    launch_half_kick_back((int)s->n, s->v.p, s->w.p, s->a.p, s->al.p, dt, s->fixed_particle, s->stream);
    LG_CUDA(cudaGetLastError());
    s->half_kicked = true;
    return LG_OK;
}

int lg_import_state_device(lg_system* s, const double* d_x, const double* d_v, const double* d_theta, const double* d_omega,
                           const double* d_a, const double* d_alpha, int kicked) {
    LG_ENTER(s);
    if (s->reorder_mode) return fail(LG_EINVAL, "lg_import_state_device: the slab-decomposition entry points address particles by internal index and are not available after lg_set_reorder");
    if (!d_x || !d_v || !d_theta || !d_omega) return fail(LG_EINVAL, "null state pointer");
    LG_TRY(sync_pairs(s));
    const size_t bytes = 3 * (size_t)s->n * sizeof(double);
    cudaStream_t st = s->stream;
    LG_CUDA(cudaMemcpyAsync(s->x.p, d_x, bytes, cudaMemcpyDeviceToDevice, st));
    LG_CUDA(cudaMemcpyAsync(s->v.p, d_v, bytes, cudaMemcpyDeviceToDevice, st));
    LG_CUDA(cudaMemcpyAsync(s->th.p, d_theta, bytes, cudaMemcpyDeviceToDevice, st));
    LG_CUDA(cudaMemcpyAsync(s->w.p, d_omega, bytes, cudaMemcpyDeviceToDevice, st));
    if (d_a) LG_CUDA(cudaMemcpyAsync(s->a.p, d_a, bytes, cudaMemcpyDeviceToDevice, st)); else LG_CUDA(cudaMemsetAsync(s->a.p, 0, bytes, st));
    if (d_alpha) LG_CUDA(cudaMemcpyAsync(s->al.p, d_alpha, bytes, cudaMemcpyDeviceToDevice, st)); else LG_CUDA(cudaMemsetAsync(s->al.p, 0, bytes, st));
    LG_CUDA(cudaStreamSynchronize(st));
    s->state_set = true;
    s->kicked = kicked != 0;
    s->half_kicked = false;
    if (s->rows_staged) LG_TRY(install_staged_rows(s));  // the discarded evaluation would have installed them
    s->grid_valid = false;  // new positions: the next rebuild measures the bounding box again
    s->pending = false;
    return LG_OK;
}

int lg_set_owned_device(lg_system* s, const uint8_t* d_owned) {
    LG_ENTER(s);
    if (s->reorder_mode) return fail(LG_EINVAL, "lg_set_owned_device: the slab-decomposition entry points address particles by internal index and are not available after lg_set_reorder");
    if (!d_owned) {
        s->has_owned = false;
        return LG_OK;
    }
    LG_CUDA(s->owned.reserve((size_t)s->n));
    LG_CUDA(cudaMemcpyAsync(s->owned.p, d_owned, (size_t)s->n, cudaMemcpyDeviceToDevice, s->stream));
    LG_CUDA(cudaStreamSynchronize(s->stream));
    s->has_owned = true;
    return LG_OK;
}

int lg_export_slots(lg_system* s, int* S, int32_t** d_sgid, double** d_spr) {
    LG_ENTER(s);
    if (s->reorder_mode) return fail(LG_EINVAL, "lg_export_slots: the slab-decomposition entry points address particles by internal index and are not available after lg_set_reorder");
    if (!S || !d_sgid || !d_spr) return fail(LG_EINVAL, "null output");
    LG_TRY(ensure_slots(s));
    LG_TRY(flush(s));
    LG_CUDA(cudaStreamSynchronize(s->stream));
    *S = s->S;
    *d_sgid = s->sgid.p;
    *d_spr = s->spr.p;
    return LG_OK;
}

int lg_import_slots_device(lg_system* s, const int32_t* d_sgid, const double* d_spr) {
    LG_ENTER(s);
    if (s->reorder_mode) return fail(LG_EINVAL, "lg_import_slots_device: the slab-decomposition entry points address particles by internal index and are not available after lg_set_reorder");
    if (!d_sgid || !d_spr) return fail(LG_EINVAL, "null slot tables");
    LG_TRY(ensure_slots(s));
    const size_t cells = (size_t)s->n * (size_t)s->S;
    LG_CUDA(cudaMemcpyAsync(s->sgid.p, d_sgid, cells * sizeof(int32_t), cudaMemcpyDeviceToDevice, s->stream));
    LG_CUDA(cudaMemcpyAsync(s->spr.p, d_spr, cells * 9 * sizeof(double), cudaMemcpyDeviceToDevice, s->stream));
    LG_CUDA(cudaStreamSynchronize(s->stream));
    LG_TRY(flush(s));  // installs rows that were rebuilt behind a pending evaluation
    if (s->n_pairs > 0 || s->allpairs_built) LG_TRY(retag(s));
    return LG_OK;
}

int lg_diagnostics_partial(lg_system* s, int pass, const double* d_com, double** d_out) {
    LG_ENTER(s);
    if (!d_out || (pass != 0 && pass != 1) || (pass == 1 && !d_com)) return fail(LG_EINVAL, "bad arguments to lg_diagnostics_partial");
    if (!s->state_set) return fail(LG_EINVAL, "lg_set_state has not been called");
    const uint8_t* owned = s->has_owned ? s->owned.p : nullptr;
    if (pass == 1) LG_CUDA(cudaMemcpyAsync(s->diag_out.p + 12, d_com, 3 * sizeof(double), cudaMemcpyDeviceToDevice, s->stream));
    launch_diag_sums((int)s->n, s->x.p, s->v.p, s->w.p, s->mp.mass, s->mp.inertia, pass == 1 ? s->diag_out.p + 12 : nullptr,
                     s->diag_partials.p, s->diag_out.p, pass, owned, s->stream);
    LG_CUDA(cudaGetLastError());
    *d_out = s->diag_out.p;
    return LG_OK;
}

// The sums of lg_diagnostics_partial for caller-owned device arrays, allocating nothing: the per-block partials go to a
// file-scope device array in kernels_neighbor.cu (see launch_diag_sums_static).  Runs on the caller's current device.
int lg_diag_sums_static_max_particles(void) { return diag_static_max_particles(); }

int lg_diag_sums_static(int64_t n, const double* d_x, const double* d_v, const double* d_w, double mass, double inertia,
                        int pass, const double* d_com, double* d_out_12, void* stream) {
    if (n < 1) return fail(LG_EINVAL, "lg_diag_sums_static: n must be >= 1");
    if (n > (int64_t)diag_static_max_particles())
        return fail(LG_ESIZE, "lg_diag_sums_static: n = " + std::to_string(n) + " exceeds the static-scratch entry point's limit of " +
                                  std::to_string(diag_static_max_particles()) + " particles (use lg_diagnostics_partial on an lg_system)");
    if (pass != 0 && pass != 1) return fail(LG_EINVAL, "lg_diag_sums_static: pass must be 0 or 1");
    if (!d_x || !d_out_12 || (pass == 0 && (!d_v || !d_w)) || (pass == 1 && !d_com))
        return fail(LG_EINVAL, "lg_diag_sums_static: null device pointer");
    int dev = -1;
    LG_CUDA(cudaGetDevice(&dev));
    const void* checked[2] = {d_x, d_out_12};
    for (const void* p : checked) {  // a host pointer or another GPU's memory would fault inside the kernel: refuse it here
        cudaPointerAttributes at{};
        LG_CUDA(cudaPointerGetAttributes(&at, p));
        if (at.type != cudaMemoryTypeDevice && at.type != cudaMemoryTypeManaged)
            return fail(LG_EINVAL, "lg_diag_sums_static: d_x and d_out_12 must be device memory");
        if (at.type == cudaMemoryTypeDevice && at.device != dev)
            return fail(LG_EINVAL, "lg_diag_sums_static: the arrays live on device " + std::to_string(at.device) +
                                       " but the current device is " + std::to_string(dev));
    }
    const char* what = "";
    const int e = launch_diag_sums_static((int)n, d_x, d_v, d_w, mass, inertia, d_com, d_out_12, pass, (cudaStream_t)stream, &what);
    if (e != 0) return fail(LG_ECUDA, std::string(what) + ": " + cudaGetErrorString((cudaError_t)e));
    return LG_OK;
}

int lg_export_wall(lg_system* s, int facet_id, double** d_wspr, uint8_t** d_wact) {
    LG_ENTER(s);
    if (s->reorder_mode) return fail(LG_EINVAL, "lg_export_wall: not available after lg_set_reorder");
    if (facet_id < 0 || facet_id >= (int)s->facets.size() || !d_wspr || !d_wact) return fail(LG_EINVAL, "bad facet id or null output");
    LG_TRY(flush(s));
    *d_wspr = s->wspr[(size_t)facet_id]->p;
    *d_wact = s->wact[(size_t)facet_id]->p;
    return LG_OK;
}

int lg_import_wall_device(lg_system* s, int facet_id, const double* d_wspr, const uint8_t* d_wact) {
    LG_ENTER(s);
    if (s->reorder_mode) return fail(LG_EINVAL, "lg_import_wall_device: not available after lg_set_reorder");
    if (facet_id < 0 || facet_id >= (int)s->facets.size() || !d_wspr || !d_wact) return fail(LG_EINVAL, "bad facet id or null input");
    LG_TRY(flush(s));
    LG_CUDA(cudaMemcpyAsync(s->wspr[(size_t)facet_id]->p, d_wspr, (size_t)s->n * 9 * sizeof(double), cudaMemcpyDeviceToDevice, s->stream));
    LG_CUDA(cudaMemcpyAsync(s->wact[(size_t)facet_id]->p, d_wact, (size_t)s->n, cudaMemcpyDeviceToDevice, s->stream));
    LG_CUDA(cudaStreamSynchronize(s->stream));
    return LG_OK;
}

int lg_export_bonds(lg_system* s, int64_t* n_ordered, uint32_t** d_bptr, int32_t** d_bj, double** d_bconn, double** d_blen) {
    LG_ENTER(s);
    if (s->reorder_mode) return fail(LG_EINVAL, "lg_export_bonds: the slab-decomposition entry points address particles by internal index and are not available after lg_set_reorder");
    if (!n_ordered || !d_bptr || !d_bj || !d_bconn || !d_blen) return fail(LG_EINVAL, "null output");
    LG_CUDA(cudaStreamSynchronize(s->stream));
    *n_ordered = s->n_bonds;
    *d_bptr = s->bptr.p;
    *d_bj = s->bj.p;
    *d_bconn = s->bconn.p;
    *d_blen = s->blen.p;
    return LG_OK;
}

int lg_import_bonds_device(lg_system* s, const lg_contact_params* contact, double k_d, double gamma_d, int64_t nb,
                           const uint32_t* d_bptr, const int32_t* d_bj, const int32_t* d_brev, const double* d_bconn,
                           const double* d_blen) {
    LG_ENTER(s);
    if (s->reorder_mode) return fail(LG_EINVAL, "lg_import_bonds_device: the slab-decomposition entry points address particles by internal index and are not available after lg_set_reorder");
    if (!contact || nb < 0 || !d_bptr || (nb > 0 && (!d_bj || !d_brev || !d_bconn || !d_blen)))
        return fail(LG_EINVAL, "bad bond arguments");
    if (s->mask & MODEL_CONTACT) return fail(LG_EINVAL, "contact model already installed; the sinter model embeds it");
    const size_t n = (size_t)s->n;
    cudaStream_t st = s->stream;
    LG_CUDA(s->bptr.reserve(n + 1));
    LG_CUDA(s->bj.reserve((size_t)nb + 16));
    LG_CUDA(s->brev.reserve((size_t)nb + 16));
    LG_CUDA(s->bconn.reserve((size_t)nb * 15 + 16));
    LG_CUDA(s->blen.reserve((size_t)nb * 5 + 16));
    LG_CUDA(cudaMemcpyAsync(s->bptr.p, d_bptr, (n + 1) * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
    std::vector<uint32_t> bptr(n + 1);
    std::vector<int32_t> bj((size_t)nb), h_brev((size_t)nb);
    std::vector<double> h_len((size_t)nb * 5);
    LG_CUDA(cudaMemcpyAsync(bptr.data(), d_bptr, (n + 1) * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    if (nb) {
        LG_CUDA(cudaMemcpyAsync(s->bj.p, d_bj, (size_t)nb * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
        LG_CUDA(cudaMemcpyAsync(s->brev.p, d_brev, (size_t)nb * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
        LG_CUDA(cudaMemcpyAsync(s->bconn.p, d_bconn, (size_t)nb * 15 * sizeof(double), cudaMemcpyDeviceToDevice, st));
        LG_CUDA(cudaMemcpyAsync(s->blen.p, d_blen, (size_t)nb * 5 * sizeof(double), cudaMemcpyDeviceToDevice, st));
        LG_CUDA(cudaMemcpyAsync(bj.data(), d_bj, (size_t)nb * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        LG_CUDA(cudaMemcpyAsync(h_brev.data(), d_brev, (size_t)nb * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        LG_CUDA(cudaMemcpyAsync(h_len.data(), d_blen, (size_t)nb * 5 * sizeof(double), cudaMemcpyDeviceToHost, st));
    }
    LG_CUDA(cudaStreamSynchronize(st));
    if (bptr[n] != (uint32_t)nb) return fail(LG_EINVAL, "bptr[n] does not match the number of bond records");
    for (size_t b = 0; b < (size_t)nb; b++)
        if (h_brev[b] < 0 || h_brev[b] >= (int32_t)nb) return fail(LG_EINVAL, "mirror index out of range");
    keep_bond_mirrors(s, std::move(h_brev), h_len);
    for (size_t i = 0; i < n; i++)
        if (bptr[i + 1] - bptr[i] > (uint32_t)kMaxBondsPerParticle) return fail(LG_ECAPACITY, "more than 120 bonds on one particle");
    to_cp(contact, s->mp.cp);
    s->mp.k_d = k_d;
    s->mp.gamma_d = gamma_d;
    s->mask |= MODEL_SINTER;
    s->n_bonds = nb;
    s->h_bptr = std::move(bptr);
    s->h_bj = std::move(bj);
    LG_TRY(flush(s));  // installs rows that were rebuilt behind a pending evaluation
    if (s->n_pairs > 0 || s->allpairs_built) LG_TRY(retag(s));
    return LG_OK;
}

int lg_halo_pack(lg_system* s, int n_part, const int32_t* d_idx, int n_rec, const int32_t* d_rec, double* d_out) {
    LG_ENTER(s);
    if (s->reorder_mode) return fail(LG_EINVAL, "lg_halo_pack: the slab-decomposition entry points address particles by internal index and are not available after lg_set_reorder");
    if (n_part < 0 || n_rec < 0 || (n_part > 0 && !d_idx) || (n_rec > 0 && !d_rec) || !d_out) return fail(LG_EINVAL, "bad halo arguments");
    launch_halo_pack(n_part, d_idx, n_rec, d_rec, s->x.p, s->v.p, s->w.p, s->bconn.p, d_out, s->stream);
    LG_CUDA(cudaGetLastError());
    return LG_OK;
}

int lg_halo_unpack(lg_system* s, int n_part, const int32_t* d_idx, int n_rec, const int32_t* d_rec, const double* d_in) {
    LG_ENTER(s);
    if (s->reorder_mode) return fail(LG_EINVAL, "lg_halo_unpack: the slab-decomposition entry points address particles by internal index and are not available after lg_set_reorder");
    if (n_part < 0 || n_rec < 0 || (n_part > 0 && !d_idx) || (n_rec > 0 && !d_rec) || !d_in) return fail(LG_EINVAL, "bad halo arguments");
    launch_halo_unpack(n_part, d_idx, n_rec, d_rec, s->x.p, s->v.p, s->w.p, s->bconn.p, d_in, s->stream);
    LG_CUDA(cudaGetLastError());
    return LG_OK;
}

// ---- peer-to-peer halo exchange (cudaIpc mappings of the neighbours' receive buffers) -------------------------
struct lg_halo {
    int device = 0;
    int n_peers = 0;
    int64_t recv_doubles[kMaxHaloPeers] = {};
    int64_t recv_off[kMaxHaloPeers] = {};   // doubles, inside one parity half
    int64_t half_doubles = 0;               // doubles per parity half
    unsigned char* base = nullptr;          // [2 halves of recv data][flags: 2 x kMaxHaloPeers u64][counters: kMaxHaloPeers u32]
    size_t flag_off = 0, counter_off = 0, bytes = 0;
    HaloSendPeer send[kMaxHaloPeers] = {};
    HaloRecvPeer recv[kMaxHaloPeers] = {};
    void* peer_base[kMaxHaloPeers] = {};
    double* remote_data[kMaxHaloPeers][2] = {};
    unsigned long long* remote_flag[kMaxHaloPeers][2] = {};
    unsigned long long seq = 0;
    // halo stores fused into step_rows2_kernel: host mirrors of the send lists, items per tile for one tile shape
    std::vector<int32_t> h_send_idx[kMaxHaloPeers], h_send_rec[kMaxHaloPeers];
    int items_shape = -1;
    unsigned int items_tiles = 0;   // tiles that hold at least one item
    DevBuf<uint32_t> d_tile_ptr, d_tile_order;
    DevBuf<uint2> d_items;
    DevBuf<unsigned int> d_done;
};

namespace {
extern "C++" {
// Items of the fused halo store for tile shape `shape`: every particle / bond record of the send lists, keyed by the tile
// that computes it (tile of its particle), with its slot in the peer's message (position in the send list).
int build_halo_items(lg_system* s, lg_halo* hl, int shape) {
    const int T = rows2_tile_particles(shape);
    const size_t n = (size_t)s->n, n_tiles = (n + (size_t)T - 1) / (size_t)T;
    std::vector<std::vector<uint2>> per_tile(n_tiles);
    for (int k = 0; k < hl->n_peers; k++) {
        for (size_t q = 0; q < hl->h_send_idx[k].size(); q++) {
            const uint32_t i = (uint32_t)hl->h_send_idx[k][q];
            if (i >= n) return fail(LG_EINVAL, "halo send index out of range");
            per_tile[i / (uint32_t)T].push_back(make_uint2(((uint32_t)k << 24) | (i % (uint32_t)T), (uint32_t)q));
        }
        for (size_t q = 0; q < hl->h_send_rec[k].size(); q++) {
            const uint32_t b = (uint32_t)hl->h_send_rec[k][q];
            if (s->h_bptr.size() != n + 1 || b >= s->h_bptr[n]) return fail(LG_EINVAL, "halo send record out of range");
            const size_t i = (size_t)(std::upper_bound(s->h_bptr.begin(), s->h_bptr.end(), b) - s->h_bptr.begin()) - 1;  // owner of record b
            const size_t t = i / (size_t)T;
            const uint32_t rel = b - s->h_bptr[t * (size_t)T];
            if (rel >= 0x800000u) return fail(LG_ECAPACITY, "too many bond records in one tile");
            per_tile[t].push_back(make_uint2(((uint32_t)k << 24) | 0x800000u | rel, (uint32_t)q));
        }
    }
    std::vector<uint32_t> tp(n_tiles + 1, 0);
    std::vector<uint2> items;
    hl->items_tiles = 0;
    for (size_t t = 0; t < n_tiles; t++) {
        tp[t] = (uint32_t)items.size();
        items.insert(items.end(), per_tile[t].begin(), per_tile[t].end());
        hl->items_tiles += per_tile[t].empty() ? 0u : 1u;
    }
    tp[n_tiles] = (uint32_t)items.size();
    std::vector<uint32_t> order;  // tiles with items first (their peer stores then overlap the interior tiles), both ascending
    order.reserve(n_tiles);
    for (size_t t = 0; t < n_tiles; t++)
        if (!per_tile[t].empty()) order.push_back((uint32_t)t);
    for (size_t t = 0; t < n_tiles; t++)
        if (per_tile[t].empty()) order.push_back((uint32_t)t);
    LG_CUDA(hl->d_tile_order.reserve(n_tiles + 1));
    LG_CUDA(cudaMemcpyAsync(hl->d_tile_order.p, order.data(), n_tiles * sizeof(uint32_t), cudaMemcpyHostToDevice, s->stream));
    LG_CUDA(hl->d_tile_ptr.reserve(n_tiles + 1));
    LG_CUDA(hl->d_items.reserve(items.size() + 1));
    if (!hl->d_done.p) {
        LG_CUDA(hl->d_done.reserve(1));
        LG_CUDA(cudaMemsetAsync(hl->d_done.p, 0, sizeof(unsigned int), s->stream));
    }
    LG_CUDA(cudaMemcpyAsync(hl->d_tile_ptr.p, tp.data(), (n_tiles + 1) * sizeof(uint32_t), cudaMemcpyHostToDevice, s->stream));
    if (!items.empty()) LG_CUDA(cudaMemcpyAsync(hl->d_items.p, items.data(), items.size() * sizeof(uint2), cudaMemcpyHostToDevice, s->stream));
    LG_CUDA(cudaStreamSynchronize(s->stream));
    hl->items_shape = shape;
    return LG_OK;
}

int fill_halo_fused(lg_system* s, ForceArgs& A) {
    static const int enabled = [] { const char* e = std::getenv("LGB200_HALO_FUSED"); return e ? std::atoi(e) : 1; }();
    lg_halo* hl = s->active_halo;
    if (!enabled || hl == nullptr || hl->n_peers == 0) return LG_OK;
    if (hl->items_shape != A.rows_shape) LG_TRY(build_halo_items(s, hl, A.rows_shape));
    if (hl->items_tiles == 0) return LG_OK;  // nothing to send from any tile: the push kernel handles (and publishes) it
    const unsigned long long seq = hl->seq + 1;  // the exchange halo_exchange is about to run
    const int par = (int)(seq & 1ull);
    HaloFusedArgs& H = A.hf;
    H.n_peers = hl->n_peers;
    for (int k = 0; k < hl->n_peers; k++) {
        const long ts = 9L * hl->send[k].n_part + 15L * hl->send[k].n_rec;
        if (hl->remote_flag[k][par] == nullptr || (ts > 0 && hl->remote_data[k][par] == nullptr)) return fail(LG_EINVAL, "halo peer not connected");
        H.remote[k] = hl->remote_data[k][par];
        H.rec_off[k] = 9LL * hl->send[k].n_part;
        H.remote_flag[k] = hl->remote_flag[k][par];
    }
    H.tile_ptr = hl->d_tile_ptr.p;
    H.tile_order = hl->d_tile_order.p;
    H.items = hl->d_items.p;
    H.done = hl->d_done.p;
    H.expected = hl->items_tiles;
    H.seq = seq;
    s->halo_stored_by_step = true;
    return LG_OK;
}
}  // extern "C++"

int halo_exchange(lg_system* s, lg_halo* hl) {
    if (hl->n_peers == 0) return LG_OK;
    hl->seq++;
    const int par = (int)(hl->seq & 1ull);
    HaloPushArgs P{};
    HaloPullArgs Q{};
    P.n_peers = Q.n_peers = hl->n_peers;
    int pb = 0, qb = 0;
    for (int k = 0; k < hl->n_peers; k++) {
        P.block_begin[k] = pb;
        Q.block_begin[k] = qb;
        P.peer[k] = hl->send[k];
        const long ts = 9L * hl->send[k].n_part + 15L * hl->send[k].n_rec;
        if (ts > 0 && hl->remote_data[k][par] == nullptr) return fail(LG_EINVAL, "halo peer not connected");
        P.peer[k].remote = hl->remote_data[k][par];
        P.peer[k].remote_flag = hl->remote_flag[k][par];
        P.peer[k].counter = (unsigned int*)(hl->base + hl->counter_off) + k;
        pb += (int)((ts + 255) / 256);
        Q.peer[k] = hl->recv[k];
        const long tr = 9L * hl->recv[k].n_part + 15L * hl->recv[k].n_rec;
        Q.peer[k].local = (const double*)hl->base + (size_t)par * hl->half_doubles + hl->recv_off[k];
        Q.peer[k].flag = (const unsigned long long*)(hl->base + hl->flag_off) + (size_t)par * kMaxHaloPeers + k;
        qb += (int)((tr + 255) / 256);
    }
    P.block_begin[hl->n_peers] = pb;
    Q.block_begin[hl->n_peers] = qb;
    P.x = s->x.p; P.v = s->v.p; P.w = s->w.p; P.bconn = s->bconn.p;
    Q.x = s->x.p; Q.v = s->v.p; Q.w = s->w.p; Q.bconn = s->bconn.p;
    P.seq = Q.seq = hl->seq;
    Q.timeout_ns = 20ull * 1000000000ull;
    Q.err = s->err.p;
    if (!s->halo_stored_by_step) launch_halo_push(P, s->stream);  // else: the step kernel has stored and published this exchange
    s->halo_stored_by_step = false;
    launch_halo_pull(Q, s->stream);
    LG_CUDA(cudaGetLastError());
    return LG_OK;
}
}  // namespace

int lg_halo_create(lg_system* s, int n_peers, const int64_t* recv_doubles, lg_halo** out) {
    LG_ENTER(s);
    if (s->reorder_mode) return fail(LG_EINVAL, "lg_halo_create: the slab-decomposition entry points address particles by internal index and are not available after lg_set_reorder");
    if (!out || n_peers < 0 || n_peers > kMaxHaloPeers || (n_peers > 0 && !recv_doubles)) return fail(LG_EINVAL, "bad halo arguments");
    std::unique_ptr<lg_halo> hl(new (std::nothrow) lg_halo());
    if (!hl) return fail(LG_ENOMEM, "host allocation failed");
    hl->device = s->device;
    hl->n_peers = n_peers;
    int64_t off = 0;
    for (int k = 0; k < n_peers; k++) {
        if (recv_doubles[k] < 0) return fail(LG_EINVAL, "negative halo size");
        hl->recv_doubles[k] = recv_doubles[k];
        hl->recv_off[k] = off;
        off += (recv_doubles[k] + 1) & ~(int64_t)1;  // 16-byte granules
    }
    hl->half_doubles = off;
    hl->flag_off = (size_t)(2 * off) * sizeof(double);
    hl->counter_off = hl->flag_off + 2 * kMaxHaloPeers * sizeof(unsigned long long);
    hl->bytes = hl->counter_off + kMaxHaloPeers * sizeof(unsigned int);
    LG_CUDA(cudaMalloc(&hl->base, hl->bytes));   // cudaMalloc (not a pool): the allocation is exported with cudaIpc
    LG_CUDA(cudaMemset(hl->base, 0, hl->bytes));
    LG_CUDA(cudaDeviceSynchronize());
    *out = hl.release();
    return LG_OK;
}

int lg_halo_set_lists(lg_halo* hl, int k, int n_send_part, const int32_t* d_send_idx, int n_send_rec, const int32_t* d_send_rec,
                      int n_recv_part, const int32_t* d_recv_idx, int n_recv_rec, const int32_t* d_recv_rec) {
    if (!hl || k < 0 || k >= hl->n_peers) return fail(LG_EINVAL, "bad halo peer");
    if (n_send_part < 0 || n_send_rec < 0 || n_recv_part < 0 || n_recv_rec < 0) return fail(LG_EINVAL, "negative halo list size");
    if (9LL * n_recv_part + 15LL * n_recv_rec > hl->recv_doubles[k]) return fail(LG_ESIZE, "halo receive list exceeds the buffer");
    {   // host mirrors of the send lists: the fused halo store of step_rows2_kernel is organised by tile on the host
        DeviceGuard guard_(hl->device);
        hl->h_send_idx[k].resize((size_t)n_send_part);
        hl->h_send_rec[k].resize((size_t)n_send_rec);
        if (n_send_part) LG_CUDA(cudaMemcpy(hl->h_send_idx[k].data(), d_send_idx, (size_t)n_send_part * sizeof(int32_t), cudaMemcpyDeviceToHost));
        if (n_send_rec) LG_CUDA(cudaMemcpy(hl->h_send_rec[k].data(), d_send_rec, (size_t)n_send_rec * sizeof(int32_t), cudaMemcpyDeviceToHost));
        hl->items_shape = -1;
    }
    hl->send[k].n_part = n_send_part; hl->send[k].idx = d_send_idx;
    hl->send[k].n_rec = n_send_rec;   hl->send[k].rec = d_send_rec;
    hl->recv[k].n_part = n_recv_part; hl->recv[k].idx = d_recv_idx;
    hl->recv[k].n_rec = n_recv_rec;   hl->recv[k].rec = d_recv_rec;
    return LG_OK;
}

int lg_halo_export(lg_halo* hl, int k, unsigned char handle[64], int64_t offs[4]) {
    if (!hl || k < 0 || k >= hl->n_peers || !handle || !offs) return fail(LG_EINVAL, "bad halo peer");
    DeviceGuard guard_(hl->device);
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    cudaIpcMemHandle_t h;
    LG_CUDA(cudaIpcGetMemHandle(&h, hl->base));
    std::memcpy(handle, &h, 64);
    for (int par = 0; par < 2; par++) {
        offs[par] = (int64_t)(((size_t)par * hl->half_doubles + hl->recv_off[k]) * sizeof(double));
        offs[2 + par] = (int64_t)(hl->flag_off + ((size_t)par * kMaxHaloPeers + k) * sizeof(unsigned long long));
    }
    return LG_OK;
}

int lg_halo_connect(lg_halo* hl, int k, const unsigned char handle[64], const int64_t offs[4]) {
    if (!hl || k < 0 || k >= hl->n_peers || !handle || !offs) return fail(LG_EINVAL, "bad halo peer");
    DeviceGuard guard_(hl->device);
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, 64);
    void* p = nullptr;
    LG_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    hl->peer_base[k] = p;
    for (int par = 0; par < 2; par++) {
        hl->remote_data[k][par] = (double*)((unsigned char*)p + offs[par]);
        hl->remote_flag[k][par] = (unsigned long long*)((unsigned char*)p + offs[2 + par]);
    }
    return LG_OK;
}

int lg_halo_disconnect(lg_halo* hl) {
    if (!hl) return LG_OK;
    DeviceGuard guard_(hl->device);
    for (int k = 0; k < hl->n_peers; k++)
        if (hl->peer_base[k]) {
            cudaIpcCloseMemHandle(hl->peer_base[k]);
            hl->peer_base[k] = nullptr;
            hl->remote_data[k][0] = hl->remote_data[k][1] = nullptr;
        }
    return LG_OK;
}

int lg_halo_destroy(lg_halo* hl) {
    if (!hl) return LG_OK;
    lg_halo_disconnect(hl);
    {
        DeviceGuard guard_(hl->device);
        if (hl->base) cudaFree(hl->base);
    }
    delete hl;
    return LG_OK;
}

int lg_halo_exchange(lg_system* s, lg_halo* hl) {
    LG_ENTER(s);
    if (!hl) return fail(LG_EINVAL, "null halo");
    return halo_exchange(s, hl);
}

int lg_run_slab(lg_system* s, lg_halo* hl, double dt, int64_t n_steps, int move_facets) {
    LG_ENTER(s);
    if (!hl) return fail(LG_EINVAL, "null halo");
    LG_TRY(prepare(s));
    s->active_halo = hl;  // fused launches of step_rows2_kernel store the halo themselves
    int rc = LG_OK;
    for (int64_t k = 0; k < n_steps && rc == LG_OK; k++) {
        rc = step(s, dt);
        if (rc == LG_OK && move_facets)  // facet.update_positions(dt) after every step, as test/surface.cpp's loop does
            for (auto& f : s->facets)
                for (int q = 0; q < 3; q++) f.vtx[q] = f.vtx[q] + f.v * dt;
        if (rc == LG_OK) rc = halo_exchange(s, hl);
    }
    s->active_halo = nullptr;
    s->halo_stored_by_step = false;
    return rc;
}

int lg_get_stream(lg_system* s, void** stream) {
    LG_ENTER(s);
    if (!stream) return fail(LG_EINVAL, "null output");
    *stream = (void*)s->stream;
    return LG_OK;
}

int lg_set_profiling(lg_system* s, int enabled) {
    LG_ENTER(s);
    LG_TRY(drain_profile(s));
    s->profiling = enabled != 0;
    s->prof_ms = 0.0;
    s->prof_count = 0;
    return LG_OK;
}

int lg_get_profile(lg_system* s, double* ms, int64_t* count) {
    LG_ENTER(s);
    if (!ms || !count) return fail(LG_EINVAL, "null output");
    LG_TRY(drain_profile(s));
    *ms = s->prof_ms;
    *count = s->prof_count;
    return LG_OK;
}

int lg_set_math_mode(lg_system* s, int mode) {
    LG_ENTER(s);
    if (mode != 0 && mode != 1) return fail(LG_EINVAL, "math mode must be LG_MATH_EXACT (0) or LG_MATH_FAST (1)");
    if (mode == 1 && s->variant < 1) return fail(LG_EINVAL, "the relaxed math mode exists for the production kernel variants (1, 2) only");
    LG_TRY(flush(s));
    s->math_mode = mode;
    return LG_OK;
}

int lg_last_step_kernel(lg_system* s, char* out, int cap) {
    LG_ENTER(s);
    if (!out || cap <= 0) return fail(LG_EINVAL, "null output");
    std::snprintf(out, (size_t)cap, "%s", s->last_kernel);
    return LG_OK;
}

int lg_set_kernel_variant(lg_system* s, int variant) {
    LG_ENTER(s);
    if (variant < 0 || variant > 2) return fail(LG_EINVAL, "variant must be 0 (validation), 1 (production) or 2 (row kernel forced)");
    if (s->simple_only && variant != 0)
        return fail(LG_EINVAL, "alt_sinter / forward Euler / fixed-particle / central-attraction systems run kernel variant 0 only");
    LG_TRY(flush(s));
    if (variant == 0 && s->math_mode != 0) return fail(LG_EINVAL, "the validation variant is bit-exact only; reset the math mode first");
    s->variant = variant;
    return LG_OK;
}

}  // extern "C"

#include "dump_impl.cuh"
