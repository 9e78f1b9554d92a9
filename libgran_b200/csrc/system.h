// system.h -- device-resident state of one granular system and the kernel launchers that act on it.
//
// Layout in HBM (all FP64 unless noted; n particles, P ordered neighbor pairs, S contact slots per
// particle, Bd ordered bonds, F facets):
//   x, v, a, theta, omega, alpha   3n doubles each, xyz-interleaved, caller's particle order
//                                  (replaces the six std::vector<Eigen::Vector3d> of
//                                  LT/rotational_system/rotational_system.h:128)
//   row_ptr  u32[n+1], cols i32[P], tag u8[P]
//                                  CSR neighbor rows, ascending j (replaces vector<vector<long>>,
//                                  rotational_binary_system_neighbors_omp.h:122); tag[p] says where the
//                                  pair's history lives: 0..S-1 contact slot, 128+k the k-th bond of i,
//                                  255 nothing
//   spr      9 doubles [n][S], sgid i32[n][S]
//                                  sparse per-ORDERED-pair contact springs owned by i (replaces the dense
//                                  N x N tuple store of contact_force.h:53,156); sgid = partner id or -1
//   bptr u32[n+1], bj i32[Bd], brev i32[Bd], bconn 15 doubles [Bd], blen 5 doubles [Bd]
//                                  bond CSR (replaces the dense tables of sinter_bridge.h:139-141 and the
//                                  particle_to_connector_map :138)
//   wspr 9 doubles [F][n], wact u8[F][n]
//                                  wall springs per facet (surface_contact_force.h:53) + "may be non-zero" flag
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

#include "physics.cuh"

namespace lg {

constexpr int kMaxFacets = 8;
constexpr int kTagNone = 255;
constexpr int kTagBond = 128;
constexpr int kMaxBondsPerParticle = 120;
constexpr int kMaxSlots = 120;

// asynchronous error flags raised by kernels (checked at synchronisation points)
enum : int { ERR_SLOT_OVERFLOW = 1, ERR_BOND_LOOKUP = 2, ERR_HALO_TIMEOUT = 4, ERR_PAIR_CAPACITY = 8 };

struct ModelParams {
    double r, mass, inertia;
    ContactParams cp;
    double hamaker_A, hamaker_h0;
    double k_d, gamma_d;
    // SURVEY.md section 8f variants
    int hamaker_kind;       // HAMAKER_*
    double hamaker_cutoff;
    int alt_bonds;          // the bond table belongs to alt_sinter_functor: no connectors, three springs per ordered bond
    double alt[8];          // its k, gamma_n, k_t, gamma_t, k_r, gamma_r, k_o, gamma_o
    double alt_dt;          // its own spring-update time step (alt_sinter_bridge.h:30)
};

// Halo stores fused into step_rows2_kernel (multi-GPU slab runs): the CTA that finishes a tile copies the tile's halo items
// -- new x, v, omega of particles a neighbouring rank holds as ghosts, new connector blocks of the bond records it mirrors --
// from the staged results straight into that rank's receive buffer over NVLink, fences, and counts the tile; the tile that
// completes the count publishes the exchange's sequence number (release, system scope).  Replaces the separate halo_push_kernel launch.
constexpr int kMaxHaloPeers = 8;
struct HaloFusedArgs {
    const uint32_t* tile_ptr;   // [n_tiles + 1] first item of every tile; nullptr = off
    const uint32_t* tile_order; // [n_tiles] the order in which the grid walks the tiles: tiles with items first
    const uint2* items;         // x: peer << 24 | kind << 23 | index inside the tile (kind 0: particle row, 1: bond record
                                //    relative to the tile's first record), y: slot in the peer's message
    int n_peers;
    double* remote[kMaxHaloPeers];              // peer memory of this exchange's parity: particle slot k at 9 k
    long long rec_off[kMaxHaloPeers];           // doubles: connector block k at rec_off + 15 k
    unsigned long long* remote_flag[kMaxHaloPeers];
    unsigned int* done;         // tiles of this launch that have stored their items
    unsigned int expected;      // tiles that hold items: the one that completes the count publishes
    unsigned long long seq;
};

struct ForceArgs {
    int n;
    const double* x;
    const double* v;
    const double* w;
    double* a;
    double* al;
    const uint32_t* row_ptr;
    const int32_t* cols;
    uint8_t* tag;
    double* spr;
    int32_t* sgid;
    int S;
    const uint32_t* bptr;
    const int32_t* bj;
    const int32_t* brev;
    const double* bconn;
    const double* blen;
    double* bspr;               // alt_sinter: 9 doubles per ordered bond record (its three springs), updated in place
    int fixed;                  // fixed_particle_step_handler: this particle's velocity increments are dropped (-1: none)
    int central_on;             // central_attraction_functor (unary): a += k_a / m * (x_i - x0)
    double central_k;
    d3 central_x0;
    const uint16_t* bwork;      // step_rows2_kernel: bridge work items per tile (api.cu: build_bridge_work)
    const uint32_t* bwork_cnt;  // items per tile
    ModelParams mp;
    int n_facets;
    Facet facets[kMaxFacets];
    double* wspr[kMaxFacets];
    uint8_t* wact[kMaxFacets];
    int* err;
    // tiled / fused kernel only
    int rows_per_block;
    int fast;           // relaxed arithmetic (lg_set_math_mode)
    int rows_shape;     // step_rows2_kernel: tile shape index (rows2_shape_for)
    HaloFusedArgs hf;   // step_rows2_kernel, fused launches of slab runs
    int use_rows;       // fast only: short rows -> a row kernel instead of the tile kernel (1 step_rows_kernel, 2 step_rows2_kernel)
    FastConsts fc;      // its host-computed constants
    double dt;          // integration step of the fused update
    double* x_next;     // next-step buffers (double-buffered state)
    double* v_next;
    double* w_next;
    double* th;         // theta is private to its particle: updated in place
    double* bconn_next;
};

struct IntegrateArgs {
    int n;
    double* x;
    double* v;
    double* th;
    double* w;
    const double* a;
    const double* al;
    const uint32_t* bptr;  // null when there are no bonds
    double* bconn;
    int64_t n_bonds;       // ordered bonds
    double dt;
    int scheme;            // 0 velocity-Verlet half (leap-frog), 1 forward Euler with the two-term Taylor position update
    int fixed;             // particle whose increment_v is dropped, -1 none
};

struct NeighborGrid {
    double lo[3];
    double cell;
    int dim[3];
    int64_t n_cells;
};

// number of kernels launched by this thread (lg_get_counters[5]); bumped by every launcher
extern thread_local int64_t g_launch_count;

// ---- launchers (kernels_*.cu) ---------------------------------------------------------------------
void launch_force(const ForceArgs& args, int mask, int variant, cudaStream_t st);
// tiled kernel: force only (fused == false) or force + integrate into the next-step buffers (fused == true)
void launch_step_tiled(const ForceArgs& args, int mask, bool fused, cudaStream_t st);
void launch_rows2(const ForceArgs& args, int mask, bool fused, cudaStream_t st);  // kernels_rows.cu
int rows2_shape_for(double k_mean, int n, bool bonds);   // tile shape index of step_rows2_kernel for a mean row length
int rows2_tile_particles(int shape);
int rows2_tile_records(int shape);
int tiled_max_rows_per_block();
int tiled_tile_pairs();
void launch_integrate(const IntegrateArgs& args, cudaStream_t st);
void launch_half_kick_back(int n, double* v, double* w, const double* a, const double* al, double dt, int fixed, cudaStream_t st);

// bounding box of x -> d_out[6] = {min xyz, max xyz}
void launch_bbox(int n, const double* x, double* d_scratch, double* d_out, cudaStream_t st);
// exclusive prefix sum of u32 (in place allowed); scratch must hold scan_scratch_elems(n) u32
size_t scan_scratch_elems(size_t n);
void launch_exclusive_scan_u32(uint32_t* data, size_t n, uint32_t* scratch, cudaStream_t st);

// device-side bond detection (kernels_bonds.cu)
void launch_bond_count(int n, const double* x, const uint32_t* row_ptr, const int32_t* cols, double r, double crit, uint32_t* deg,
                       uint32_t* up, cudaStream_t st);
void launch_bond_fill(int n, const double* x, const uint32_t* row_ptr, const int32_t* cols, double r, double crit, const uint32_t* bptr,
                      const uint32_t* upoff, const double* uv, int64_t n_uv, int32_t* bj, int32_t* brev, double* bconn, double* blen,
                      int* flags, cudaStream_t st);
// spatial renumbering (kernels_neighbor.cu)
size_t morton_cells();
void launch_morton_order(int n, const double* x, const double lo[3], double cell, uint32_t* key, uint32_t* count, uint32_t* cursor,
                         uint32_t* scan_scratch, int32_t* tmp, int32_t* perm, int32_t* inv, cudaStream_t st);
void launch_gather3(int n, const int32_t* idx, const double* src, double* dst, cudaStream_t st);
void launch_cell_count(int n, const double* x, NeighborGrid g, int32_t* cell_of, uint32_t* cell_count, cudaStream_t st);
void launch_cell_fill(int n, const int32_t* cell_of, const uint32_t* cell_start, uint32_t* cell_cursor, int32_t* order,
                      cudaStream_t st);
// hits: n * neighbor_hit_list_entries() ints of scratch that carries the pairs found by the count pass to the fill pass
// (which then does not sweep the cells again), or nullptr
int neighbor_hit_list_entries();
void launch_row_count(int n, const double* x, NeighborGrid g, const uint32_t* cell_start, const int32_t* order,
                      double r_verlet, uint32_t* row_count, int32_t* hits, const uint8_t* owned, cudaStream_t st);
void launch_row_fill(int n, const double* x, NeighborGrid g, const uint32_t* cell_start, const int32_t* order,
                     double r_verlet, const uint32_t* row_ptr, int32_t* cols, const int32_t* hits, const uint8_t* owned,
                     cudaStream_t st);
void launch_allpairs_fill(int n, uint32_t* row_ptr, int32_t* cols, cudaStream_t st);
void launch_pair_guard(int n, uint32_t* row_ptr, uint32_t cap, uint32_t* d_total, int* err, cudaStream_t st);
// recompute tag[] for every row from the slot table and the bond CSR; frees slots whose partner left the row
void launch_retag(int n, const uint32_t* row_ptr, const int32_t* cols, uint8_t* tag, int32_t* sgid, double* spr, int S,
                  const uint32_t* bptr, const int32_t* bj, int* err, cudaStream_t st);

void launch_count_slots(int n, const int32_t* sgid, int S, unsigned long long* d_out, cudaStream_t st);
// out[16]: sums for lg_diagnostics (pass 0) ; pass 1 adds sum |x - com|^2
void launch_diag_sums(int n, const double* x, const double* v, const double* w, double mass, double inertia,
                      const double* com, double* d_partials, double* d_out, int pass, const uint8_t* owned, cudaStream_t st);
// the same two kernels with the partial sums in a file-scope device array (no allocation; lg_diag_sums_static); returns a
// cudaError_t as int and names the failing call in *what
int launch_diag_sums_static(int n, const double* x, const double* v, const double* w, double mass, double inertia,
                            const double* com, double* d_out, int pass, cudaStream_t st, const char** what);
int diag_static_max_particles();

// slab decomposition: ghost staging (kernels_halo.cu)
void launch_halo_pack(int n_part, const int32_t* idx, int n_rec, const int32_t* rec, const double* x, const double* v,
                      const double* w, const double* bconn, double* out, cudaStream_t st);
void launch_halo_unpack(int n_part, const int32_t* idx, int n_rec, const int32_t* rec, double* x, double* v, double* w,
                        double* bconn, const double* in, cudaStream_t st);

// ---- peer-to-peer halo exchange over NVLink (kernels_halo.cu) ------------------------------------------------
// One rank's view of the exchange with up to kMaxHaloPeers neighbouring ranks.  The SEND side gathers the owners'
// x, v, omega (9 doubles per particle) and the connector blocks of cross-slab bond records (15 doubles each) and
// stores them straight into the receiving rank's buffer (a cudaIpc mapping of peer memory), then publishes a
// sequence number there; the RECEIVE side polls its own flag and scatters the buffer into the ghost slots.
struct HaloSendPeer {
    int n_part, n_rec;
    const int32_t* idx;
    const int32_t* rec;
    double* remote;                    // peer memory (this step's parity)
    unsigned long long* remote_flag;   // peer memory
    unsigned int* counter;             // local: blocks of this peer that have finished storing
};
struct HaloRecvPeer {
    int n_part, n_rec;
    const int32_t* idx;
    const int32_t* rec;
    const double* local;               // this step's parity
    const unsigned long long* flag;
};
struct HaloPushArgs {
    int n_peers;
    int block_begin[kMaxHaloPeers + 1];
    HaloSendPeer peer[kMaxHaloPeers];
    const double *x, *v, *w, *bconn;
    unsigned long long seq;
};
struct HaloPullArgs {
    int n_peers;
    int block_begin[kMaxHaloPeers + 1];
    HaloRecvPeer peer[kMaxHaloPeers];
    double *x, *v, *w, *bconn;
    unsigned long long seq;
    unsigned long long timeout_ns;
    int* err;
};
void launch_halo_push(const HaloPushArgs& args, cudaStream_t st);
void launch_halo_pull(const HaloPullArgs& args, cudaStream_t st);

}  // namespace lg
