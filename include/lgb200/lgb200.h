/*
 * lgb200.h -- the C ABI of libgran-b200 (libgran_b200/liblgb200.so).
 *
 * This is the drop-in boundary: plain pointers and sizes, `extern "C"`, int return codes, no C++ or
 * torch types.  The reference has no FFI -- its hot path sits behind a C++ template API -- so each
 * entry point below names the reference interface it stands in for (paths relative to the reference
 * repository root; LT = deps/libtimestep/include/libtimestep, LG = include/libgran).  The header
 * templates under include/libgran/ and include/libtimestep/ of THIS repository (same class names and
 * constructor signatures as the reference) call nothing but these functions; so do the Python
 * bindings in libgran_b200/system.py used by tests/ and bench.py.
 *
 * Conventions
 *   - every function returns LG_OK (0) or a negative LG_E* code; lg_last_error() returns a
 *     thread-local human-readable message for the last failure.  No exception crosses the ABI.
 *   - all particle fields are FP64, array-of-xyz, length 3*n doubles, in the caller's particle
 *     order (the order of the vectors handed to the reference's granular_system constructor).
 *   - host pointers unless a parameter is named `d_*` (device pointer on the system's device).
 *   - a system handle is not thread-safe; it owns one CUDA stream on one device.
 *   - there is NO CPU fallback: lg_create fails with LG_ENODEVICE when no CUDA device is usable.
 */
#ifndef LGB200_H
#define LGB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LGB200_ABI_VERSION 1

enum {
    LG_OK = 0,
    LG_EINVAL = -1,     /* bad argument / call order                                    */
    LG_ENODEVICE = -2,  /* no usable CUDA device (there is no CPU fallback)             */
    LG_ECUDA = -3,      /* a CUDA runtime call or kernel failed                         */
    LG_ENOMEM = -4,     /* device or host allocation failed                             */
    LG_ESIZE = -5,      /* size mismatch -- the ABI's form of LT/exception/exception.h:11-24
                           SizeMismatchException (LT/rotational_system/rotational_system.h:69-70) */
    LG_ECAPACITY = -6   /* a fixed-capacity table overflowed (contact slots per particle,
                           bonds per particle, pairs >= 2^32); see lg_set_max_contacts  */
};

/* field ids: the six state vectors of LT/rotational_system/rotational_system.h:94-121,128 */
enum { LG_X = 0, LG_V = 1, LG_A = 2, LG_THETA = 3, LG_OMEGA = 4, LG_ALPHA = 5 };

/* Constructor arguments of LG/contact_force/contact_force.h:10-30 (and of the wall twin
 * LG/contact_force/surface_contact_force.h:10-30) that are not common to the whole system. */
typedef struct lg_contact_params {
    double k, gamma_n;
    double k_t, gamma_t, mu_s, phi_d;
    double k_r, gamma_r, mu_r, phi_r;
    double k_o, gamma_o, mu_o, phi_o;
    double dt; /* the functor's own spring-update time step (contact_force.h:28,138) */
} lg_contact_params;

typedef struct lg_system lg_system;

/* ---- lifetime ------------------------------------------------------------------------------- */

/* granular_system / granular_system_neighbor_list constructor, part 1
 * (LG/granular_system/granular_system.h:53-64, granular_system_neighbor_list.h:54-64):
 * n particles of equal radius r_part, mass and moment of inertia (the constants every reference
 * functor receives, e.g. contact_force.h:25-27).  device = CUDA ordinal, or -1 for the current one. */
int lg_create(int64_t n, double r_part, double mass, double inertia, int device, lg_system** out);
int lg_destroy(lg_system* s);
const char* lg_last_error(void);
int lg_abi_version(void);

/* ---- state ---------------------------------------------------------------------------------- */

/* The x0, v0, theta0, omega0 vectors moved into rotational_generic_system
 * (LT/rotational_system/rotational_system.h:52-66).  Copies host -> device; resets a, alpha to zero
 * (:80,83-86) and re-arms the first-step half kick (LT/rotational_integrator/
 * rotational_velocity_verlet_half.h:45-60). `n_check` must equal the n given to lg_create, else
 * LG_ESIZE (rotational_system.h:69-70).  A second call starts a new trajectory exactly as constructing a new
 * reference system would: contact and wall history springs are cleared, the neighbor list is empty again until
 * lg_update_neighbor_list, the step counter of lg_run restarts; installed models and bond tables stay. */
int lg_set_state(lg_system* s, int64_t n_check, const double* x, const double* v, const double* theta,
                 const double* omega);

/* Spatial renumbering (no reference counterpart).  mode 1: at the first lg_set_state the
 * particles are renumbered internally along a Morton curve (counting sort by cell on the device), so that the neighbours of
 * consecutive particles are mostly consecutive -- what the staged position window of the row kernel needs on inputs whose
 * caller order has no spatial locality (random packings, fractal aggregates).  Invisible at this boundary: lg_set_state,
 * lg_get_field, bond tables, lg_get_neighbor_rows, lg_get_contact_springs, lg_get_bond ... all speak the caller's order.
 * The numbering is fixed by the first lg_set_state.  Rows are then accumulated in ascending INTERNAL index, so results are
 * no longer bit-identical to the reference's ascending-j sums (they stay within rounding of them): meant for LG_MATH_FAST.
 * Must precede the first lg_set_state; the slab-decomposition entry points below refuse renumbered systems. */
int lg_set_reorder(lg_system* s, int mode);

/* get_x / get_v / get_a / get_theta / get_omega / get_alpha (rotational_system.h:94-121): copies
 * device -> host, in the caller's particle order.  Implies a stream synchronisation. */
int lg_get_field(lg_system* s, int field, double* out_3n);

/* Device-resident view of a field for callers that keep data on the GPU (3*n doubles, xyz-interleaved,
 * in the caller's particle order as long as no reordering is active).  Valid until lg_destroy. */
int lg_device_field(lg_system* s, int field, double** d_ptr);

/* ---- binary force models (the members of LG/granular_system/binary_force_container.h:17-34) --- */

/* contact_force_functor (LG/contact_force/contact_force.h:8-157). */
int lg_set_contact(lg_system* s, const lg_contact_params* p);

/* hamaker_functor (LG/hamaker_force/hamaker_force.h:19-48). */
int lg_set_hamaker(lg_system* s, double A, double h0);

/* sinter_functor constructor (LG/sinter_bridge/sinter_bridge.h:101-212): bonds every pair with
 * |dist - 2 r| <= critical_separation using the positions of the last lg_set_state, drawing one unit
 * vector per bond, in lexicographic (i < j) order, from `gen` (the random_unit_vector_generator
 * argument, :124; redrawn while collinear with the bond axis, :171-175).  Also installs the embedded
 * contact model (:134-135) from `contact` and arms the connector update of sinter_step_handler
 * (:24-42,55-75).  n_bonds_out (nullable) receives the number of unordered bonds. */
typedef void (*lg_unit_vector_fn)(void* user, double out_xyz[3]);
int lg_build_sinter_bonds(lg_system* s, const lg_contact_params* contact, double k_d, double gamma_d,
                          double critical_separation, lg_unit_vector_fn gen, void* user, int64_t* n_bonds_out);

/* Same tables supplied explicitly (what sinter_bridge.h:139-141,185-209 hold after construction):
 * n_ordered records (i, j), both orientations present, each with the 5 connectors attached to i
 * (15 doubles) and 5 rest lengths. */
int lg_set_bonds(lg_system* s, const lg_contact_params* contact, double k_d, double gamma_d, int64_t n_ordered,
                 const int64_t* ij_pairs, const double* connectors_15, const double* rest_lengths_5);

/* The same construction split in two so that it can run where the reference runs it -- inside the
 * sinter_functor constructor, before any system exists (test/sintered.cpp:106-107 builds the functor first,
 * the system at :122-127).  lg_bond_table_create is pure host code (no CUDA device needed): it scans
 * x (3 * n doubles) for bonds exactly like sinter_bridge.h:148-211 and consumes `gen` in the same order.
 * lg_set_bond_table installs the result (with the embedded contact model, :134-135) into a system;
 * lg_bond_table_records copies out 2 * count ordered records in the layout lg_set_bonds takes. */
typedef struct lg_bond_table lg_bond_table;
int lg_bond_table_create(int64_t n, const double* x_3n, double r_part, double critical_separation,
                         lg_unit_vector_fn gen, void* user, lg_bond_table** out);
/* lg_bond_table_create on the GPU (device-side bond detection): candidate partners from the cell
 * list, the reference's predicate and construction, bond number k in lexicographic (i < j) order takes unit vector k of
 * `unit_vectors_3m` (the draws of the reference's generator, in order).  Same table as the host scan, bit for bit; a vector
 * collinear with its bond axis (the reference draws again, sinter_bridge.h:171-175) makes it fall back to the host scan.
 * device = CUDA ordinal or -1. */
int lg_bond_table_create_device(int64_t n, const double* x_3n, double r_part, double critical_separation, int64_t n_unit_vectors,
                                const double* unit_vectors_3m, int device, lg_bond_table** out);
int lg_bond_table_count(const lg_bond_table* t, int64_t* n_bonds);
int lg_bond_table_records(const lg_bond_table* t, int64_t* ij_pairs, double* connectors_15, double* rest_lengths_5);
int lg_bond_table_destroy(lg_bond_table* t);
int lg_set_bond_table(lg_system* s, const lg_contact_params* contact, double k_d, double gamma_d,
                      const lg_bond_table* t);

/* ---- model, integrator and step-handler variants ------------------------------------------------
 * Systems that use alt_sinter bonds, forward Euler, a fixed particle or the central attraction run the reference's schedule
 * literally (kernel variant 0: integrate kernel + one-thread-per-particle force kernel, exact arithmetic); the Hamaker
 * variants run in every kernel. */

/* limited_range_hamaker_functor (LG/hamaker_force/limited_range_hamaker_force.h:19-45: zero when the centre distance
 * exceeds r_cutoff, :34-35) and simplified_hamaker_functor (LG/hamaker_force/simplified_hamaker_force.h:10-34:
 * F = 2 A r / h^2, :27).  Call after lg_set_hamaker, which supplies A and h0. */
enum { LG_HAMAKER_STANDARD = 0, LG_HAMAKER_LIMITED_RANGE = 1, LG_HAMAKER_SIMPLIFIED = 2 };
int lg_set_hamaker_variant(lg_system* s, int kind, double r_cutoff);

/* The integrator template argument of granular_system: rotational_velocity_verlet_half (default) or
 * rotational_forward_euler (LT/rotational_integrator/rotational_forward_euler.h:44-63: evaluate, then
 * x += v dt + 0.5 a dt dt, v += a dt, theta += omega dt + 0.5 alpha dt dt, omega += alpha dt).  Before the first step. */
enum { LG_INTEGRATOR_VERLET_HALF = 0, LG_INTEGRATOR_FORWARD_EULER = 1 };
int lg_set_integrator(lg_system* s, int integrator);

/* fixed_particle_step_handler (Writerside/topics/Custom-step-handler.md:48-146): increment_v of particle `index` is
 * dropped, every other increment goes to the base step handler.  -1 = none. */
int lg_set_fixed_particle(lg_system* s, int64_t index);

/* central_attraction_functor (Writerside/topics/Custom-unary-force-model.md:44-121): a_i = k_a / mass * (x_i - x0),
 * alpha_i = 0; x0 is what its update_center_of_mass() last computed -- the caller passes it (the header template sums
 * get_x() in index order on the host, as the tutorial does). */
int lg_set_central_attraction(lg_system* s, double k_a, const double x0[3]);

/* alt_sinter_functor (LG/sinter_bridge/alt_sinter_bridge.h:16-176).  lg_alt_bond_table_create is its constructor's bond
 * scan (:56-89; host only): pairs with |dist - 2 r| < critical_separation (strict) in lexicographic i < j order, a bond
 * that would close a cycle is skipped.  lg_set_alt_bond_table installs the table, the embedded contact model (:33,95-96)
 * and the bond's own coefficients bond_params = {k, gamma_n, k_t, gamma_t, k_r, gamma_r, k_o, gamma_o, dt} (:18-30).
 * lg_get_alt_bond copies contact_springs[i * n + j] of a bonded pair (:162-164). */
int lg_alt_bond_table_create(int64_t n, const double* x_3n, double r_part, double critical_separation, lg_bond_table** out);
int lg_set_alt_bond_table(lg_system* s, const lg_contact_params* contact, const double bond_params[9], const lg_bond_table* t);
int lg_get_alt_bond(lg_system* s, int64_t i, int64_t j, int* bonded, double springs_9[9]);

/* Capacity of the per-particle table of simultaneously overlapping non-bonded neighbours
 * (replaces the dense N x N spring store of contact_force.h:53; default 16, kissing number is 12).
 * Must be called before the first step. Exceeding it at run time makes the step return LG_ECAPACITY. */
int lg_set_max_contacts(lg_system* s, int max_contacts);

/* ---- unary force models (LG/granular_system/unary_force_container.h:17-34) --------------------- */

/* triangular_facet<surface_contact_force_functor, surface_hamaker_functor>
 * (LG/surface_force/triangular_facet.h:18-29; LG/contact_force/surface_contact_force.h:8-30;
 * LG/hamaker_force/surface_hamaker_force.h:8-15).  surface_contact may be NULL (no wall contact);
 * has_hamaker = 0 omits the wall van der Waals term.  Facets are folded in the order added. */
int lg_add_facet(lg_system* s, const double vertices_9[9], const double v_facet[3],
                 const lg_contact_params* surface_contact, int has_hamaker, double A, double h0, int* facet_id);
/* triangular_facet::v_facet is a public data member (triangular_facet.h:69) */
int lg_facet_set_velocity(lg_system* s, int facet_id, const double v_facet[3]);
/* overwrite the three vertices (the header-template triangular_facet keeps its own host copy, moved by
 * update_positions, and pushes it before every step) */
int lg_facet_set_vertices(lg_system* s, int facet_id, const double vertices_9[9]);
/* triangular_facet::update_positions(dt) (triangular_facet.h:62-66) */
int lg_facet_update_positions(lg_system* s, int facet_id, double dt);
/* triangular_facet::get_vertices() (:57-59) */
int lg_facet_get_vertices(lg_system* s, int facet_id, double out_9[9]);

/* ---- pair loop and integrator ------------------------------------------------------------------ */

/* Selects the neighbor-list pair loop (LT/rotational_system/rotational_binary_system_neighbors_omp.h)
 * with Verlet radius r_verlet; without this call the system is the all-pairs loop
 * (LT/rotational_system/rotational_binary_system_omp.h:69-90).  As in the reference the list is
 * empty until the first lg_update_neighbor_list (:66).  r_verlet must be >= 2 * r_part (LG_EINVAL otherwise):
 * history springs of a pair are dropped when it leaves the list, which equals the reference's reset on a visited
 * non-overlapping evaluation (contact_force.h:68-71) only while every overlapping pair is listed. */
int lg_enable_neighbor_list(lg_system* s, double r_verlet);

/* update_neighbor_list() (rotational_binary_system_neighbors_omp.h:102-116): rows
 * { j != i : |x_i - x_j| < r_verlet }, ascending j, bit-identical predicate. */
int lg_update_neighbor_list(lg_system* s);

/* do_step(dt) (rotational_system.h:89-91 -> rotational_velocity_verlet_half.h:43-76). */
int lg_do_step(lg_system* s, double dt);

/* The driver loop of test/hamaker_neighbor.cpp:120-124 run without returning to the host:
 *   for k in 0..n_steps: if (rebuild_every > 0 && step % rebuild_every == 0) update_neighbor_list();
 *                        do_step(dt); if (move_facets) every facet.update_positions(dt);
 * `step` is the system's own running count of do_step calls. */
int lg_run(lg_system* s, double dt, int64_t n_steps, int64_t rebuild_every, int move_facets);

/* reset_acceleration_buffers() (rotational_system.h:83-86). */
int lg_reset_acceleration_buffers(lg_system* s);

/* Blocks until all queued work of this system has finished; reports asynchronous errors
 * (LG_ECUDA, LG_ECAPACITY). */
int lg_synchronize(lg_system* s);

/* ---- introspection used by the parity tests ---------------------------------------------------- */

int lg_get_neighbor_count(lg_system* s, int64_t* n_pairs);
/* row_ptr: n + 1 entries, cols: n_pairs entries (the private neighbor_list of
 * rotational_binary_system_neighbors_omp.h:122 flattened to CSR). */
int lg_get_neighbor_rows(lg_system* s, int64_t* row_ptr, int64_t* cols);
/* the three springs of ordered pair (i, j): contact_springs[i * n + j] of contact_force.h:100,156 */
int lg_get_contact_springs(lg_system* s, int64_t i, int64_t j, double out_9[9]);
/* returns 1 in *bonded if (i, j) is bonded and copies spring_connectors[i * n + j] /
 * spring_lengths[i * n + j] (sinter_bridge.h:139-140) */
int lg_get_bond(lg_system* s, int64_t i, int64_t j, int* bonded, double conn_15[15], double len_5[5]);

/* counters: out[0] ordered neighbor pairs, [1] ordered pairs holding a contact-spring record,
 * [2] ordered bonded pairs, [3] do_step calls so far, [4] neighbor rebuilds so far,
 * [5] kernels launched so far, [6] number of cells of the last rebuild, [7] reserved */
int lg_get_counters(lg_system* s, int64_t out_8[8]);

/* Device reductions restating test/compute_energy.cpp:7-58 and test/mass_distribution.cpp:7-26
 * (fixed-shape tree, deterministic run to run):
 *   out[0] total kinetic energy   out[1] translational KE   out[2] rotational KE
 *   out[3] |linear momentum|      out[4] |angular momentum|
 *   out[5..7] centre of mass      out[8] radius of gyration */
int lg_diagnostics(lg_system* s, double out_9[9]);

/* Arithmetic of the force laws in the tiled kernel.
 *   LG_MATH_EXACT (default): the reference's operand order, IEEE division and square root, no FMA contraction --
 *       x, v, theta, omega, a, alpha are bit-identical to the reference built with -ffp-contract=off.
 *   LG_MATH_FAST: reciprocal square roots shared between the components of a normalisation, fused multiply-adds,
 *       the four Hamaker fractions (hamaker_force.h:36-40) combined into one (-64 r^6 / (h^2 (2r+h)^3 (4r+h)^2)),
 *       multiplication by 1/mass, 1/inertia.  Per-step forces and torques stay within 1e-10 relative of the
 *       reference (the tolerance BASELINE.json states; measured ~1e-13), neighbor and contact SETS stay bit-exact
 *       (overlap decisions inside a rounding band are re-taken on the exact expression), and pairs that are in
 *       contact run the exact Luding law so their history springs evolve as in the reference. */
enum { LG_MATH_EXACT = 0, LG_MATH_FAST = 1 };
int lg_set_math_mode(lg_system* s, int mode);

/* ---- initial-condition tooling (host only, no device needed) ------------------------
 * What a libgran driver does before the time loop (main.cpp:56-93, test/hamaker_3.cpp:60-87), for the synthetic
 * inputs BASELINE.json's configs name.  Deterministic in (seed, n): std::mt19937_64 as in test/sintered.cpp:30-39. */

/* Particle-cluster sticking: particle k sits at exactly 2 r from a uniformly chosen earlier particle, in a uniform
 * random direction, rejected if within 2 r (1 + 1e-9) of any other particle.  x_out_3n: 3 * n doubles. */
int lg_gen_fractal_aggregate(int64_t n, uint64_t seed, double r_part, double* x_out_3n);
/* Random sequential addition of spheres to `volume_fraction` (<= 0.36) in a box aspect : 1 : 1 with its corner at the
 * origin; min_gap_factor in (0.5, 1] scales the excluded diameter (1 = no overlaps).  box_out (nullable) = edges. */
int lg_gen_random_packing(int64_t n, double volume_fraction, double aspect, double min_gap_factor, uint64_t seed,
                          double r_part, double* x_out_3n, double box_out[3]);

/* ---- dump pipeline ------------------------------------------------------------------------
 * The reference's writer.cpp (VTK ASCII: write_particles :51-118, write_necks :120-170, write_spring_connectors :13-49; file
 * numbering from 1 per kind, :9-11) and the tutorial's CSV dump (Writerside/topics/Custom-unary-force-model.md:339-355),
 * byte for byte.  lg_dump_* does not stall the time loop: it snapshots the needed arrays device-to-device on the system's
 * stream and returns; a second stream copies the snapshot to pinned host memory and a background thread formats and writes the
 * file.  Two snapshots may be in flight; a third call blocks until one has been written.  Files are complete after
 * lg_dumper_wait (or lg_dumper_destroy).  `dir` must exist. */
typedef struct lg_dumper lg_dumper;
int lg_dumper_create(lg_system* s, const char* dir, lg_dumper** out);
int lg_dump_particles(lg_system* s, lg_dumper* d);          /* <dir>/particles_<k>.vtk: x / r_part, orientation triad from theta */
int lg_dump_necks(lg_system* s, lg_dumper* d);              /* <dir>/necks_<k>.vtk: midpoint and normal of every bond i < j */
int lg_dump_spring_connectors(lg_system* s, lg_dumper* d);  /* <dir>/spring_connectors_<k>.vtk: 5 connectors per ordered bond */
int lg_dump_particles_csv(lg_system* s, lg_dumper* d, int64_t count); /* <dir>/particles_<count>.csv */
int lg_dumper_wait(lg_dumper* d);
int lg_dumper_destroy(lg_dumper* d);
/* The same formatter for host arrays -- the argument lists of the reference's functions (writer.h:16-23), with the dense
 * N x N tables replaced by the bond list: ij_pairs = n_bonds (i, j) pairs (either or both orientations), connectors_15 = the
 * 5 connectors of every ordered record in (i, j) order.  Pure host code. */
int lg_write_particles(const char* path, int64_t n, const double* x_3n, const double* theta_3n, double r_part);
int lg_write_necks(const char* path, int64_t n, const double* x_3n, double r_part, int64_t n_bonds, const int64_t* ij_pairs);
int lg_write_spring_connectors(const char* path, double r_part, int64_t n_records, const double* connectors_15);

/* ---- slab decomposition (no reference counterpart -- the reference is single-process) ----
 * One lg_system per rank holds that rank's owned particles and its ghosts, ordered by global particle id.  The
 * functions below move device-resident tables in and out of a system (d_* = device pointers on its device) so that
 * the launcher (libgran_b200/slab.py, or any C++ host with its own transport) can repartition and exchange halos. */

/* Materialises a pending closing force evaluation so that a, alpha and the contact
 * springs correspond to the current x, v, omega. */
int lg_flush(lg_system* s);
/* First half of the very first do_step (rotational_velocity_verlet_half.h:45-57): evaluate, then v -= a dt/2,
 * omega -= alpha dt/2 for the particles with rows.  The caller refreshes ghost v, omega and then calls lg_do_step,
 * which resumes at the second evaluation (:59). */
int lg_first_half_kick(lg_system* s, double dt);
/* lg_set_state from device memory, including a and alpha (nullable = zero) and the integrator phase:
 * kicked != 0 means the first-step half kick has already been applied to this state. */
int lg_import_state_device(lg_system* s, const double* d_x, const double* d_v, const double* d_theta,
                           const double* d_omega, const double* d_a, const double* d_alpha, int kicked);
/* n bytes: 1 = owned (gets neighbor rows, is integrated by force), 0 = ghost (rows stay empty).  NULL clears. */
int lg_set_owned_device(lg_system* s, const uint8_t* d_owned);
/* The sparse contact-spring table: S slots per particle; sgid[i*S+k] = partner index or -1, spr[(i*S+k)*9..]. */
int lg_export_slots(lg_system* s, int* S, int32_t** d_sgid, double** d_spr);
int lg_import_slots_device(lg_system* s, const int32_t* d_sgid, const double* d_spr);
/* Wall history of facet `facet_id` (surface_contact_force.h:53): springs [n][9] and active flags [n], device views /
 * device-to-device import -- wall springs migrate with their particle at a repartition like contact slots do. */
int lg_export_wall(lg_system* s, int facet_id, double** d_wspr, uint8_t** d_wact);
int lg_import_wall_device(lg_system* s, int facet_id, const double* d_wspr, const uint8_t* d_wact);
/* The sums of lg_diagnostics over the particles this system OWNS (ghosts excluded), left on the device for a caller that
 * reduces them across ranks (queued on the system's stream, no synchronisation).  *d_out = 12 doubles:
 * pass 0 fills [0..10] = {translational KE, rotational KE, sum v (3), sum (I omega + x cross m v) (3), sum x (3)};
 * pass 1 (d_com = device pointer to the global centre of mass) fills [11] = sum |x - com|^2. */
int lg_diagnostics_partial(lg_system* s, int pass, const double* d_com, double** d_out);
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The sums of lg_diagnostics_partial
 * over caller-owned DEVICE arrays d_x, d_v, d_w ([n][3] doubles, pass 1 reads only d_x) into the caller's d_out_12
 * (12 device doubles, same slots as above; pass 0 writes [0..10], pass 1 writes [11] and needs d_com = 3 device
 * doubles).  Queued on `stream` (a cudaStream_t, NULL = the default stream) of the CURRENT device, no synchronisation.
 * It allocates nothing -- no cudaMalloc / cudaFree, no lg_system: the per-block partial sums live in a 48 KiB
 * file-scope __device__ array that every GPU loading the library holds until exit.  Limits and caveats:
 *   - n <= lg_diag_sums_static_max_particles() (= 67 108 864); larger n returns LG_ESIZE with a message;
 *   - every call on a device shares that array (it keeps the last call's contents; the part a call reads is cleared
 *     and rewritten first), so calls that OVERLAP on the GPU -- different streams, or two host threads with their own
 *     streams -- race on it; calls on one stream are ordered and safe. */
int lg_diag_sums_static(int64_t n, const double* d_x, const double* d_v, const double* d_w, double mass, double inertia,
                        int pass, const double* d_com, double* d_out_12, void* stream);
int lg_diag_sums_static_max_particles(void);
/* The bond CSR (rows ascending in partner index): bptr[n+1], bj[nb], connectors[nb][15], rest lengths[nb][5];
 * brev[b] = record index of the mirror (j, i), or b itself when the mirror is not held locally. */
int lg_export_bonds(lg_system* s, int64_t* n_ordered, uint32_t** d_bptr, int32_t** d_bj, double** d_bconn,
                    double** d_blen);
int lg_import_bonds_device(lg_system* s, const lg_contact_params* contact, double k_d, double gamma_d,
                           int64_t n_ordered, const uint32_t* d_bptr, const int32_t* d_bj, const int32_t* d_brev,
                           const double* d_bconn, const double* d_blen);
/* Halo staging on the system's stream: gathers x, v, omega of particles d_idx[0..n_part) (9 doubles each) followed
 * by the connector blocks of bond records d_rec[0..n_rec) (15 doubles each) into d_out; unpack scatters the same
 * layout into the ghost slots. */
int lg_halo_pack(lg_system* s, int n_part, const int32_t* d_idx, int n_rec, const int32_t* d_rec, double* d_out);
int lg_halo_unpack(lg_system* s, int n_part, const int32_t* d_idx, int n_rec, const int32_t* d_rec,
                   const double* d_in);

/* Peer-to-peer halo exchange over NVLink, without a host transport in the time loop.  Each rank allocates ONE
 * receive buffer (two parity halves, a flag per peer and half) and exports it with cudaIpcGetMemHandle; a neighbour
 * maps it (lg_halo_connect) and its pack kernel stores x, v, omega and connector blocks straight into it, then
 * publishes a sequence number with a system-scope release store.  lg_halo_exchange queues, on the system's stream,
 * the push to every peer and a pull that waits for the peers' sequence numbers (bounded: ERR -> LG_ECUDA at the next
 * synchronisation) and scatters into the ghost slots.  The launcher exchanges the 64-byte handles and offsets between
 * ranks (libgran_b200/slab.py does it over torch.distributed).  Message layout as lg_halo_pack.
 * Order of teardown across ranks: every rank lg_halo_disconnect, barrier, then lg_halo_destroy. */
typedef struct lg_halo lg_halo;
int lg_halo_create(lg_system* s, int n_peers, const int64_t* recv_doubles, lg_halo** out);
int lg_halo_set_lists(lg_halo* h, int peer, int n_send_part, const int32_t* d_send_idx, int n_send_rec,
                      const int32_t* d_send_rec, int n_recv_part, const int32_t* d_recv_idx, int n_recv_rec,
                      const int32_t* d_recv_rec);
int lg_halo_export(lg_halo* h, int peer, unsigned char handle[64], int64_t offs[4]);
int lg_halo_connect(lg_halo* h, int peer, const unsigned char handle[64], const int64_t offs[4]);
int lg_halo_exchange(lg_system* s, lg_halo* h);
/* n_steps x (lg_do_step, [every facet.update_positions(dt)], lg_halo_exchange) without returning to the caller (no neighbor
 * rebuild inside).  Fused launches of step_rows2_kernel store the halo themselves (the CTA that finishes a tile writes its
 * halo items into the neighbours' buffers, the last CTA publishes the sequence number): no separate push kernel. */
int lg_run_slab(lg_system* s, lg_halo* h, double dt, int64_t n_steps, int move_facets);
int lg_halo_disconnect(lg_halo* h);
int lg_halo_destroy(lg_halo* h);

/* The CUDA stream (cudaStream_t) all work of this system is queued on, so that callers can time it with
 * events or order their own copies against it. */
int lg_get_stream(lg_system* s, void** stream);

/* Per-kernel device timing for roofline reports: while enabled, every launch of the pair-force kernel is
 * bracketed by CUDA events on the system's stream.  lg_get_profile synchronises and returns the summed
 * duration in milliseconds and the number of launches since profiling was enabled. */
int lg_set_profiling(lg_system* s, int enabled);
int lg_get_profile(lg_system* s, double* force_ms_total, int64_t* force_launches);

/* Selects the force-kernel organisation: 0 = one thread per particle, unfused (validation; the reference's schedule
 * literally), 1 = production (default): the fused step kernels, pair-tile kernel or -- in LG_MATH_FAST with short
 * rows -- the row kernel, 2 = as 1 but the row kernel whatever the row length (LG_MATH_FAST only; coverage tests).
 * Variants 0 and 1 give bit-identical results in LG_MATH_EXACT. */
int lg_set_kernel_variant(lg_system* s, int variant);
/* Tile shape of step_rows2_kernel (0 = 96 particles x 2 lanes, 1 = 64 x 4, 2 = 32 x 8, 3 = 48 x 2: shape 0's summation order on half-size tiles, for
 * the small per-rank systems of slab runs).  By default a deterministic
 * rule on the mean row length picks it; lg_set_autotune(1) measures the three shapes on the first fused launches (CUDA events,
 * three launches each) and keeps the fastest -- results then depend on the shape chosen (summation order inside a row),
 * within the LG_MATH_FAST tolerance.  lg_get_rows_shape reports the shape in use. */
/* lg_set_rows_shape pins the shape (-1 returns to the rule); lg_rows_shape_rule is the rule itself (returns the shape index)
 * -- the slab driver evaluates it on GLOBAL statistics and pins the result on every rank, so that a decomposed run sums each
 * row in the same order as one GPU. */
int lg_set_rows_shape(lg_system* s, int shape);
int lg_rows_shape_rule(double mean_row_length, int64_t n, int has_bonds);
int lg_set_autotune(lg_system* s, int enabled);
int lg_get_rows_shape(lg_system* s, int* shape);
/* Name of the kernel the most recent force evaluation launched ("step_rows_kernel", "step_tiled_kernel",
 * "force_simple_kernel"; "" before the first), copied into out[cap]. */
int lg_last_step_kernel(lg_system* s, char* out, int cap);

#ifdef __cplusplus
}
#endif
#endif /* LGB200_H */
