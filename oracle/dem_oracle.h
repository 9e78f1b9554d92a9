/*
 * dem_oracle.h -- C entry points of oracle/libdem_oracle.so, the CPU restatement of libgran's
 * per-timestep hot path with sparse, exposed pair state.  TEST INFRASTRUCTURE ONLY: the product
 * library never links or loads it.  Parity is PINNED: tests/test_oracle_vs_reference.py proves the
 * restatement bit-identical to the unmodified reference (oracle/_ref/libref_oracle.so, built with
 * -ffp-contract=off) on every model combination, and tests/test_golden.py checks it against the
 * committed reference outputs under tests/golden/ where /root/reference is not available.
 */
#ifndef LGB200_ORACLE_DEM_ORACLE_H
#define LGB200_ORACLE_DEM_ORACLE_H

#include <stdint.h>
#include "dem_config.h"

#ifdef __cplusplus
extern "C" {
#endif

void* orc_create(const dem_config* cfg, const double* x0, const double* v0, const double* theta0, const double* omega0);
void orc_destroy(void* h);
void orc_update_neighbor_list(void* h);
void orc_do_step(void* h, double dt);
void orc_run(void* h, double dt, int64_t n_steps, int64_t rebuild_every);
void orc_facet_update_positions(void* h, double dt);
void orc_get_field(void* h, int field, double* out_3n);
int64_t orc_n_pairs(void* h);
void orc_get_rows(void* h, int64_t* row_ptr_np1, int64_t* cols);
void orc_get_contact_springs(void* h, int64_t i, int64_t j, double* out9);
int orc_get_bond(void* h, int64_t i, int64_t j, double* conn15, double* len5);
void orc_get_facet_vertices(void* h, double* out9);
int64_t orc_n_bonds(void* h);      /* ordered bonded pairs */
int64_t orc_n_contacts(void* h);   /* ordered pairs currently holding a spring record */
void orc_set_brute_force_neighbors(void* h, int on);
void orc_unit_vectors(uint64_t seed, int64_t count, double* out_3count);
/* out8 = { KE, |P|, |L|, com.x, com.y, com.z, Rg, 0 } */
void orc_diagnostics(const double* x, const double* v, const double* omega, int64_t n, double mass, double inertia,
                     double* out8);
int orc_max_threads(void);
void orc_set_threads(int n);

#ifdef __cplusplus
}
#endif
#endif
