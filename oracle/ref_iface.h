// ref_iface.h -- TEST INFRASTRUCTURE ONLY. Type-erased view of one reference system instance
// (see ref_harness.cpp) and the C entry points exported by oracle/_ref/libref_oracle.so.
#ifndef LGB200_ORACLE_REF_IFACE_H
#define LGB200_ORACLE_REF_IFACE_H

#include <stdint.h>
#include "dem_config.h"

#ifdef __cplusplus
struct RefSystem {
    virtual ~RefSystem() = default;
    virtual void update_neighbor_list() = 0;
    virtual void do_step(double dt) = 0;
    virtual void run(double dt, int64_t n_steps, int64_t rebuild_every) = 0;
    virtual void facet_update_positions(double dt) = 0;
    virtual void get_field(int field, double* out) const = 0;
    virtual int64_t n_pairs() const = 0;
    virtual void get_rows(int64_t* row_ptr, int64_t* cols) const = 0;
    virtual void get_contact_springs(int64_t i, int64_t j, double* out9) const = 0;
    virtual int get_bond(int64_t i, int64_t j, double* conn15, double* len5) const = 0;
    virtual void get_facet_vertices(double* out9) const = 0;
};

RefSystem* ref_make_binary_0(const dem_config*, const double*, const double*, const double*, const double*);
RefSystem* ref_make_binary_1(const dem_config*, const double*, const double*, const double*, const double*);
RefSystem* ref_make_binary_2(const dem_config*, const double*, const double*, const double*, const double*);
RefSystem* ref_make_binary_3(const dem_config*, const double*, const double*, const double*, const double*);
RefSystem* ref_make_binary_4(const dem_config*, const double*, const double*, const double*, const double*);
RefSystem* ref_make_binary_5(const dem_config*, const double*, const double*, const double*, const double*);
RefSystem* ref_make_binary_6(const dem_config*, const double*, const double*, const double*, const double*);
RefSystem* ref_make_binary_7(const dem_config*, const double*, const double*, const double*, const double*);
RefSystem* ref_make_binary_fe_1(const dem_config*, const double*, const double*, const double*, const double*);
RefSystem* ref_make_binary_fe_3(const dem_config*, const double*, const double*, const double*, const double*);

extern "C" {
#endif

void* ref_create(const dem_config* cfg, const double* x0, const double* v0, const double* theta0, const double* omega0);
void ref_destroy(void* h);
void ref_update_neighbor_list(void* h);
void ref_do_step(void* h, double dt);
/* n_steps of: [if step % rebuild_every == 0: update_neighbor_list()] do_step(dt)  (test/hamaker_neighbor.cpp:120-124) */
void ref_run(void* h, double dt, int64_t n_steps, int64_t rebuild_every);
void ref_facet_update_positions(void* h, double dt);
void ref_get_field(void* h, int field, double* out_3n);
int64_t ref_n_pairs(void* h);
void ref_get_rows(void* h, int64_t* row_ptr_np1, int64_t* cols);
void ref_get_contact_springs(void* h, int64_t i, int64_t j, double* out9);
int ref_get_bond(void* h, int64_t i, int64_t j, double* conn15, double* len5);
void ref_get_facet_vertices(void* h, double* out9);
/* diagnostics: test/compute_energy.cpp:7-58, test/mass_distribution.cpp:7-26 compiled as they are */
double ref_total_kinetic_energy(const double* v, const double* omega, int64_t n, double mass, double inertia);
double ref_linear_momentum(const double* v, int64_t n, double mass);
double ref_angular_momentum(const double* x, const double* v, const double* omega, int64_t n, double mass, double inertia);
void ref_center_of_mass(const double* x, int64_t n, double* out3);
double ref_radius_of_gyration(const double* x, int64_t n);
int ref_max_threads(void);
void ref_set_threads(int n);

#ifdef __cplusplus
}
#endif
#endif
