/*
 * dem_config.h -- plain-C description of one DEM set-up, shared by the two CPU
 * checkers under oracle/ (TEST INFRASTRUCTURE ONLY -- never linked into the
 * product library):
 *
 *   oracle/ref_harness.cpp   the unmodified reference headers, driven in lock-step
 *   oracle/dem_oracle.cpp    the sparse-state CPU restatement of the same path
 *
 * The field order of `contact[]` follows the constructor argument order of the
 * reference functor (include/libgran/contact_force/contact_force.h:10-30):
 *   k, gamma_n, k_t, gamma_t, mu_s, phi_d, k_r, gamma_r, mu_r, phi_r,
 *   k_o, gamma_o, mu_o, phi_o, dt
 */
#ifndef LGB200_ORACLE_DEM_CONFIG_H
#define LGB200_ORACLE_DEM_CONFIG_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    DEM_MODEL_CONTACT = 1, /* contact_force_functor            */
    DEM_MODEL_HAMAKER = 2, /* hamaker_functor                  */
    DEM_MODEL_SINTER  = 4, /* sinter_functor (embeds contact)  */
    DEM_MODEL_FACET   = 8, /* one triangular_facet<surface contact, surface hamaker> */
    DEM_MODEL_ALT_SINTER = 16 /* alt_sinter_functor (embeds contact); excludes CONTACT and SINTER */
};

enum { DEM_HAMAKER_STANDARD = 0,      /* hamaker_functor                                    */
       DEM_HAMAKER_LIMITED_RANGE = 1, /* limited_range_hamaker_functor (hamaker_cutoff)     */
       DEM_HAMAKER_SIMPLIFIED = 2 };  /* simplified_hamaker_functor                         */
enum { DEM_INTEGRATOR_VERLET_HALF = 0, DEM_INTEGRATOR_FORWARD_EULER = 1 };

enum { DEM_X = 0, DEM_V = 1, DEM_A = 2, DEM_THETA = 3, DEM_OMEGA = 4, DEM_ALPHA = 5 };

typedef struct dem_config {
    int64_t n;
    double r_part, mass, inertia;
    int32_t model_mask;
    int32_t use_neighbor_list;
    double r_verlet;
    double contact[15];
    double hamaker_A, hamaker_h0;
    double k_d, gamma_d, critical_separation;
    uint64_t rng_seed;            /* mt19937_64 seed of the unit-vector generator (test/sintered.cpp:29-39) */
    double facet_vertices[9];     /* three vertices, xyz each */
    double facet_v[3];
    double surf_contact[15];
    double surf_A, surf_h0;
    /* ---- SURVEY.md section 8f variants ---- */
    double alt_bond[8];           /* alt_sinter_functor's own k, gamma_n, k_t, gamma_t, k_r, gamma_r, k_o, gamma_o
                                     (alt_sinter_bridge.h:18-26); its embedded contact model takes contact[] */
    int32_t hamaker_kind;         /* DEM_HAMAKER_* */
    int32_t integrator;           /* DEM_INTEGRATOR_* */
    double hamaker_cutoff;        /* limited_range_hamaker_functor's r_cutoff (centre distance) */
    int64_t fixed_particle;       /* fixed_particle_step_handler's index, -1 = none (Custom-step-handler.md:48-146) */
} dem_config;

#ifdef __cplusplus
}
#endif
#endif
