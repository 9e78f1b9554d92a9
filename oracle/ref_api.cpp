// ref_api.cpp -- TEST INFRASTRUCTURE ONLY. C entry points of oracle/_ref/libref_oracle.so: dispatch
// to the per-combination translation units of ref_harness.cpp, plus the reference's own diagnostics
// (test/compute_energy.cpp, test/mass_distribution.cpp are compiled from /root/reference as they
// are and linked in; see oracle/Makefile).
#include <cstdint>
#include <vector>
#include <omp.h>
#include <Eigen/Eigen>

#include "ref_iface.h"

// declarations of the reference's diagnostics (test/compute_energy.h, test/mass_distribution.h)
double compute_total_kinetic_energy(std::vector<Eigen::Vector3d> const& v, std::vector<Eigen::Vector3d> const& omega,
                                    double mass, double inertia);
double compute_linear_momentum(std::vector<Eigen::Vector3d> const& vs, double mass);
double compute_angular_momentum(std::vector<Eigen::Vector3d> const& xs, std::vector<Eigen::Vector3d> const& vs, double m,
                                std::vector<Eigen::Vector3d> const& omegas, double inertia);
Eigen::Vector3d center_of_mass(std::vector<Eigen::Vector3d> const& x);
double radius_of_gyration(std::vector<Eigen::Vector3d> const& x);

static std::vector<Eigen::Vector3d> as_vec(const double* p, int64_t n) {
    std::vector<Eigen::Vector3d> out((size_t)n);
    for (int64_t i = 0; i < n; i++) out[(size_t)i] = Eigen::Vector3d(p[3 * i], p[3 * i + 1], p[3 * i + 2]);
    return out;
}

extern "C" {

void* ref_create(const dem_config* c, const double* x0, const double* v0, const double* th0, const double* om0) {
    const int bin = c->model_mask & (DEM_MODEL_CONTACT | DEM_MODEL_HAMAKER | DEM_MODEL_SINTER | DEM_MODEL_ALT_SINTER);
    try {
        if (c->fixed_particle >= 0) return nullptr;  // the tutorial's step handler is not part of the reference headers
        if (c->integrator == DEM_INTEGRATOR_FORWARD_EULER) {
            if (c->hamaker_kind != DEM_HAMAKER_STANDARD) return nullptr;
            if (bin == (DEM_MODEL_CONTACT | DEM_MODEL_HAMAKER)) return ref_make_binary_fe_1(c, x0, v0, th0, om0);
            if (bin == (DEM_MODEL_SINTER | DEM_MODEL_HAMAKER)) return ref_make_binary_fe_3(c, x0, v0, th0, om0);
            return nullptr;
        }
        if (bin == DEM_MODEL_ALT_SINTER) return ref_make_binary_5(c, x0, v0, th0, om0);
        if (c->hamaker_kind != DEM_HAMAKER_STANDARD) {
            if (bin != (DEM_MODEL_CONTACT | DEM_MODEL_HAMAKER)) return nullptr;
            return c->hamaker_kind == DEM_HAMAKER_LIMITED_RANGE ? ref_make_binary_6(c, x0, v0, th0, om0)
                                                                : ref_make_binary_7(c, x0, v0, th0, om0);
        }
        switch (bin) {
            case DEM_MODEL_CONTACT: return ref_make_binary_0(c, x0, v0, th0, om0);
            case DEM_MODEL_CONTACT | DEM_MODEL_HAMAKER: return ref_make_binary_1(c, x0, v0, th0, om0);
            case DEM_MODEL_SINTER: return ref_make_binary_2(c, x0, v0, th0, om0);
            case DEM_MODEL_SINTER | DEM_MODEL_HAMAKER: return ref_make_binary_3(c, x0, v0, th0, om0);
            case DEM_MODEL_HAMAKER: return ref_make_binary_4(c, x0, v0, th0, om0);
            default: return nullptr;
        }
    } catch (...) {
        return nullptr;
    }
}
void ref_destroy(void* h) { delete static_cast<RefSystem*>(h); }
void ref_update_neighbor_list(void* h) { static_cast<RefSystem*>(h)->update_neighbor_list(); }
void ref_do_step(void* h, double dt) { static_cast<RefSystem*>(h)->do_step(dt); }
void ref_run(void* h, double dt, int64_t n, int64_t every) { static_cast<RefSystem*>(h)->run(dt, n, every); }
void ref_facet_update_positions(void* h, double dt) { static_cast<RefSystem*>(h)->facet_update_positions(dt); }
void ref_get_field(void* h, int f, double* out) { static_cast<RefSystem*>(h)->get_field(f, out); }
int64_t ref_n_pairs(void* h) { return static_cast<RefSystem*>(h)->n_pairs(); }
void ref_get_rows(void* h, int64_t* p, int64_t* c) { static_cast<RefSystem*>(h)->get_rows(p, c); }
void ref_get_contact_springs(void* h, int64_t i, int64_t j, double* o) { static_cast<RefSystem*>(h)->get_contact_springs(i, j, o); }
int ref_get_bond(void* h, int64_t i, int64_t j, double* c, double* l) { return static_cast<RefSystem*>(h)->get_bond(i, j, c, l); }
void ref_get_facet_vertices(void* h, double* o) { static_cast<RefSystem*>(h)->get_facet_vertices(o); }

double ref_total_kinetic_energy(const double* v, const double* w, int64_t n, double m, double I) {
    return compute_total_kinetic_energy(as_vec(v, n), as_vec(w, n), m, I);
}
double ref_linear_momentum(const double* v, int64_t n, double m) { return compute_linear_momentum(as_vec(v, n), m); }
double ref_angular_momentum(const double* x, const double* v, const double* w, int64_t n, double m, double I) {
    return compute_angular_momentum(as_vec(x, n), as_vec(v, n), m, as_vec(w, n), I);
}
void ref_center_of_mass(const double* x, int64_t n, double* out3) {
    Eigen::Vector3d c = center_of_mass(as_vec(x, n));
    out3[0] = c[0]; out3[1] = c[1]; out3[2] = c[2];
}
double ref_radius_of_gyration(const double* x, int64_t n) { return radius_of_gyration(as_vec(x, n)); }
int ref_max_threads(void) { return omp_get_max_threads(); }
void ref_set_threads(int n) { omp_set_num_threads(n); }

}  // extern "C"
