// dem_oracle.cpp -- CPU restatement of libgran's per-timestep DEM hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under libgran_b200/ or include/ links, loads or calls this file;
// it exists so that tests/ can check the CUDA path at sizes where the reference's dense N x N tables
// (72 B * N^2 contact springs, contact_force.h:53; 232 B * N^2 with sinter bridges,
// sinter_bridge.h:139-141) do not fit in memory.  Parity is PINNED: the restatement is proven
// bit-identical to the unmodified reference (oracle/_ref/libref_oracle.so) by
// tests/test_oracle.py (live lock-step against the reference and against the committed golden vectors).
//
// Compile with -ffp-contract=off (oracle/Makefile): every expression below spells out the
// evaluation order of the corresponding Eigen expression in the reference, including
//   * 3-term reductions (dot / squaredNorm) as (a0*b0 + a1*b1) + a2*b2 -- Eigen's packet-of-two
//     linear-vectorised redux for a 3-vector (deps/eigen/Eigen/src/Core/Redux.h:360-381),
//   * normalized(): v / sqrt(v.v) component-wise, v unchanged if v.v == 0 (Dot.h:113-123),
//   * cross(): textbook formula (Geometry/OrthoMethods.h).
//
// Restated reference code (paths relative to /root/reference):
//   include/libgran/contact_force/contact_force.h:57-147           -> contact_pair, shear_contribution
//   include/libgran/hamaker_force/hamaker_force.h:28-43             -> hamaker_pair
//   include/libgran/sinter_bridge/sinter_bridge.h:24-42,55-75       -> integrate (connector update)
//   include/libgran/sinter_bridge/sinter_bridge.h:148-211           -> build_bonds
//   include/libgran/sinter_bridge/sinter_bridge.h:214-259           -> sinter_pair, bonded_pair
//   include/libgran/surface_force/triangular_facet.h:18-137         -> Facet
//   include/libgran/contact_force/surface_contact_force.h:57-145    -> surface_contact
//   include/libgran/hamaker_force/surface_hamaker_force.h:18-30     -> surface_hamaker
//   include/libgran/granular_system/{binary,unary}_force_container.h:20-32 -> fold order
//   deps/libtimestep/.../rotational_binary_system_omp.h:60-91       -> evaluate (all pairs)
//   deps/libtimestep/.../rotational_binary_system_neighbors_omp.h:69-116 -> evaluate, update_neighbor_list
//   deps/libtimestep/.../rotational_velocity_verlet_half.h:43-76    -> do_step
//   test/compute_energy.cpp:7-58, test/mass_distribution.cpp:7-26   -> orc_diagnostics
// SURVEY.md section 8f variants:
//   include/libgran/sinter_bridge/alt_sinter_bridge.h:56-156        -> build_alt_bonds, alt_pair
//   include/libgran/hamaker_force/limited_range_hamaker_force.h:29-47, simplified_hamaker_force.h:18-30 -> hamaker_pair
//   deps/libtimestep/.../rotational_forward_euler.h:44-63           -> do_step_forward_euler
//   Writerside/topics/Custom-step-handler.md:48-146 (fixed_particle_step_handler: increment_v skipped) -> fixed_particle
#include "dem_oracle.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <random>
#include <unordered_map>
#include <vector>

#include <omp.h>

namespace {

struct V3 {
    double c[3];
    double& operator[](int k) { return c[k]; }
    double operator[](int k) const { return c[k]; }
};

inline V3 mk(double a, double b, double c) { return V3{{a, b, c}}; }
inline V3 zero3() { return mk(0.0, 0.0, 0.0); }
inline V3 add(V3 a, V3 b) { return mk(a[0] + b[0], a[1] + b[1], a[2] + b[2]); }
inline V3 sub(V3 a, V3 b) { return mk(a[0] - b[0], a[1] - b[1], a[2] - b[2]); }
inline V3 neg(V3 a) { return mk(-a[0], -a[1], -a[2]); }
inline V3 scale(double s, V3 a) { return mk(s * a[0], s * a[1], s * a[2]); }
inline V3 divs(V3 a, double s) { return mk(a[0] / s, a[1] / s, a[2] / s); }
inline double dot(V3 a, V3 b) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }
inline double norm(V3 a) { return std::sqrt(dot(a, a)); }
inline V3 normalized(V3 a) {
    double z = dot(a, a);
    if (z > 0.0) return divs(a, std::sqrt(z));
    return a;
}
inline V3 cross(V3 a, V3 b) {
    return mk(a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]);
}
inline double square(double x) { return x * x; }
inline double cube(double x) { return x * x * x; }

struct Accel {
    V3 a, alpha;
};
inline Accel operator+(Accel const& l, Accel const& r) { return Accel{add(l.a, r.a), add(l.alpha, r.alpha)}; }

struct ContactParams {
    double k, gamma_n, k_t, gamma_t, mu_s, phi_d, k_r, gamma_r, mu_r, phi_r, k_o, gamma_o, mu_o, phi_o, dt;
    void load(const double* p) {
        k = p[0]; gamma_n = p[1]; k_t = p[2]; gamma_t = p[3]; mu_s = p[4]; phi_d = p[5]; k_r = p[6]; gamma_r = p[7];
        mu_r = p[8]; phi_r = p[9]; k_o = p[10]; gamma_o = p[11]; mu_o = p[12]; phi_o = p[13]; dt = p[14];
    }
};

struct Springs {
    V3 xi[3];  // 0 sliding/sticking, 1 rolling, 2 torsion
};

struct SpringRec {
    int64_t j;
    Springs s;
};

// contact_force.h:105-147 (identical in surface_contact_force.h:103-145)
V3 shear_contribution(V3& xi, V3 n, double stiffness, double damping, double normal_force, double mu_static,
                      double phi_dynamic, V3 rel_v, double dt) {
    double mu_dynamic = mu_static * phi_dynamic;
    V3 xi_new = sub(xi, scale(dot(xi, n), n));
    if (norm(xi_new) > 0.0) {
        xi_new = divs(xi_new, std::sqrt(dot(xi_new, xi_new)));  // normalize()
        xi_new = scale(norm(xi), xi_new);                        // *= xi.norm()  (component * scalar)
    }
    xi = xi_new;
    V3 f_0 = sub(scale(-stiffness, xi), scale(damping, rel_v));
    double static_friction = mu_static * normal_force;
    double dynamic_friction = mu_dynamic * normal_force;
    V3 t;
    if (norm(f_0) > 0.0)
        t = normalized(f_0);
    else
        t = zero3();
    double f_selected;
    if (norm(f_0) <= static_friction) {
        xi = add(xi, scale(dt, rel_v));
        f_selected = static_friction;
    } else {
        xi = scale(-1.0 / stiffness, add(scale(dynamic_friction, t), scale(damping, rel_v)));
        f_selected = dynamic_friction;
    }
    return scale(std::min(f_selected, norm(f_0)), t);
}

struct Facet {
    V3 vtx[3];
    V3 v_facet;
    V3 unit_normal;
    double area;

    // triangular_facet.h:18-29
    void init(const double* verts, const double* vf) {
        for (int k = 0; k < 3; k++) vtx[k] = mk(verts[3 * k], verts[3 * k + 1], verts[3 * k + 2]);
        v_facet = mk(vf[0], vf[1], vf[2]);
        V3 u = sub(vtx[1], vtx[0]);
        V3 v = sub(vtx[2], vtx[0]);
        area = 1.0 / 2.0 * norm(cross(u, v));
        unit_normal = normalized(cross(u, v));
    }
    // :62-66
    void update_positions(double dt) {
        for (int k = 0; k < 3; k++) vtx[k] = add(vtx[k], scale(dt, v_facet));
    }
    // :73-78
    V3 closest_point_on_plane(V3 p) const {
        V3 offset = scale(dot(vtx[0], unit_normal), unit_normal);
        V3 projection = sub(p, scale(dot(p, unit_normal), unit_normal));
        return add(projection, offset);
    }
    // :80-90
    bool is_point_on_triangle(V3 p) const {
        double a = 0.0;
        V3 pa = sub(vtx[0], p), pb = sub(vtx[1], p), pc = sub(vtx[2], p);
        a += 1.0 / 2.0 * norm(cross(pa, pb));
        a += 1.0 / 2.0 * norm(cross(pb, pc));
        a += 1.0 / 2.0 * norm(cross(pc, pa));
        return !(a > area);
    }
    // :92-107
    V3 closest_point_on_edge(int i, int j, V3 p) const {
        V3 line = sub(vtx[j], vtx[i]);
        V3 unit = normalized(line);
        V3 tp = sub(p, vtx[i]);
        double len = norm(line);
        double proj = dot(tp, unit);
        if (proj < 0.0) return vtx[i];
        if (proj > len) return vtx[j];
        return add(vtx[i], scale(proj, unit));
    }
    // :109-137
    V3 closest_point_on_triangle(V3 p) const {
        V3 cp = closest_point_on_plane(p);
        if (is_point_on_triangle(cp)) return cp;
        V3 e1 = closest_point_on_edge(0, 1, p);
        V3 e2 = closest_point_on_edge(1, 2, p);
        V3 e3 = closest_point_on_edge(2, 0, p);
        V3 d1 = sub(p, e1), d2 = sub(p, e2), d3 = sub(p, e3);
        double s1 = dot(d1, d1), s2 = dot(d2, d2), s3 = dot(d3, d3);
        if (s1 < s2) {
            if (s1 < s3) return e1;
        } else {
            if (s2 < s3) return e2;
        }
        return e3;
    }
};

struct Bond {
    int64_t j;      // partner
    int64_t rev;    // index of the (j, i) record in bonds[j]
    V3 conn[5];     // connectors attached to i
    double len[5];  // rest lengths
};

struct Oracle {
    dem_config cfg;
    int64_t n;
    double r, mass, inertia;
    ContactParams cp, sp;
    bool has_contact, has_hamaker, has_sinter, has_facet, has_alt, neigh;
    std::vector<std::vector<SpringRec>> alt_springs;  // per-i springs of the alt-bonded ordered pairs (i, j), ascending j
    bool brute_force = false;
    std::vector<V3> x, v, a, theta, omega, alpha;
    std::vector<std::vector<int64_t>> rows;      // neighbor rows (ascending j); all-pairs mode leaves them empty
    std::vector<std::vector<SpringRec>> springs; // per-i sparse spring records
    std::vector<std::vector<Bond>> bonds;        // per-i bonds, ascending j
    std::vector<Springs> wall_springs;           // per particle (surface_contact_force.h:53)
    std::vector<char> wall_active;
    Facet facet;
    bool velocities_initialized = false;
    int64_t step_count = 0;

    Oracle(const dem_config& c, const double* x0, const double* v0, const double* th0, const double* om0)
        : cfg(c), n(c.n), r(c.r_part), mass(c.mass), inertia(c.inertia) {
        cp.load(c.contact);
        sp.load(c.surf_contact);
        has_contact = (c.model_mask & DEM_MODEL_CONTACT) != 0;
        has_hamaker = (c.model_mask & DEM_MODEL_HAMAKER) != 0;
        has_sinter = (c.model_mask & DEM_MODEL_SINTER) != 0;
        has_facet = (c.model_mask & DEM_MODEL_FACET) != 0;
        has_alt = (c.model_mask & DEM_MODEL_ALT_SINTER) != 0;
        neigh = c.use_neighbor_list != 0;
        auto load = [&](std::vector<V3>& dst, const double* src) {
            dst.resize((size_t)n);
            for (int64_t i = 0; i < n; i++) dst[(size_t)i] = mk(src[3 * i], src[3 * i + 1], src[3 * i + 2]);
        };
        load(x, x0); load(v, v0); load(theta, th0); load(omega, om0);
        a.assign((size_t)n, zero3());
        alpha.assign((size_t)n, zero3());
        rows.resize((size_t)n);
        springs.resize((size_t)n);
        bonds.resize((size_t)n);
        if (has_facet) {
            facet.init(c.facet_vertices, c.facet_v);
            Springs z{{zero3(), zero3(), zero3()}};
            wall_springs.assign((size_t)n, z);
        }
        if (has_sinter) build_bonds();
        if (has_alt) build_alt_bonds();
    }

    // ---- alt_sinter_bridge.h:56-89: bonds with |dist - 2 r| < critical_separation (strict), i < j in lexicographic
    //      order, a bond that would close a cycle is skipped (:70-73; the subset relabelling of :83-87 is a union-find
    //      "same component?" test) --------------------------------------------------------------------------------------
    void build_alt_bonds() {
        alt_springs.resize((size_t)n);
        const double crit = cfg.critical_separation;
        std::vector<std::vector<int64_t>> cand((size_t)n);
        if (n <= 2048 || brute_force) {
            for (int64_t i = 0; i < n - 1; i++)
                for (int64_t j = i + 1; j < n; j++) cand[(size_t)i].push_back(j);
        } else {
            for_each_near(2.0 * r + crit, [&](int64_t i, int64_t j) {
                if (j > i) cand[(size_t)i].push_back(j);
            });
            for (auto& row : cand) std::sort(row.begin(), row.end());
        }
        std::vector<int64_t> parent((size_t)n);
        for (int64_t i = 0; i < n; i++) parent[(size_t)i] = i;
        auto find = [&](int64_t a0) {
            int64_t root = a0;
            while (parent[(size_t)root] != root) root = parent[(size_t)root];
            while (parent[(size_t)a0] != root) {
                int64_t nx = parent[(size_t)a0];
                parent[(size_t)a0] = root;
                a0 = nx;
            }
            return root;
        };
        for (int64_t i = 0; i < n - 1; i++)
            for (int64_t j : cand[(size_t)i]) {
                if (!(std::fabs(norm(sub(x[(size_t)i], x[(size_t)j])) - 2.0 * r) < crit)) continue;
                const int64_t ri = find(i), rj = find(j);
                if (ri == rj) continue;  // "preventing neck insertion to avoid a cycle"
                parent[(size_t)rj] = ri;
                alt_springs[(size_t)i].push_back(SpringRec{j, Springs{{zero3(), zero3(), zero3()}}});
                alt_springs[(size_t)j].push_back(SpringRec{i, Springs{{zero3(), zero3(), zero3()}}});
            }
        for (auto& row : alt_springs) std::sort(row.begin(), row.end(), [](SpringRec const& p, SpringRec const& q) { return p.j < q.j; });
    }

    // alt_sinter_bridge.h:141-160: spring rotated into the tangent plane keeping its length, test force, spring += v dt
    V3 alt_shear(V3& xi, V3 nrm, double stiffness, double damping, V3 rel_v) const {
        V3 xi_new = sub(xi, scale(dot(xi, nrm), nrm));
        if (norm(xi_new) > 0.0) {
            xi_new = divs(xi_new, std::sqrt(dot(xi_new, xi_new)));
            xi_new = scale(norm(xi), xi_new);
        }
        xi = xi_new;
        V3 f_0 = sub(scale(-stiffness, xi), scale(damping, rel_v));
        xi = add(xi, scale(cp.dt, rel_v));
        return f_0;
    }

    // ---- alt_sinter_bridge.h:92-131 ------------------------------------------------------------------------------------
    Accel alt_pair(int64_t i, int64_t j) {
        auto& row = alt_springs[(size_t)i];
        auto it = std::lower_bound(row.begin(), row.end(), j, [](SpringRec const& p, int64_t key) { return p.j < key; });
        if (it == row.end() || it->j != j) return contact_pair(i, j);
        Springs& s = it->s;
        const double* b = cfg.alt_bond;  // k, gamma_n, k_t, gamma_t, k_r, gamma_r, k_o, gamma_o
        V3 dx = sub(x[(size_t)i], x[(size_t)j]);
        V3 nrm = normalized(dx);
        double overlap = 2.0 * r - dot(dx, nrm);
        double r_prime = r - overlap / 2.0;
        V3 vi = v[(size_t)i], vj = v[(size_t)j], wi = omega[(size_t)i], wj = omega[(size_t)j];
        double v_n = -dot(sub(vi, vj), nrm);
        double f_n = b[0] * overlap + b[1] * v_n;
        V3 ncwi = cross(nrm, wi), ncwj = cross(nrm, wj);
        V3 v_ij = add(add(sub(vi, vj), scale(r_prime, ncwi)), scale(r_prime, ncwj));
        V3 v_t = sub(v_ij, scale(dot(v_ij, nrm), nrm));
        V3 v_r = scale(-r_prime, sub(ncwi, ncwj));
        V3 v_o = scale(r * (dot(nrm, wi) - dot(nrm, wj)), nrm);
        V3 f_t = alt_shear(s.xi[0], nrm, b[2], b[3], v_t);
        V3 f_r = alt_shear(s.xi[1], nrm, b[4], b[5], v_r);
        V3 f_o = alt_shear(s.xi[2], nrm, b[6], b[7], v_o);
        V3 tau_t = scale(r_prime, cross(nrm, f_t));
        V3 tau_r = scale(r, cross(nrm, f_r));
        V3 tau_o = scale(r, f_o);
        return Accel{divs(add(scale(f_n, nrm), f_t), mass), divs(add(add(neg(tau_t), tau_r), tau_o), inertia)};
    }

    // ---- candidate pairs within `range` of each other, via a uniform grid (any range) ------------
    template <typename F>
    void for_each_near(double range, F&& fn) const {
        // fn(i, j) for all ordered (i, j), i != j, whose cells are adjacent; caller applies the exact predicate
        V3 lo = x[0], hi = x[0];
        for (int64_t i = 1; i < n; i++)
            for (int k = 0; k < 3; k++) {
                lo[k] = std::min(lo[k], x[(size_t)i][k]);
                hi[k] = std::max(hi[k], x[(size_t)i][k]);
            }
        double cell = range * 1.0001;
        int64_t dim[3];
        for (int k = 0; k < 3; k++) dim[k] = std::max<int64_t>(1, (int64_t)std::floor((hi[k] - lo[k]) / cell) + 1);
        // cap the table; a coarser grid only yields more candidates
        while ((double)dim[0] * (double)dim[1] * (double)dim[2] > 64.0 * (double)n + 1024.0) {
            cell *= 1.5;
            for (int k = 0; k < 3; k++) dim[k] = std::max<int64_t>(1, (int64_t)std::floor((hi[k] - lo[k]) / cell) + 1);
        }
        auto cell_of = [&](V3 p, int64_t* ci) {
            for (int k = 0; k < 3; k++) {
                int64_t t = (int64_t)std::floor((p[k] - lo[k]) / cell);
                ci[k] = std::min(std::max<int64_t>(t, 0), dim[k] - 1);
            }
        };
        int64_t ncell = dim[0] * dim[1] * dim[2];
        std::vector<int64_t> start((size_t)ncell + 1, 0), order((size_t)n);
        std::vector<int64_t> cidx((size_t)n);
        for (int64_t i = 0; i < n; i++) {
            int64_t ci[3];
            cell_of(x[(size_t)i], ci);
            cidx[(size_t)i] = (ci[0] * dim[1] + ci[1]) * dim[2] + ci[2];
            start[(size_t)cidx[(size_t)i] + 1]++;
        }
        for (int64_t c = 0; c < ncell; c++) start[(size_t)c + 1] += start[(size_t)c];
        std::vector<int64_t> cur(start.begin(), start.end() - 1);
        for (int64_t i = 0; i < n; i++) order[(size_t)cur[(size_t)cidx[(size_t)i]]++] = i;
#pragma omp parallel for schedule(dynamic, 256)
        for (int64_t i = 0; i < n; i++) {
            int64_t ci[3];
            cell_of(x[(size_t)i], ci);
            for (int64_t dx = -1; dx <= 1; dx++)
                for (int64_t dy = -1; dy <= 1; dy++)
                    for (int64_t dz = -1; dz <= 1; dz++) {
                        int64_t cx = ci[0] + dx, cy = ci[1] + dy, cz = ci[2] + dz;
                        if (cx < 0 || cy < 0 || cz < 0 || cx >= dim[0] || cy >= dim[1] || cz >= dim[2]) continue;
                        int64_t c = (cx * dim[1] + cy) * dim[2] + cz;
                        for (int64_t s = start[(size_t)c]; s < start[(size_t)c + 1]; s++) {
                            int64_t j = order[(size_t)s];
                            if (j != i) fn(i, j);
                        }
                    }
        }
    }

    // ---- sinter_bridge.h:148-211 ---------------------------------------------------------------
    void build_bonds() {
        std::mt19937_64 mt(cfg.rng_seed);
        std::uniform_real_distribution<double> dist(-1.0, 1.0);
        auto random_unit_vector = [&]() {  // test/sintered.cpp:33-39
            V3 vec;
            do {
                double a0 = dist(mt), a1 = dist(mt), a2 = dist(mt);
                vec = mk(a0, a1, a2);
            } while (norm(vec) == 0);
            return normalized(vec);
        };
        const double crit = cfg.critical_separation;
        // candidate (i < j) pairs in lexicographic order
        std::vector<std::vector<int64_t>> cand((size_t)n);
        if (n <= 2048 || brute_force) {
            for (int64_t i = 0; i < n - 1; i++)
                for (int64_t j = i + 1; j < n; j++) cand[(size_t)i].push_back(j);
        } else {
            for_each_near(2.0 * r + crit, [&](int64_t i, int64_t j) {
                if (j > i) cand[(size_t)i].push_back(j);  // row i is written only by iteration i
            });
            for (auto& row : cand) std::sort(row.begin(), row.end());
        }
        for (int64_t i = 0; i < n - 1; i++) {
            for (int64_t j : cand[(size_t)i]) {
                V3 part_i = x[(size_t)i], part_j = x[(size_t)j];
                double separation = norm(sub(part_j, part_i)) - 2.0 * r;
                if (std::fabs(separation) > crit) continue;
                V3 n_i = normalized(sub(part_j, part_i));
                V3 n_j = neg(n_i);
                V3 tv, db;
                do {
                    tv = random_unit_vector();
                    db = cross(tv, n_j);
                } while (norm(db) == 0.0);
                V3 ds = sub(tv, scale(dot(tv, n_i), n_i));
                db = divs(db, std::sqrt(dot(db, db)));
                ds = divs(ds, std::sqrt(dot(ds, ds)));
                Bond bi, bj;
                bi.j = j;
                bj.j = i;
                bi.conn[0] = add(part_i, scale(r, ds));  // ds * r_part : commutative product per component
                bi.conn[1] = sub(part_i, scale(r, ds));
                bi.conn[2] = add(part_i, scale(r, db));
                bi.conn[3] = sub(part_i, scale(r, db));
                bi.conn[4] = add(part_i, scale(r, n_i));
                bj.conn[0] = add(part_j, scale(r, ds));
                bj.conn[1] = sub(part_j, scale(r, ds));
                bj.conn[2] = add(part_j, scale(r, db));
                bj.conn[3] = sub(part_j, scale(r, db));
                bj.conn[4] = add(part_j, scale(r, n_j));
                for (int k = 0; k < 4; k++) bi.len[k] = bj.len[k] = 2.0 * r;
                bi.len[4] = bj.len[4] = 0.0;
                bi.rev = (int64_t)bonds[(size_t)j].size();
                bj.rev = (int64_t)bonds[(size_t)i].size();
                bonds[(size_t)i].push_back(bi);
                bonds[(size_t)j].push_back(bj);
            }
        }
        // bonds[i] was filled with partners in ascending order for j > i and ascending i for j < i,
        // i.e. already ascending overall (all partners < i are appended before any partner > i).
    }

    const Bond* find_bond(int64_t i, int64_t j) const {
        for (auto const& b : bonds[(size_t)i])
            if (b.j == j) return &b;
        return nullptr;
    }

    // ---- contact_force.h:57-96 -----------------------------------------------------------------
    Accel contact_pair(int64_t i, int64_t j) {
        V3 dx = sub(x[(size_t)i], x[(size_t)j]);
        V3 nrm = normalized(dx);
        double overlap = 2.0 * r - dot(dx, nrm);
        auto& row = springs[(size_t)i];
        auto it = std::find_if(row.begin(), row.end(), [j](SpringRec const& s) { return s.j == j; });
        if (overlap <= 0) {
            if (it != row.end()) row.erase(it);  // reset_springs: a zero record == no record
            return Accel{zero3(), zero3()};
        }
        if (it == row.end()) {
            row.push_back(SpringRec{j, Springs{{zero3(), zero3(), zero3()}}});
            it = row.end() - 1;
        }
        Springs& s = it->s;
        V3 vi = v[(size_t)i], vj = v[(size_t)j], wi = omega[(size_t)i], wj = omega[(size_t)j];
        double r_prime = r - overlap / 2.0;
        double v_n = -dot(sub(vi, vj), nrm);
        double f_n = cp.k * overlap + cp.gamma_n * v_n;
        V3 ncwi = cross(nrm, wi), ncwj = cross(nrm, wj);
        V3 v_ij = add(add(sub(vi, vj), scale(r_prime, ncwi)), scale(r_prime, ncwj));
        V3 v_t = sub(v_ij, scale(dot(v_ij, nrm), nrm));
        V3 v_r = scale(-r_prime, sub(ncwi, ncwj));
        V3 v_o = scale(r * (dot(nrm, wi) - dot(nrm, wj)), nrm);
        V3 f_t = shear_contribution(s.xi[0], nrm, cp.k_t, cp.gamma_t, f_n, cp.mu_s, cp.phi_d, v_t, cp.dt);
        V3 f_r = shear_contribution(s.xi[1], nrm, cp.k_r, cp.gamma_r, f_n, cp.mu_r, cp.phi_r, v_r, cp.dt);
        V3 f_o = shear_contribution(s.xi[2], nrm, cp.k_o, cp.gamma_o, f_n, cp.mu_o, cp.phi_o, v_o, cp.dt);
        V3 tau_t = scale(r_prime, cross(nrm, f_t));
        V3 tau_r = scale(r, cross(nrm, f_r));
        V3 tau_o = scale(r, f_o);
        return Accel{divs(add(scale(f_n, nrm), f_t), mass), divs(add(add(neg(tau_t), tau_r), tau_o), inertia)};
    }

    // ---- hamaker_force.h:28-43 -----------------------------------------------------------------
    Accel hamaker_pair(int64_t i, int64_t j) const {
        if (cfg.hamaker_kind == DEM_HAMAKER_LIMITED_RANGE) {  // limited_range_hamaker_force.h:34-35
            if (norm(sub(x[(size_t)i], x[(size_t)j])) > cfg.hamaker_cutoff) return Accel{zero3(), zero3()};
        }
        V3 rv = sub(x[(size_t)j], x[(size_t)i]);
        double h = norm(rv) - 2.0 * r;
        if (h < cfg.hamaker_h0) h = cfg.hamaker_h0;
        double A = cfg.hamaker_A;
        if (cfg.hamaker_kind == DEM_HAMAKER_SIMPLIFIED) {  // simplified_hamaker_force.h:27
            V3 fs = scale(2.0 * A * r / square(h), normalized(rv));
            return Accel{divs(fs, mass), zero3()};
        }
        double coef = -A / 6.0 *
                      ((4.0 * r + 2.0 * h) / (4.0 * r + h) / h - 2.0 / (2.0 * r + h) - 4.0 * square(r) / cube(2.0 * r + h) -
                       2.0 * square(r) * (4.0 * r + 2.0 * h) / square(4.0 * r + h) / square(h));
        V3 f = scale(coef, normalized(rv));
        return Accel{divs(f, mass), zero3()};
    }

    // ---- sinter_bridge.h:235-259 ---------------------------------------------------------------
    Accel bonded_pair(int64_t i, int64_t j, Bond const& bi) const {
        Bond const& bj = bonds[(size_t)j][(size_t)bi.rev];
        V3 force = zero3(), torque = zero3();
        V3 xi_ = x[(size_t)i], xj_ = x[(size_t)j];
        for (int k = 0; k < 5; k++) {
            V3 spring = sub(bj.conn[k], bi.conn[k]);
            double spring_length = norm(spring);
            double eq = bi.len[k];
            V3 ci_v = add(v[(size_t)i], cross(omega[(size_t)i], sub(bi.conn[k], xi_)));
            V3 cj_v = add(v[(size_t)j], cross(omega[(size_t)j], sub(bj.conn[k], xj_)));
            V3 rel_v = sub(cj_v, ci_v);
            V3 dforce = add(scale(-(eq - spring_length) * cfg.k_d, normalized(spring)), scale(cfg.gamma_d, rel_v));
            force = add(force, dforce);
            V3 arm = sub(bi.conn[k], xi_);
            torque = add(torque, cross(arm, dforce));
        }
        return Accel{divs(force, mass), divs(torque, inertia)};
    }

    // ---- sinter_bridge.h:214-225 ---------------------------------------------------------------
    Accel sinter_pair(int64_t i, int64_t j) {
        if (const Bond* b = find_bond(i, j)) return bonded_pair(i, j, *b);
        return contact_pair(i, j);
    }

    // ---- binary_force_container.h:20-32 : (f1(i,j) + (f2(i,j) + ...)) in container order ----------
    Accel binary(int64_t i, int64_t j) {
        // at most {contact | sinter | alt_sinter} and {hamaker}: a two-term fold, contact/sinter first
        if (has_alt) {
            Accel s = alt_pair(i, j);
            return has_hamaker ? s + hamaker_pair(i, j) : s;
        }
        if (has_sinter) {
            Accel s = sinter_pair(i, j);
            return has_hamaker ? s + hamaker_pair(i, j) : s;
        }
        if (has_contact) {
            Accel c = contact_pair(i, j);
            return has_hamaker ? c + hamaker_pair(i, j) : c;
        }
        return hamaker_pair(i, j);
    }

    // ---- surface_contact_force.h:57-96 ---------------------------------------------------------
    Accel surface_contact(int64_t i, V3 x_facet, V3 v_facet) {
        V3 dx = sub(x[(size_t)i], x_facet);
        V3 nrm = normalized(dx);
        double overlap = r - dot(dx, nrm);
        Springs& s = wall_springs[(size_t)i];
        if (overlap <= 0) {
            s = Springs{{zero3(), zero3(), zero3()}};
            return Accel{zero3(), zero3()};
        }
        V3 vi = v[(size_t)i], wi = omega[(size_t)i];
        double r_prime = r - overlap;
        double v_n = -dot(sub(vi, v_facet), nrm);
        double f_n = sp.k * overlap + sp.gamma_n * v_n;
        V3 ncw = cross(nrm, wi);
        V3 v_ij = add(sub(vi, v_facet), scale(r_prime, ncw));
        V3 v_t = sub(v_ij, scale(dot(v_ij, nrm), nrm));
        V3 v_r = scale(-r_prime, ncw);
        V3 v_o = scale(r * dot(nrm, wi), nrm);
        V3 f_t = shear_contribution(s.xi[0], nrm, sp.k_t, sp.gamma_t, f_n, sp.mu_s, sp.phi_d, v_t, sp.dt);
        V3 f_r = shear_contribution(s.xi[1], nrm, sp.k_r, sp.gamma_r, f_n, sp.mu_r, sp.phi_r, v_r, sp.dt);
        V3 f_o = shear_contribution(s.xi[2], nrm, sp.k_o, sp.gamma_o, f_n, sp.mu_o, sp.phi_o, v_o, sp.dt);
        V3 tau_t = scale(r_prime, cross(nrm, f_t));
        V3 tau_r = scale(r, cross(nrm, f_r));
        V3 tau_o = scale(r, f_o);
        return Accel{divs(add(scale(f_n, nrm), f_t), mass), divs(add(add(neg(tau_t), tau_r), tau_o), inertia)};
    }

    // ---- surface_hamaker_force.h:18-30 ---------------------------------------------------------
    Accel surface_hamaker(int64_t i, V3 x_facet) const {
        V3 rv = sub(x_facet, x[(size_t)i]);
        double h = norm(rv) - r;
        if (h < cfg.surf_h0) h = cfg.surf_h0;
        double coef = 2.0 * cfg.surf_A / 3.0 * (cube(r) / square(h) / square(h + 2.0 * r));
        V3 f = scale(coef, normalized(rv));
        return Accel{divs(f, mass), zero3()};
    }

    // ---- triangular_facet.h:32-51 --------------------------------------------------------------
    Accel unary(int64_t i) {
        V3 x_facet = facet.closest_point_on_triangle(x[(size_t)i]);
        return surface_contact(i, x_facet, facet.v_facet) + surface_hamaker(i, x_facet);
    }

    // ---- rotational_binary_system[_neighbors]_omp.h operator() -----------------------------------
    void evaluate() {
#pragma omp parallel for schedule(dynamic, 64)
        for (int64_t i = 0; i < n; i++) {
            V3 ai = zero3(), ali = zero3();
            if (neigh) {
                for (int64_t j : rows[(size_t)i]) {
                    Accel c = binary(i, j);
                    ai = add(ai, c.a);
                    ali = add(ali, c.alpha);
                }
            } else {
                for (int64_t j = 0; j < n; j++) {
                    if (j == i) continue;
                    Accel c = binary(i, j);
                    ai = add(ai, c.a);
                    ali = add(ali, c.alpha);
                }
            }
            if (has_facet) {
                Accel c = unary(i);
                ai = add(ai, c.a);
                ali = add(ali, c.alpha);
            }
            a[(size_t)i] = ai;
            alpha[(size_t)i] = ali;
        }
    }

    // ---- rotational_binary_system_neighbors_omp.h:102-116 ----------------------------------------
    void update_neighbor_list() {
        if (!neigh) return;
        const double rv = cfg.r_verlet;
        if (n <= 2048 || brute_force) {
#pragma omp parallel for schedule(dynamic, 16)
            for (int64_t i = 0; i < n; i++) {
                auto& row = rows[(size_t)i];
                row.clear();
                for (int64_t j = 0; j < n; j++) {
                    if (i == j) continue;
                    double distance = norm(sub(x[(size_t)i], x[(size_t)j]));
                    if (distance < rv) row.push_back(j);
                }
            }
        } else {
            for (auto& row : rows) row.clear();
            for_each_near(rv, [&](int64_t i, int64_t j) {
                double distance = norm(sub(x[(size_t)i], x[(size_t)j]));
                if (distance < rv) rows[(size_t)i].push_back(j);
            });
#pragma omp parallel for schedule(dynamic, 256)
            for (int64_t i = 0; i < n; i++) std::sort(rows[(size_t)i].begin(), rows[(size_t)i].end());
        }
    }

    // ---- rotational_velocity_verlet_half.h:43-76 with rotational_step_handler.h:19-68 or
    //      sinter_step_handler (sinter_bridge.h:24-86) ---------------------------------------------
    // ---- rotational_forward_euler.h:44-63 (two-term Taylor expansion in x and theta; v, omega copied BEFORE the
    //      increments, a / alpha by reference) with the step handlers as below ---------------------------------------------
    void do_step_forward_euler(double dt) {
        evaluate();
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; i++) {
            size_t s = (size_t)i;
            const V3 vc = v[s];
            const V3 wc = omega[s];
            V3 dx = add(scale(dt, vc), scale(dt, scale(dt, scale(0.5, a[s]))));  // v*dt + ((0.5*a)*dt)*dt
            x[s] = add(x[s], dx);
            for (auto& b : bonds[s])
                for (int k = 0; k < 5; k++) b.conn[k] = add(b.conn[k], dx);
            if (i != cfg.fixed_particle) v[s] = add(v[s], scale(dt, a[s]));
            V3 dth = add(scale(dt, wc), scale(dt, scale(dt, scale(0.5, alpha[s]))));
            theta[s] = add(theta[s], dth);
            for (auto& b : bonds[s])
                for (int k = 0; k < 5; k++) {
                    V3 sv = sub(b.conn[k], x[s]);
                    b.conn[k] = add(b.conn[k], cross(dth, sv));
                }
            omega[s] = add(omega[s], scale(dt, alpha[s]));
        }
    }

    void do_step(double dt) {
        if (cfg.integrator == DEM_INTEGRATOR_FORWARD_EULER) {
            do_step_forward_euler(dt);
            return;
        }
        if (!velocities_initialized) {
            evaluate();
            velocities_initialized = true;
            for (int64_t i = 0; i < n; i++) {
                if (i != cfg.fixed_particle)  // increment_v of the fixed particle is swallowed by its step handler
                    v[(size_t)i] = add(v[(size_t)i], divs(scale(dt, neg(a[(size_t)i])), 2.0));
                omega[(size_t)i] = add(omega[(size_t)i], divs(scale(dt, neg(alpha[(size_t)i])), 2.0));
            }
            evaluate();
        }
#pragma omp parallel for schedule(static)
        for (int64_t i = 0; i < n; i++) {
            size_t s = (size_t)i;
            if (i != cfg.fixed_particle) v[s] = add(v[s], scale(dt, a[s]));
            V3 dx = scale(dt, v[s]);
            x[s] = add(x[s], dx);
            for (auto& b : bonds[s])
                for (int k = 0; k < 5; k++) b.conn[k] = add(b.conn[k], dx);
            omega[s] = add(omega[s], scale(dt, alpha[s]));
            V3 dth = scale(dt, omega[s]);
            theta[s] = add(theta[s], dth);
            for (auto& b : bonds[s])
                for (int k = 0; k < 5; k++) {
                    V3 sv = sub(b.conn[k], x[s]);
                    b.conn[k] = add(b.conn[k], cross(dth, sv));
                }
        }
        evaluate();
    }

    void run(double dt, int64_t n_steps, int64_t rebuild_every) {
        for (int64_t s = 0; s < n_steps; s++, step_count++) {
            if (neigh && rebuild_every > 0 && step_count % rebuild_every == 0) update_neighbor_list();
            do_step(dt);
        }
    }
};

}  // namespace

extern "C" {

void* orc_create(const dem_config* cfg, const double* x0, const double* v0, const double* th0, const double* om0) {
    const int bin = cfg->model_mask & (DEM_MODEL_CONTACT | DEM_MODEL_HAMAKER | DEM_MODEL_SINTER | DEM_MODEL_ALT_SINTER);
    const int excl = ((bin & DEM_MODEL_CONTACT) != 0) + ((bin & DEM_MODEL_SINTER) != 0) + ((bin & DEM_MODEL_ALT_SINTER) != 0);
    if (bin == 0 || excl > 1) return nullptr;  // sinter and alt_sinter embed the contact model
    return new Oracle(*cfg, x0, v0, th0, om0);
}
void orc_destroy(void* h) { delete static_cast<Oracle*>(h); }
void orc_update_neighbor_list(void* h) { static_cast<Oracle*>(h)->update_neighbor_list(); }
void orc_do_step(void* h, double dt) { static_cast<Oracle*>(h)->do_step(dt); }
void orc_run(void* h, double dt, int64_t n, int64_t every) { static_cast<Oracle*>(h)->run(dt, n, every); }
void orc_facet_update_positions(void* h, double dt) {
    Oracle* o = static_cast<Oracle*>(h);
    if (o->has_facet) o->facet.update_positions(dt);
}
void orc_get_field(void* h, int field, double* out) {
    Oracle* o = static_cast<Oracle*>(h);
    const std::vector<V3>* src = nullptr;
    switch (field) {
        case DEM_X: src = &o->x; break;
        case DEM_V: src = &o->v; break;
        case DEM_A: src = &o->a; break;
        case DEM_THETA: src = &o->theta; break;
        case DEM_OMEGA: src = &o->omega; break;
        case DEM_ALPHA: src = &o->alpha; break;
        default: return;
    }
    std::memcpy(out, src->data(), src->size() * sizeof(V3));
}
int64_t orc_n_pairs(void* h) {
    Oracle* o = static_cast<Oracle*>(h);
    if (!o->neigh) return o->n * (o->n - 1);
    int64_t p = 0;
    for (auto const& row : o->rows) p += (int64_t)row.size();
    return p;
}
void orc_get_rows(void* h, int64_t* row_ptr, int64_t* cols) {
    Oracle* o = static_cast<Oracle*>(h);
    int64_t p = 0;
    for (int64_t i = 0; i < o->n; i++) {
        row_ptr[i] = p;
        if (o->neigh) {
            for (int64_t j : o->rows[(size_t)i]) cols[p++] = j;
        } else {
            for (int64_t j = 0; j < o->n; j++)
                if (j != i) cols[p++] = j;
        }
    }
    row_ptr[o->n] = p;
}
void orc_get_contact_springs(void* h, int64_t i, int64_t j, double* out9) {
    Oracle* o = static_cast<Oracle*>(h);
    std::fill(out9, out9 + 9, 0.0);
    for (auto const& s : o->springs[(size_t)i])
        if (s.j == j) std::memcpy(out9, s.s.xi, 72);
    if (o->has_alt)
        for (auto const& s : o->alt_springs[(size_t)i])
            if (s.j == j) std::memcpy(out9, s.s.xi, 72);
}
int orc_get_bond(void* h, int64_t i, int64_t j, double* conn15, double* len5) {
    Oracle* o = static_cast<Oracle*>(h);
    if (o->has_alt) {  // no connectors in this model: only the flag
        std::fill(conn15, conn15 + 15, 0.0);
        std::fill(len5, len5 + 5, 0.0);
        for (auto const& s : o->alt_springs[(size_t)i])
            if (s.j == j) return 1;
        return 0;
    }
    const Bond* b = o->find_bond(i, j);
    if (!b) return 0;
    std::memcpy(conn15, b->conn, 120);
    std::memcpy(len5, b->len, 40);
    return 1;
}
void orc_get_facet_vertices(void* h, double* out9) {
    Oracle* o = static_cast<Oracle*>(h);
    if (o->has_facet)
        std::memcpy(out9, o->facet.vtx, 72);
    else
        std::fill(out9, out9 + 9, 0.0);
}
int64_t orc_n_bonds(void* h) {
    Oracle* o = static_cast<Oracle*>(h);
    int64_t c = 0;
    for (auto const& b : o->bonds) c += (int64_t)b.size();
    for (auto const& b : o->alt_springs) c += (int64_t)b.size();
    return c;
}
int64_t orc_n_contacts(void* h) {
    Oracle* o = static_cast<Oracle*>(h);
    int64_t c = 0;
    for (auto const& s : o->springs) c += (int64_t)s.size();
    return c;
}
void orc_set_brute_force_neighbors(void* h, int on) { static_cast<Oracle*>(h)->brute_force = on != 0; }

// test/sintered.cpp:29-39
void orc_unit_vectors(uint64_t seed, int64_t count, double* out) {
    std::mt19937_64 mt(seed);
    std::uniform_real_distribution<double> dist(-1.0, 1.0);
    for (int64_t k = 0; k < count; k++) {
        V3 vec;
        do {
            double a0 = dist(mt), a1 = dist(mt), a2 = dist(mt);
            vec = mk(a0, a1, a2);
        } while (norm(vec) == 0);
        vec = normalized(vec);
        out[3 * k] = vec[0]; out[3 * k + 1] = vec[1]; out[3 * k + 2] = vec[2];
    }
}

// test/compute_energy.cpp:7-58, test/mass_distribution.cpp:7-26
void orc_diagnostics(const double* xs, const double* vs, const double* ws, int64_t n, double mass, double inertia,
                     double* out8) {
    auto at = [](const double* p, int64_t i) { return mk(p[3 * i], p[3 * i + 1], p[3 * i + 2]); };
    double energy = 0.0;
    for (int64_t i = 0; i < n; i++) {
        energy += mass * dot(at(vs, i), at(vs, i)) / 2.0;
        energy += inertia * dot(at(ws, i), at(ws, i)) / 2.0;
    }
    V3 mom = zero3();
    for (int64_t i = 0; i < n; i++) mom = add(mom, at(vs, i));
    V3 ang = zero3();
    for (int64_t i = 0; i < n; i++) ang = add(ang, scale(inertia, at(ws, i)));
    for (int64_t i = 0; i < n; i++) ang = add(ang, cross(at(xs, i), scale(mass, at(vs, i))));
    V3 center = zero3();
    for (int64_t i = 0; i < n; i++) center = add(center, at(xs, i));
    center = divs(center, double(n));
    double rg = 0.0;
    for (int64_t i = 0; i < n; i++) {
        V3 d = sub(at(xs, i), center);
        rg += dot(d, d);
    }
    out8[0] = energy;
    out8[1] = mass * norm(mom);
    out8[2] = norm(ang);
    out8[3] = center[0]; out8[4] = center[1]; out8[5] = center[2];
    out8[6] = std::sqrt(rg / double(n));
    out8[7] = 0.0;
}
int orc_max_threads(void) { return omp_get_max_threads(); }
void orc_set_threads(int n) { omp_set_num_threads(n); }

}  // extern "C"
