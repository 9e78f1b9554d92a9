// physics.cuh -- the force laws of libgran's hot path as inlinable device functions.
//
// Written from the formulas of the reference (restated in SURVEY.md appendix A.4), with the operand
// order of every expression kept so that, compiled with --fmad=false, each function returns the same
// bits as the reference functor built with -ffp-contract=off:
//   contact_force()      LG/contact_force/contact_force.h:57-96   (Luding normal + 3 history springs)
//   shear_spring()       LG/contact_force/contact_force.h:105-147 (also surface_contact_force.h:103-145)
//   hamaker_accel()      LG/hamaker_force/hamaker_force.h:28-43
//   bonded_spring()      LG/sinter_bridge/sinter_bridge.h:244-256 (one of the five bridge springs)
//   facet_closest_point() LG/surface_force/triangular_facet.h:73-137
//   wall_contact_force() LG/contact_force/surface_contact_force.h:57-96
//   wall_hamaker_accel() LG/hamaker_force/surface_hamaker_force.h:18-30
// The functions are stateless: callers own where springs / connectors live.
#pragma once

#include "vec3.cuh"

namespace lg {

struct ContactParams {
    double k, gamma_n;
    double k_t, gamma_t, mu_s, phi_d;
    double k_r, gamma_r, mu_r, phi_r;
    double k_o, gamma_o, mu_o, phi_o;
    double dt;
};

struct Springs {
    d3 xi[3];  // 0: sliding/sticking, 1: rolling, 2: torsion
};

struct Accel {
    d3 a, alpha;
};

LG_HD Accel operator+(Accel l, Accel r) { return Accel{l.a + r.a, l.alpha + r.alpha}; }
LG_HD Accel zero_accel() { return Accel{zero3(), zero3()}; }

enum : int { MODEL_CONTACT = 1, MODEL_HAMAKER = 2, MODEL_SINTER = 4, MODEL_FACET = 8, MODEL_ALT_SINTER = 16 };
enum : int { HAMAKER_STANDARD = 0, HAMAKER_LIMITED_RANGE = 1, HAMAKER_SIMPLIFIED = 2 };

// Separation geometry shared by the contact and Hamaker laws of one ordered pair (i, j):
// contact uses n = normalized(x_i - x_j) (contact_force.h:65), Hamaker uses normalized(x_j - x_i)
// (hamaker_force.h:32,40), which is exactly -n, and both norms are sqrt of the same sum of squares.
struct PairGeom {
    d3 dx;       // x_i - x_j
    d3 n;        // normalized(dx)
    double len;  // |dx|
};

LG_HD PairGeom pair_geometry(d3 xi, d3 xj) {
    PairGeom g;
    g.dx = xi - xj;
    double z = dot(g.dx, g.dx);
    g.len = sqrt(z);
    g.n = (z > 0.0) ? g.dx / g.len : g.dx;
    return g;
}

// contact_force.h:66 / surface_contact_force.h:67 (two_r = 2 r for a pair, r for a wall)
LG_HD double contact_overlap(const PairGeom& g, double two_r) { return two_r - dot(g.dx, g.n); }

// One history spring: rotate into the tangent plane keeping its length, test force, Coulomb switch.
LG_HD d3 shear_spring(d3& xi, d3 n, double stiffness, double damping, double normal_force, double mu_static,
                      double phi_dynamic, d3 rel_v, double dt) {
    double mu_dynamic = mu_static * phi_dynamic;
    d3 xi_new = xi - dot(xi, n) * n;
    double z_new = dot(xi_new, xi_new);
    if (sqrt(z_new) > 0.0) {
        xi_new = xi_new / sqrt(z_new);
        xi_new = xi_new * norm(xi);
    }
    xi = xi_new;
    d3 f_0 = (-stiffness) * xi - damping * rel_v;
    double static_friction = mu_static * normal_force;
    double dynamic_friction = mu_dynamic * normal_force;
    double z0 = dot(f_0, f_0);
    double f0_len = sqrt(z0);
    d3 t = (f0_len > 0.0) ? f_0 / f0_len : zero3();
    double f_selected;
    if (f0_len <= static_friction) {
        xi = xi + rel_v * dt;
        f_selected = static_friction;
    } else {
        xi = (-1.0 / stiffness) * (dynamic_friction * t + damping * rel_v);
        f_selected = dynamic_friction;
    }
    double f_mag = (f0_len < f_selected) ? f0_len : f_selected;  // std::min(f_selected, |f_0|)
    return f_mag * t;
}

// Particle-particle contact for overlap > 0.  r_eff_sub = overlap / 2 (contact_force.h:73).
LG_HD Accel contact_force(const ContactParams& p, double r, double mass, double inertia, const PairGeom& g, double overlap,
                          d3 vi, d3 vj, d3 wi, d3 wj, Springs& s) {
    d3 n = g.n;
    double r_prime = r - overlap / 2.0;
    double v_n = -dot(vi - vj, n);
    double f_n = p.k * overlap + p.gamma_n * v_n;
    d3 ncwi = cross(n, wi), ncwj = cross(n, wj);
    d3 v_ij = ((vi - vj) + r_prime * ncwi) + r_prime * ncwj;
    d3 v_t = v_ij - dot(v_ij, n) * n;
    d3 v_r = (-r_prime) * (ncwi - ncwj);
    d3 v_o = (r * (dot(n, wi) - dot(n, wj))) * n;
    d3 f_t = shear_spring(s.xi[0], n, p.k_t, p.gamma_t, f_n, p.mu_s, p.phi_d, v_t, p.dt);
    d3 f_r = shear_spring(s.xi[1], n, p.k_r, p.gamma_r, f_n, p.mu_r, p.phi_r, v_r, p.dt);
    d3 f_o = shear_spring(s.xi[2], n, p.k_o, p.gamma_o, f_n, p.mu_o, p.phi_o, v_o, p.dt);
    d3 tau_t = r_prime * cross(n, f_t);
    d3 tau_r = r * cross(n, f_r);
    d3 tau_o = r * f_o;
    return Accel{(f_n * n + f_t) / mass, (((-tau_t) + tau_r) + tau_o) / inertia};
}

// Particle-wall contact for overlap > 0 (one-body kinematics, r' = r - overlap).
LG_HD Accel wall_contact_force(const ContactParams& p, double r, double mass, double inertia, const PairGeom& g,
                               double overlap, d3 vi, d3 v_facet, d3 wi, Springs& s) {
    d3 n = g.n;
    double r_prime = r - overlap;
    double v_n = -dot(vi - v_facet, n);
    double f_n = p.k * overlap + p.gamma_n * v_n;
    d3 ncw = cross(n, wi);
    d3 v_ij = (vi - v_facet) + r_prime * ncw;
    d3 v_t = v_ij - dot(v_ij, n) * n;
    d3 v_r = (-r_prime) * ncw;
    d3 v_o = (r * dot(n, wi)) * n;
    d3 f_t = shear_spring(s.xi[0], n, p.k_t, p.gamma_t, f_n, p.mu_s, p.phi_d, v_t, p.dt);
    d3 f_r = shear_spring(s.xi[1], n, p.k_r, p.gamma_r, f_n, p.mu_r, p.phi_r, v_r, p.dt);
    d3 f_o = shear_spring(s.xi[2], n, p.k_o, p.gamma_o, f_n, p.mu_o, p.phi_o, v_o, p.dt);
    d3 tau_t = r_prime * cross(n, f_t);
    d3 tau_r = r * cross(n, f_r);
    d3 tau_o = r * f_o;
    return Accel{(f_n * n + f_t) / mass, (((-tau_t) + tau_r) + tau_o) / inertia};
}

// Bonded pair of alt_sinter_functor (LG/sinter_bridge/alt_sinter_bridge.h:92-160): the contact kinematics with the
// bond's own stiffnesses, NO overlap test (the bond also pulls), no Coulomb limit -- the three springs always grow by
// v_rel * dt after the test force is taken.  b = {k, gamma_n, k_t, gamma_t, k_r, gamma_r, k_o, gamma_o}.
LG_HD d3 alt_shear_spring(d3& xi, d3 n, double stiffness, double damping, d3 rel_v, double dt) {
    d3 xi_new = xi - dot(xi, n) * n;
    double z_new = dot(xi_new, xi_new);
    if (sqrt(z_new) > 0.0) {
        xi_new = xi_new / sqrt(z_new);
        xi_new = xi_new * norm(xi);
    }
    xi = xi_new;
    d3 f_0 = (-stiffness) * xi - damping * rel_v;
    xi = xi + rel_v * dt;
    return f_0;
}
LG_HD Accel alt_bonded_force(const double* b, double dt, double r, double mass, double inertia, const PairGeom& g, d3 vi, d3 vj,
                             d3 wi, d3 wj, Springs& s) {
    d3 n = g.n;
    double overlap = 2.0 * r - dot(g.dx, n);
    double r_prime = r - overlap / 2.0;
    double v_n = -dot(vi - vj, n);
    double f_n = b[0] * overlap + b[1] * v_n;
    d3 ncwi = cross(n, wi), ncwj = cross(n, wj);
    d3 v_ij = ((vi - vj) + r_prime * ncwi) + r_prime * ncwj;
    d3 v_t = v_ij - dot(v_ij, n) * n;
    d3 v_r = (-r_prime) * (ncwi - ncwj);
    d3 v_o = (r * (dot(n, wi) - dot(n, wj))) * n;
    d3 f_t = alt_shear_spring(s.xi[0], n, b[2], b[3], v_t, dt);
    d3 f_r = alt_shear_spring(s.xi[1], n, b[4], b[5], v_r, dt);
    d3 f_o = alt_shear_spring(s.xi[2], n, b[6], b[7], v_o, dt);
    d3 tau_t = r_prime * cross(n, f_t);
    d3 tau_r = r * cross(n, f_r);
    d3 tau_o = r * f_o;
    return Accel{(f_n * n + f_t) / mass, (((-tau_t) + tau_r) + tau_o) / inertia};
}

// Hamaker sphere-sphere attraction on i from j; g is the geometry of (i, j) so x_j - x_i = -g.dx.
// kind: hamaker_functor | limited_range_hamaker_functor (zero beyond `cutoff` between centres,
// limited_range_hamaker_force.h:34-35) | simplified_hamaker_functor (2 A r / h^2, simplified_hamaker_force.h:27).
LG_HD d3 hamaker_accel(double A, double h0, double r, double mass, const PairGeom& g, int kind = HAMAKER_STANDARD,
                       double cutoff = 0.0) {
    if (kind == HAMAKER_LIMITED_RANGE && g.len > cutoff) return zero3();
    double h = g.len - 2.0 * r;
    if (h < h0) h = h0;
    if (kind == HAMAKER_SIMPLIFIED) {
        d3 fs = (2.0 * A * r / square(h)) * (-g.n);
        return fs / mass;
    }
    double coef = -A / 6.0 *
                  ((4.0 * r + 2.0 * h) / (4.0 * r + h) / h - 2.0 / (2.0 * r + h) - 4.0 * square(r) / cube(2.0 * r + h) -
                   2.0 * square(r) * (4.0 * r + 2.0 * h) / square(4.0 * r + h) / square(h));
    d3 f = coef * (-g.n);
    return f / mass;
}

// Particle-half-space Hamaker; g is the geometry of (x_i, x_facet).
LG_HD d3 wall_hamaker_accel(double A, double h0, double r, double mass, const PairGeom& g) {
    double h = g.len - r;
    if (h < h0) h = h0;
    double coef = 2.0 * A / 3.0 * (cube(r) / square(h) / square(h + 2.0 * r));
    d3 f = coef * (-g.n);
    return f / mass;
}

// One of the five springs of a sinter bridge: connector c_i on i, c_j on j, rest length eq.
// Accumulates into force / torque exactly like the loop body of sinter_bridge.h:244-256.
LG_HD void bonded_spring(double k_d, double gamma_d, double eq, d3 ci, d3 cj, d3 xi, d3 xj, d3 vi, d3 vj, d3 wi, d3 wj,
                         d3& force, d3& torque) {
    d3 spring = cj - ci;
    double spring_length = norm(spring);
    d3 arm = ci - xi;
    d3 ci_v = vi + cross(wi, arm);
    d3 cj_v = vj + cross(wj, cj - xj);
    d3 rel_v = cj_v - ci_v;
    d3 dforce = ((-(eq - spring_length)) * k_d) * normalized(spring) + gamma_d * rel_v;
    force += dforce;
    torque += cross(arm, dforce);
}

// ---- relaxed arithmetic (lg_set_math_mode(LG_MATH_FAST)) ------------------------------------------------------
// Same laws, evaluated with shared reciprocals, fused multiply-adds and an algebraically combined Hamaker term.
// Results differ from the reference's by a few ulp per pair (<= 1e-13 relative where the reference's own
// expression cancels), far inside the 1e-10 relative tolerance BASELINE.json's north_star states for forces;
// the DECISIONS that define contact sets (overlap <= 0) are still taken on the reference's exact expression
// whenever the fast value is within a rounding band of zero, so neighbor and contact sets stay bit-exact.
#if defined(__CUDACC__)
__device__ __forceinline__ double dot_fma(d3 a, d3 b) { return fma(a.z, b.z, fma(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ d3 cross_fma(d3 a, d3 b) {
    return d3{fma(a.y, b.z, -(a.z * b.y)), fma(a.z, b.x, -(a.x * b.z)), fma(a.x, b.y, -(a.y * b.x))};
}

// Reciprocal and reciprocal square root for NORMAL, positive operands (squared distances, products of gaps): the
// hardware seed (MUFU.RCP64H / MUFU.RSQ64H, >= 20 bits) refined by Newton steps to ~1 ulp.  Unlike 1.0 / x and rsqrt(x)
// there is no special-case slow path, hence no CALL in the instruction stream: a call inside the pair loop makes the
// compiler park the loop's accumulators in local memory (profiles/README.md r01f).
// ITER Newton steps: the seed carries >= 20 good bits, each step doubles them (20 -> 40 -> 80), so two steps reach the
// last bit or two of a double -- enough wherever 1e-10 is asked (directions, Hamaker); three where the result feeds a
// cancelling difference (bridge spring lengths, refined once more there).
#ifndef LG_FAST_NR_ITERS
#define LG_FAST_NR_ITERS 2
#endif
#ifndef LG_BRIDGE_NR_ITERS
#define LG_BRIDGE_NR_ITERS 3
#endif
template <int ITER = LG_FAST_NR_ITERS>
__device__ __forceinline__ double rcp_nr(double x) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
#pragma unroll
    for (int k = 0; k < ITER; k++) {
        const double e = fma(-x, y, 1.0);
        y = fma(y, e, y);
    }
    return y;
}
template <int ITER = LG_FAST_NR_ITERS>
__device__ __forceinline__ double rsqrt_nr(double x) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double hx = 0.5 * x;
#pragma unroll
    for (int k = 0; k < ITER; k++) {
        const double e = fma(-hx * y, y, 0.5);   // 0.5 (1 - x y^2)
        y = fma(y, e, y);
    }
    return y;
}
// sqrt(z) and 1 / sqrt(z) for z >= 0 without the library's special-case path: g = z y refined by the residual z - g^2
// (the closing step of the correctly rounded square root; the result is the correctly rounded one except for arguments
// within ~2^-100 of a rounding tie).  z == 0 -> both 0 (a coincident connector pair contributes no force).
__device__ __forceinline__ void sqrt_and_rsqrt(double z, double& root, double& inv) {
    const double y = z > 0.0 ? rsqrt_nr<LG_BRIDGE_NR_ITERS>(z) : 0.0;
    const double g = z * y;
    root = fma(0.5 * y, fma(-g, g, z), g);
    inv = y;
}

__device__ __forceinline__ PairGeom pair_geometry_fast(d3 xi, d3 xj) {
    PairGeom g;
    g.dx = xi - xj;
    const double z = dot_fma(g.dx, g.dx);
    const double inv = z > 0.0 ? rsqrt_nr(z) : 0.0;
    g.len = z * inv;
    g.n = g.dx * inv;
    return g;
}

#endif
// constants of the relaxed laws, computed once on the host (api.cu: evaluate)
struct FastConsts {
    double c0m;  // (A / 6) * 64 r^6 / mass
    double r2, r4, h0;
    double two_r, band;  // contact decision: |2r - len| < band is re-taken on the exact expression
    double inv_mass, inv_inertia;
    int kind;            // HAMAKER_*
    double cutoff;       // limited range: zero beyond this centre distance
    double c_simple;     // simplified: 2 A r / mass
};
inline FastConsts fast_consts(double A, double h0, double r, double mass, double inertia, int kind = HAMAKER_STANDARD,
                              double cutoff = 0.0) {
    const double r3 = r * r * r;
    return FastConsts{A / 6.0 * 64.0 * r3 * r3 / mass, 2.0 * r, 4.0 * r, h0, 2.0 * r, 1e-12 * r, 1.0 / mass, 1.0 / inertia,
                      kind, cutoff, 2.0 * A * r / mass};
}
typedef FastConsts HamakerFast;
#if defined(__CUDACC__)
// (4r+2h)/((4r+h)h) - 2/(2r+h) - 4r^2/(2r+h)^3 - 2r^2(4r+2h)/((4r+h)^2 h^2)  ==  -64 r^6 / (h^2 (2r+h)^3 (4r+h)^2)
__device__ __forceinline__ d3 hamaker_accel_fast(const HamakerFast& hf, const PairGeom& g) {
    double h = g.len - hf.r2;
    if (h < hf.h0) h = hf.h0;
    if (hf.kind != HAMAKER_STANDARD) {  // uniform over the launch: no divergence
        if (hf.kind == HAMAKER_LIMITED_RANGE) {
            if (g.len > hf.cutoff) return zero3();
        } else {
            return (-(hf.c_simple * rcp_nr(h * h))) * g.n;
        }
    }
    const double s = hf.r2 + h, q = hf.r4 + h;
    const double hq = h * q;
    const double den = (hq * hq) * (s * s * s);
    const double coef = hf.c0m * rcp_nr(den);   // > 0: attraction, along -n (towards j)
    return (-coef) * g.n;
}

template <bool DAMPED>
__device__ __forceinline__ void bonded_spring_fast(double k_d, double gamma_d, double eq, d3 ci, d3 cj, d3 xi, d3 xj, d3 vi,
                                                   d3 vj, d3 wi, d3 wj, d3& force, d3& torque) {
    // The stretch (length - eq) cancels to ~1e-5 of the length near equilibrium, so the length is computed to the
    // reference's value (same sum order, correctly rounded square root: sqrt_and_rsqrt); the direction shares its reciprocal.
    const d3 spring = cj - ci;
    const double z = dot(spring, spring);
    double spring_length, inv;
    sqrt_and_rsqrt(z, spring_length, inv);
    const d3 arm = ci - xi;
    d3 dforce = (((-(eq - spring_length)) * k_d) * inv) * spring;
    if constexpr (DAMPED) {
        const d3 ci_v = vi + cross_fma(wi, arm);
        const d3 cj_v = vj + cross_fma(wj, cj - xj);
        const d3 rel_v = cj_v - ci_v;
        dforce = d3{fma(gamma_d, rel_v.x, dforce.x), fma(gamma_d, rel_v.y, dforce.y), fma(gamma_d, rel_v.z, dforce.z)};
    }
    force += dforce;
    torque += cross_fma(arm, dforce);
}

// Both orientations of one bridge in one pass (step_rows2_kernel, both ends in the same tile): the springs of (j, i) are
// those of (i, j) reversed, so dF_ji = -dF_ij exactly; force_j = -force_i, torque_j = sum (c_j - x_j) x (-dF).
template <bool DAMPED>
__device__ __forceinline__ void bonded_spring_pair_fast(double k_d, double gamma_d, double eq, d3 ci, d3 cj, d3 xi, d3 xj, d3 vi,
                                                        d3 vj, d3 wi, d3 wj, d3& force, d3& torque_i, d3& torque_j) {
    const d3 spring = cj - ci;
    const double z = dot(spring, spring);
    double spring_length, inv;
    sqrt_and_rsqrt(z, spring_length, inv);
    const d3 arm_i = ci - xi, arm_j = cj - xj;
    d3 dforce = (((-(eq - spring_length)) * k_d) * inv) * spring;
    if constexpr (DAMPED) {
        const d3 rel_v = (vj + cross_fma(wj, arm_j)) - (vi + cross_fma(wi, arm_i));
        dforce = d3{fma(gamma_d, rel_v.x, dforce.x), fma(gamma_d, rel_v.y, dforce.y), fma(gamma_d, rel_v.z, dforce.z)};
    }
    force += dforce;
    torque_i += cross_fma(arm_i, dforce);
    torque_j = torque_j - cross_fma(arm_j, dforce);
}
#endif

// ---- triangular facet ---------------------------------------------------------------------------
struct Facet {
    d3 vtx[3];
    d3 v;            // v_facet
    d3 unit_normal;  // fixed at construction (triangular_facet.h:28), not updated by update_positions
    double area;     // fixed at construction (:26)
    ContactParams sp;
    double A, h0;
    int has_contact, has_hamaker;
};

LG_HD void facet_init_geometry(Facet& f) {
    d3 u = f.vtx[1] - f.vtx[0];
    d3 v = f.vtx[2] - f.vtx[0];
    f.area = 1.0 / 2.0 * norm(cross(u, v));
    f.unit_normal = normalized(cross(u, v));
}

LG_HD d3 facet_edge_point(const Facet& f, int i, int j, d3 p) {
    d3 line = f.vtx[j] - f.vtx[i];
    d3 unit = normalized(line);
    d3 tp = p - f.vtx[i];
    double len = norm(line);
    double proj = dot(tp, unit);
    if (proj < 0.0) return f.vtx[i];
    if (proj > len) return f.vtx[j];
    return f.vtx[i] + proj * unit;
}

LG_HD d3 facet_closest_point(const Facet& f, d3 p) {
    d3 offset = dot(f.vtx[0], f.unit_normal) * f.unit_normal;
    d3 projection = p - dot(p, f.unit_normal) * f.unit_normal;
    d3 cp = projection + offset;
    {
        double area = 0.0;
        d3 pa = f.vtx[0] - cp, pb = f.vtx[1] - cp, pc = f.vtx[2] - cp;
        area += 1.0 / 2.0 * norm(cross(pa, pb));
        area += 1.0 / 2.0 * norm(cross(pb, pc));
        area += 1.0 / 2.0 * norm(cross(pc, pa));
        if (!(area > f.area)) return cp;
    }
    d3 e1 = facet_edge_point(f, 0, 1, p);
    d3 e2 = facet_edge_point(f, 1, 2, p);
    d3 e3 = facet_edge_point(f, 2, 0, p);
    d3 d1 = p - e1, d2 = p - e2, d3_ = p - e3;
    double s1 = dot(d1, d1), s2 = dot(d2, d2), s3 = dot(d3_, d3_);
    if (s1 < s2) {
        if (s1 < s3) return e1;
    } else {
        if (s2 < s3) return e2;
    }
    return e3;
}

}  // namespace lg
