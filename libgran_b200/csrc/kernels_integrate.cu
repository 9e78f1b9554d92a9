// kernels_integrate.cu -- the stand-alone leap-frog ("velocity Verlet half") update, in place.
//
// Used on the first step, after a neighbor rebuild and by the validation path; in steady state the update is
// fused into the step kernel (kernels_force_tiled.cu).  Replaces the serial loop of
// LT/rotational_integrator/rotational_velocity_verlet_half.h:63-73 with the default step handler
// (LT/rotational_step_handler/rotational_step_handler.h:19-68) or, when bonds exist, the sinter step handler
// (LG/sinter_bridge/sinter_bridge.h:24-42,55-75):
//   v += a dt ; x += v dt (updated v) ; omega += alpha dt ; theta += omega dt (updated omega)
//   each connector c of each bond of the particle: c += dx ; c += dtheta x (c - x_new)
// and the first-call half kick back v -= a dt/2, omega -= alpha dt/2 (:51-57).
//
// SURVEY.md section 8f variants: rotational_forward_euler (LT/rotational_integrator/rotational_forward_euler.h:44-63:
//   x += v dt + 0.5 a dt dt (v copied before it is incremented) ; v += a dt ; theta += omega dt + 0.5 alpha dt dt ; omega += alpha dt)
// and the fixed-particle step handler of Writerside/topics/Custom-step-handler.md:48-146 (increment_v of one particle is dropped).
#include "system.h"

namespace lg {

// flat over the 3n components: every operation of the particle update is component-wise
__global__ void __launch_bounds__(256) integrate_particles_kernel(int n3, double* x, double* v, double* th, double* w,
                                                                  const double* a, const double* al, double dt, int fixed) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n3) return;
    const double vk = (k / 3 != fixed) ? v[k] + a[k] * dt : v[k];  // increment_v(n, a * dt)
    v[k] = vk;
    x[k] = x[k] + vk * dt;  // increment_x(n, v * dt) with v already updated (:64,69-70)
    const double wk = w[k] + al[k] * dt;  // increment_omega
    w[k] = wk;
    th[k] = th[k] + wk * dt;  // increment_theta with omega already updated (:66,71-72)
}

// one thread per connector; runs after integrate_particles_kernel, so v, x, w already hold the new values and
// dx = v_new * dt, dtheta = omega_new * dt are recomputed with the very same operands
__global__ void __launch_bounds__(256) integrate_connectors_kernel(int n, int64_t n_conn, const uint32_t* bptr, double* bconn,
                                                                   const double* x, const double* v, const double* w, double dt) {
    const int64_t cq = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (cq >= n_conn) return;
    const uint32_t b = (uint32_t)(cq / 5);
    int lo = 0, hi = n;  // owner: largest i with bptr[i] <= b
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (bptr[mid] <= b) lo = mid; else hi = mid - 1;
    }
    const size_t o = 3 * (size_t)lo;
    const d3 dx = load3(v + o) * dt;
    const d3 dth = load3(w + o) * dt;
    double* c = bconn + 3 * (size_t)cq;
    d3 ck = load3(c) + dx;                // sinter_bridge.h:36-41
    const d3 sv = ck - load3(x + o);      // :70-71, x already incremented
    ck = ck + cross(dth, sv);
    store3(c, ck);
}

// rotational_forward_euler.h:51-61, flat over the 3n components
__global__ void __launch_bounds__(256) euler_particles_kernel(int n3, double* x, double* v, double* th, double* w,
                                                              const double* a, const double* al, double dt, int fixed) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n3) return;
    const double vk = v[k], wk = w[k];
    x[k] = x[k] + (vk * dt + ((0.5 * a[k]) * dt) * dt);  // increment_x(n, v*dt + 0.5*a*dt*dt)
    if (k / 3 != fixed) v[k] = vk + a[k] * dt;
    th[k] = th[k] + (wk * dt + ((0.5 * al[k]) * dt) * dt);
    w[k] = wk + al[k] * dt;
}

// connectors under forward Euler; runs BEFORE euler_particles_kernel (the increments are functions of the old state)
__global__ void __launch_bounds__(256) euler_connectors_kernel(int n, int64_t n_conn, const uint32_t* bptr, double* bconn,
                                                               const double* x, const double* v, const double* w,
                                                               const double* a, const double* al, double dt) {
    const int64_t cq = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (cq >= n_conn) return;
    const uint32_t b = (uint32_t)(cq / 5);
    int lo = 0, hi = n;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (bptr[mid] <= b) lo = mid; else hi = mid - 1;
    }
    const size_t o = 3 * (size_t)lo;
    const d3 dx = load3(v + o) * dt + ((0.5 * load3(a + o)) * dt) * dt;
    const d3 dth = load3(w + o) * dt + ((0.5 * load3(al + o)) * dt) * dt;
    const d3 xn = load3(x + o) + dx;
    double* c = bconn + 3 * (size_t)cq;
    d3 ck = load3(c) + dx;
    const d3 sv = ck - xn;
    ck = ck + cross(dth, sv);
    store3(c, ck);
}

void launch_integrate(const IntegrateArgs& args, cudaStream_t st) {
    if (args.n <= 0) return;
    const int threads = 256;
    const int n3 = 3 * args.n;
    if (args.scheme == 1) {
        if (args.bptr != nullptr && args.n_bonds > 0) {
            const int64_t n_conn = args.n_bonds * 5;
            euler_connectors_kernel<<<(unsigned)((n_conn + threads - 1) / threads), threads, 0, st>>>(
                args.n, n_conn, args.bptr, args.bconn, args.x, args.v, args.w, args.a, args.al, args.dt);
            g_launch_count++;
        }
        euler_particles_kernel<<<(n3 + threads - 1) / threads, threads, 0, st>>>(n3, args.x, args.v, args.th, args.w, args.a,
                                                                               args.al, args.dt, args.fixed);
        g_launch_count++;
        return;
    }
    integrate_particles_kernel<<<(n3 + threads - 1) / threads, threads, 0, st>>>(n3, args.x, args.v, args.th, args.w, args.a,
                                                                               args.al, args.dt, args.fixed);
    g_launch_count++;
    if (args.bptr != nullptr && args.n_bonds > 0) {
        const int64_t n_conn = args.n_bonds * 5;
        integrate_connectors_kernel<<<(unsigned)((n_conn + threads - 1) / threads), threads, 0, st>>>(
            args.n, n_conn, args.bptr, args.bconn, args.x, args.v, args.w, args.dt);
        g_launch_count++;
    }
}

// rotational_velocity_verlet_half.h:51-57: increment_v(n, -a * dt / 2.0), increment_omega(n, -alpha * dt / 2.0)
__global__ void __launch_bounds__(256)
    half_kick_back_kernel(int n3, double* v, double* w, const double* a, const double* al, double dt, int fixed) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n3) return;
    if (k / 3 != fixed) v[k] = v[k] + ((-a[k]) * dt) / 2.0;  // fixed_particle_step_handler drops increment_v of one particle
    w[k] = w[k] + ((-al[k]) * dt) / 2.0;
}

void launch_half_kick_back(int n, double* v, double* w, const double* a, const double* al, double dt, int fixed,
                           cudaStream_t st) {
    if (n <= 0) return;
    const int n3 = 3 * n, threads = 256;
    half_kick_back_kernel<<<(n3 + threads - 1) / threads, threads, 0, st>>>(n3, v, w, a, al, dt, fixed);
    g_launch_count++;
}

}  // namespace lg
