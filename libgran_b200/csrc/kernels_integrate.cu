// kernels_integrate.cu -- the leap-frog ("velocity Verlet half") update.
//
// Replaces the serial loop of LT/rotational_integrator/rotational_velocity_verlet_half.h:63-73 with the
// default step handler (LT/rotational_step_handler/rotational_step_handler.h:19-68) or, when bonds
// exist, the sinter step handler (LG/sinter_bridge/sinter_bridge.h:24-42,55-75):
//   v += a dt ; x += v dt (updated v) ; omega += alpha dt ; theta += omega dt (updated omega)
//   each connector c of each bond of the particle: c += dx ; c += dtheta x (c - x_new)
// and the first-call half kick back v -= a dt/2, omega -= alpha dt/2 (:51-57).
// Every particle touches only its own rows, so one thread per particle is race free.
#include "system.h"

namespace lg {

__global__ void __launch_bounds__(256) integrate_kernel(const IntegrateArgs A) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= A.n) return;
    const size_t o = 3 * (size_t)i;
    const double dt = A.dt;
    d3 v = load3(A.v + o) + load3(A.a + o) * dt;  // increment_v(n, a * dt)
    store3(A.v + o, v);
    const d3 dx = v * dt;  // increment_x(n, v * dt) with v already updated (:64,69-70)
    const d3 x = load3(A.x + o) + dx;
    store3(A.x + o, x);
    d3 w = load3(A.w + o) + load3(A.al + o) * dt;  // increment_omega
    store3(A.w + o, w);
    const d3 dth = w * dt;  // increment_theta with omega already updated (:66,71-72)
    store3(A.th + o, load3(A.th + o) + dth);
    if (A.bptr != nullptr) {
        const uint32_t b0 = A.bptr[i], b1 = A.bptr[i + 1];
        for (uint32_t b = b0; b < b1; b++) {
            double* c = A.bconn + (size_t)b * 15;
#pragma unroll
            for (int k = 0; k < 5; k++) {
                d3 ck = load3(c + 3 * k) + dx;  // sinter_bridge.h:36-41
                const d3 sv = ck - x;           // :70-71, x already incremented
                ck = ck + cross(dth, sv);
                store3(c + 3 * k, ck);
            }
        }
    }
}

void launch_integrate(const IntegrateArgs& args, cudaStream_t st) {
    if (args.n <= 0) return;
    const int threads = 256;
    integrate_kernel<<<(args.n + threads - 1) / threads, threads, 0, st>>>(args);
    g_launch_count++;
}

// rotational_velocity_verlet_half.h:51-57: increment_v(n, -a * dt / 2.0), increment_omega(n, -alpha * dt / 2.0)
__global__ void __launch_bounds__(256)
    half_kick_back_kernel(int n3, double* v, double* w, const double* a, const double* al, double dt) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n3) return;
    v[k] = v[k] + ((-a[k]) * dt) / 2.0;
    w[k] = w[k] + ((-al[k]) * dt) / 2.0;
}

void launch_half_kick_back(int n, double* v, double* w, const double* a, const double* al, double dt, cudaStream_t st) {
    if (n <= 0) return;
    const int n3 = 3 * n, threads = 256;
    half_kick_back_kernel<<<(n3 + threads - 1) / threads, threads, 0, st>>>(n3, v, w, a, al, dt);
    g_launch_count++;
}

}  // namespace lg
