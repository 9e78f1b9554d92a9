// kernels_halo.cu -- ghost-particle exchange staging for the slab decomposition (the
// reference has no distributed path, so nothing here replaces reference code).
//
// A rank's local system holds owned particles and ghosts (copies of other ranks' particles near the slab or bonded
// to an owned particle).  After every step the owner's new x, v, omega (72 B) and the connector blocks of the
// ghost's bond records (120 B each) are packed into ONE contiguous send buffer per neighbour rank, moved over
// NVLink (NCCL send/recv), and scattered into the ghost slots on the receiving side.  Index lists are fixed
// between repartitions; they are built by libgran_b200/slab.py.
//   buffer layout: [n_part][9] doubles (x, v, omega)  then  [n_rec][15] doubles (5 connectors)
#include "system.h"

namespace lg {

__global__ void __launch_bounds__(256)
    halo_pack_kernel(int n_part, const int32_t* idx, int n_rec, const int32_t* rec, const double* x, const double* v,
                     const double* w, const double* bconn, double* out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int n9 = 9 * n_part;
    if (t < n9) {
        const int k = t / 9, c = t - 9 * k;
        const size_t src = 3 * (size_t)idx[k] + (c % 3);
        out[t] = c < 3 ? x[src] : (c < 6 ? v[src] : w[src]);
    } else if (t < n9 + 15 * n_rec) {
        const int q = t - n9, k = q / 15, c = q - 15 * k;
        out[t] = bconn[15 * (size_t)rec[k] + c];
    }
}

__global__ void __launch_bounds__(256)
    halo_unpack_kernel(int n_part, const int32_t* idx, int n_rec, const int32_t* rec, double* x, double* v, double* w,
                       double* bconn, const double* in) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    const int n9 = 9 * n_part;
    if (t < n9) {
        const int k = t / 9, c = t - 9 * k;
        const size_t dst = 3 * (size_t)idx[k] + (c % 3);
        const double val = in[t];
        if (c < 3) x[dst] = val; else if (c < 6) v[dst] = val; else w[dst] = val;
    } else if (t < n9 + 15 * n_rec) {
        const int q = t - n9, k = q / 15, c = q - 15 * k;
        bconn[15 * (size_t)rec[k] + c] = in[t];
    }
}

void launch_halo_pack(int n_part, const int32_t* idx, int n_rec, const int32_t* rec, const double* x, const double* v,
                      const double* w, const double* bconn, double* out, cudaStream_t st) {
    const long total = 9L * n_part + 15L * n_rec;
    if (total <= 0) return;
    halo_pack_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(n_part, idx, n_rec, rec, x, v, w, bconn, out);
    g_launch_count++;
}

void launch_halo_unpack(int n_part, const int32_t* idx, int n_rec, const int32_t* rec, double* x, double* v, double* w,
                        double* bconn, const double* in, cudaStream_t st) {
    const long total = 9L * n_part + 15L * n_rec;
    if (total <= 0) return;
    halo_unpack_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(n_part, idx, n_rec, rec, x, v, w, bconn, in);
    g_launch_count++;
}

}  // namespace lg
