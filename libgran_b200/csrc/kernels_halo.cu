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

// ---- peer-to-peer variant: the pack kernel stores directly into the neighbour's memory over NVLink -------------
// push: one block handles 256 consecutive doubles of ONE peer's message.  After its stores a block fences at system
// scope and bumps the peer's block counter; the block that completes the count publishes the sequence number in
// the peer's memory (release, system scope), so a receiver that observes the flag observes the whole message.
__global__ void __launch_bounds__(256) halo_push_kernel(const __grid_constant__ HaloPushArgs A) {
    int k = 0;
    while (k + 1 < A.n_peers && (int)blockIdx.x >= A.block_begin[k + 1]) k++;
    const HaloSendPeer& P = A.peer[k];
    const int nblocks = A.block_begin[k + 1] - A.block_begin[k];
    const int t = ((int)blockIdx.x - A.block_begin[k]) * 256 + (int)threadIdx.x;
    const int n9 = 9 * P.n_part;
    if (t < n9) {
        const int q = t / 9, c = t - 9 * q;
        const size_t src = 3 * (size_t)P.idx[q] + (c % 3);
        P.remote[t] = c < 3 ? A.x[src] : (c < 6 ? A.v[src] : A.w[src]);
    } else if (t < n9 + 15 * P.n_rec) {
        const int u = t - n9, q = u / 15, c = u - 15 * q;
        P.remote[t] = A.bconn[15 * (size_t)P.rec[q] + c];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int prev = atomicAdd(P.counter, 1u);
        if (prev == (unsigned int)nblocks - 1u) {
            *P.counter = 0u;  // next exchange (stream-ordered after this kernel)
            __threadfence_system();
            asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(P.remote_flag), "l"(A.seq) : "memory");
        }
    }
}

// pull: waits (bounded) until the neighbour has published this exchange's sequence number, then scatters.
__global__ void __launch_bounds__(256) halo_pull_kernel(const __grid_constant__ HaloPullArgs A) {
    int k = 0;
    while (k + 1 < A.n_peers && (int)blockIdx.x >= A.block_begin[k + 1]) k++;
    const HaloRecvPeer& P = A.peer[k];
    if (threadIdx.x == 0) {
        unsigned long long t0, now, seen;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (;;) {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(P.flag) : "memory");
            if (seen >= A.seq) break;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (now - t0 > A.timeout_ns) {
                atomicOr(A.err, ERR_HALO_TIMEOUT);
                break;
            }
            __nanosleep(64);
        }
    }
    __syncthreads();
    const int t = ((int)blockIdx.x - A.block_begin[k]) * 256 + (int)threadIdx.x;
    const int n9 = 9 * P.n_part;
    if (t < n9) {
        const int q = t / 9, c = t - 9 * q;
        const size_t dst = 3 * (size_t)P.idx[q] + (c % 3);
        const double val = __ldcv(P.local + t);
        if (c < 3) A.x[dst] = val; else if (c < 6) A.v[dst] = val; else A.w[dst] = val;
    } else if (t < n9 + 15 * P.n_rec) {
        const int u = t - n9, q = u / 15, c = u - 15 * q;
        A.bconn[15 * (size_t)P.rec[q] + c] = __ldcv(P.local + t);
    }
}

void launch_halo_push(const HaloPushArgs& args, cudaStream_t st) {
    const int blocks = args.block_begin[args.n_peers];
    if (blocks <= 0) return;
    halo_push_kernel<<<(unsigned)blocks, 256, 0, st>>>(args);
    g_launch_count++;
}

void launch_halo_pull(const HaloPullArgs& args, cudaStream_t st) {
    const int blocks = args.block_begin[args.n_peers];
    if (blocks <= 0) return;
    halo_pull_kernel<<<(unsigned)blocks, 256, 0, st>>>(args);
    g_launch_count++;
}

}  // namespace lg
