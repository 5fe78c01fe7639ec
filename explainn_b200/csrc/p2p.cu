// p2p.cu -- data-parallel optimizer step as ONE kernel over NVLink peer memory.
//
// Replaces, for world sizes 2..8 on one NVSwitch box, the pair "NCCL all-reduce of the flat gradient buffer" +
// "fused Adam" (explainn/train/train.py:65-67 run under DDP semantics): every rank reads the gradient slices of all
// ranks straight out of their HBM (peer pointers opened through CUDA IPC by the Python side), sums them in rank order
// -- the same order on every rank, so the replicas stay bit-identical -- and applies Adam to its full parameter copy.
// Per GPU that is (world-1) x 4.64 MB of NVLink reads (5 us at 2 GPUs, ~36 us at 8) instead of a ring all-reduce
// followed by a second pass over the parameters.
//
// Synchronisation is two flag barriers in peer memory, both inside the kernel:
//   A  "my backward of step s is complete" -- written to every peer by block 0 at kernel start (the kernel is stream-
//      ordered after the backward); every block waits for all ranks' A flags before it reads peer gradients.
//   B  "I have finished reading everybody's gradients of step s" -- written by the last block of a rank; that block then
//      waits for all ranks' B flags, so a rank's kernel (and with it the stream) does not complete before every peer is
//      done with its gradient buffer, which the next backward may then overwrite.
// Blocks never wait for other blocks of their own grid (only the ticket counter), so the waits cannot deadlock
// whatever the residency; a wait that exceeds ~2 s traps instead of hanging the GPU.
#include "common.cuh"
#include <string.h>
#include "../../include/explainn_b200.h"

namespace {

constexpr int P2P_MAX_WORLD = 8;

struct P2pArgs {
    float *p, *m, *v;
    const float *g[P2P_MAX_WORLD];                 // gradient buffer of every rank (g[rank] is local)
    unsigned long long *flags[P2P_MAX_WORLD];      // flag array of every rank: [2][P2P_MAX_WORLD] u64
    unsigned int *ticket;                          // local block counter
    const float *step;                             // Adam step count (already incremented)
    long long n;
    int world, rank;
    unsigned long long seq;                        // step sequence number, identical on all ranks, > 0, increasing
    float lr, b1, b2, eps;
};

__device__ __forceinline__ unsigned long long ld_sys(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_sys(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ void wait_flags(const unsigned long long *mine, int phase, int world, unsigned long long seq)
{
    for (int q = 0; q < world; ++q) {
        const unsigned long long *f = mine + phase * P2P_MAX_WORLD + q;
        long long spins = 0;
        while (ld_sys(f) < seq) {
            __nanosleep(64);
            if (++spins > (1LL << 24)) __trap();       // ~2 s: a peer died; fail loudly instead of hanging
        }
    }
    __threadfence_system();
}

__global__ void __launch_bounds__(256) p2p_adam_kernel(const P2pArgs a)
{
    // ---- barrier A ----
    if (threadIdx.x == 0) {
        if (blockIdx.x == 0) {
            __threadfence_system();
            for (int q = 0; q < a.world; ++q) st_sys(a.flags[q] + 0 * P2P_MAX_WORLD + a.rank, a.seq);
        }
        wait_flags(a.flags[a.rank], 0, a.world, a.seq);
    }
    __syncthreads();

    const double t = (double)*a.step;
    const double bc1 = 1.0 - pow((double)a.b1, t), bc2 = 1.0 - pow((double)a.b2, t);
    const float step_size = (float)((double)a.lr / bc1);
    const float bc2s = (float)sqrt(bc2);
    const float w1 = 1.0f - a.b1, w2 = 1.0f - a.b2;
    const float scale = 1.0f / (float)a.world;
    const long long n4 = a.n >> 2;             // the flat buffers are 16-byte aligned and padded to multiples of 4
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        float4 gs = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int q = 0; q < P2P_MAX_WORLD; ++q) {
            if (q < a.world) {                 // rank order: the same summation order on every rank
                const float4 gq = __ldcv(reinterpret_cast<const float4 *>(a.g[q]) + i);
                gs.x += gq.x; gs.y += gq.y; gs.z += gq.z; gs.w += gq.w;
            }
        }
        float4 pi = reinterpret_cast<float4 *>(a.p)[i], mi = reinterpret_cast<float4 *>(a.m)[i], vi = reinterpret_cast<float4 *>(a.v)[i];
        const float gv[4] = {gs.x * scale, gs.y * scale, gs.z * scale, gs.w * scale};
        float *pp = &pi.x, *mm = &mi.x, *vv = &vi.x;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float mk = fmaf(gv[k] - mm[k], w1, mm[k]);
            const float vk = fmaf(gv[k] * gv[k], w2, vv[k] * a.b2);
            mm[k] = mk;
            vv[k] = vk;
            pp[k] = pp[k] - step_size * (mk / (sqrtf(vk) / bc2s + a.eps));
        }
        reinterpret_cast<float4 *>(a.p)[i] = pi;
        reinterpret_cast<float4 *>(a.m)[i] = mi;
        reinterpret_cast<float4 *>(a.v)[i] = vi;
    }

    // ---- barrier B: the last block of this rank tells everybody that it is done reading, then waits for the peers ----
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned int done = atomicAdd(a.ticket, 1u);
        if (done == gridDim.x - 1) {
            *a.ticket = 0u;
            __threadfence_system();
            for (int q = 0; q < a.world; ++q) st_sys(a.flags[q] + 1 * P2P_MAX_WORLD + a.rank, a.seq);
            wait_flags(a.flags[a.rank], 1, a.world, a.seq);
        }
    }
}

}  // namespace

namespace {
__global__ void p2p_tick_kernel(float *step) { *step += 1.0f; }
}  // namespace

// Peer buffers.  CUDA IPC exports whole cudaMalloc allocations, so the two buffers peers read (flat gradients, flags)
// are the one exception to "PyTorch owns every buffer": they are plain cudaMalloc allocations made here, wrapped into
// torch tensors by the Python side (explainn_b200/parallel.py) and freed by it.
extern "C" int explainn_p2p_alloc(long long bytes, void **out)
{
    XB_REQUIRE(out && bytes > 0, "p2p_alloc: bad argument");
    cudaError_t e = cudaMalloc(out, (size_t)bytes);
    if (e == cudaSuccess) e = cudaMemset(*out, 0, (size_t)bytes);
    if (e != cudaSuccess) {
        xb_set_error("p2p_alloc(%lld): %s", bytes, cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    return XB_OK;
}

extern "C" int explainn_p2p_free(void *p)
{
    const cudaError_t e = p ? cudaFree(p) : cudaSuccess;
    if (e != cudaSuccess) {
        xb_set_error("p2p_free: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    return XB_OK;
}

// handle: 64 bytes (cudaIpcMemHandle_t) that another process of the same node passes to explainn_p2p_open
extern "C" int explainn_p2p_export(const void *p, unsigned char *handle)
{
    XB_REQUIRE(p && handle, "p2p_export: null pointer");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    cudaIpcMemHandle_t h;
    const cudaError_t e = cudaIpcGetMemHandle(&h, const_cast<void *>(p));
    if (e != cudaSuccess) {
        xb_set_error("p2p_export: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    memcpy(handle, &h, sizeof(h));
    return XB_OK;
}

// maps a peer's allocation into this process for the CURRENT device (peer access is enabled by the driver)
extern "C" int explainn_p2p_open(const unsigned char *handle, void **out)
{
    XB_REQUIRE(handle && out, "p2p_open: null pointer");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    const cudaError_t e = cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        xb_set_error("p2p_open: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    return XB_OK;
}

extern "C" int explainn_p2p_close(void *p)
{
    const cudaError_t e = p ? cudaIpcCloseMemHandle(p) : cudaSuccess;
    if (e != cudaSuccess) {
        xb_set_error("p2p_close: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    return XB_OK;
}

extern "C" long long explainn_p2p_flag_bytes(void) { return 2LL * P2P_MAX_WORLD * sizeof(unsigned long long) + 64; }

extern "C" int explainn_p2p_adam_step(float *param, float *exp_avg, float *exp_avg_sq, long long n, float *step,
                                      const float *const *grads, unsigned long long *const *flags, int world, int rank,
                                      unsigned long long seq, float lr, float beta1, float beta2, float eps,
                                      cudaStream_t stream)
{
    XB_REQUIRE(param && exp_avg && exp_avg_sq && step && grads && flags && n >= 0, "p2p_adam_step: null pointer");
    XB_REQUIRE(world >= 1 && world <= P2P_MAX_WORLD && rank >= 0 && rank < world && seq > 0, "p2p_adam_step: world %d rank %d seq %llu",
               world, rank, seq);
    XB_REQUIRE((n & 3) == 0 && (((uintptr_t)param | (uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15) == 0,
               "p2p_adam_step: buffers must be 16-byte aligned and padded to a multiple of 4 elements");
    P2pArgs a{};
    a.p = param; a.m = exp_avg; a.v = exp_avg_sq; a.n = n; a.step = step; a.world = world; a.rank = rank; a.seq = seq;
    a.lr = lr; a.b1 = beta1; a.b2 = beta2; a.eps = eps;
    for (int q = 0; q < world; ++q) {
        XB_REQUIRE(grads[q] && flags[q] && ((uintptr_t)grads[q] & 15) == 0, "p2p_adam_step: bad peer pointer for rank %d", q);
        a.g[q] = grads[q];
        a.flags[q] = flags[q];
    }
    // the ticket counter lives behind the two flag rows of this rank's own flag buffer
    a.ticket = reinterpret_cast<unsigned int *>(flags[rank] + 2 * P2P_MAX_WORLD);
    p2p_tick_kernel<<<1, 1, 0, stream>>>(step);
    long long blocks = (n / 4 + 255) / 256;
    const long long cap = 2LL * xb_num_sms();
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    p2p_adam_kernel<<<(unsigned)blocks, 256, 0, stream>>>(a);
    return xb_check_launch("p2p_adam_step");
}
