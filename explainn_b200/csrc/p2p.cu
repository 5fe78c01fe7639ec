// p2p.cu -- data-parallel optimizer step as ONE kernel over NVLink peer memory (reduce-scatter + Adam + all-gather).
//
// Replaces, for world sizes 2..8 on one NVSwitch box, the pair "NCCL all-reduce of the flat gradient buffer" +
// "fused Adam" (explainn/train/train.py:65-67 run under DDP semantics).  Rank r owns slice r of the flat parameter
// space: it reads that slice of every rank's gradient straight out of the peers' HBM (plain loads through pointers
// opened with CUDA IPC), sums them in rank order, applies Adam to its slice of its own parameters and moments and
// stores the updated parameters into every peer's parameter buffer.  Per GPU that is (world-1)/world x 4.64 MB of
// NVLink reads and the same amount of writes per step (the all-reduce form of round 1 read (world-1) x 4.64 MB and
// redid the whole Adam on every rank: 104 us at 8 GPUs against 22 us for the single-GPU Adam).  One rank computes
// each parameter, so the replicas stay bit-identical.  The Adam moments of a slice live on its owner only
// (train.FusedTrainer gathers them when the optimizer state is asked for).
//
// Synchronisation is two flag barriers in peer memory, both inside the kernel:
//   A  "my backward of step s is complete and nobody of mine reads my parameters any more" -- written to every peer by
//      block 0 at kernel start (the kernel is stream-ordered after the backward); every block waits for all ranks' A
//      flags before it reads peer gradients or writes peer parameters.
//   B  "my slice is stored everywhere" -- written by the last block of a rank after a system-scope fence; that block
//      then waits for all ranks' B flags, so a rank's kernel (and with it its stream) does not complete before its own
//      parameter buffer is complete and every peer is done with its gradient buffer.
// Every block waits for its own rank's A flags (set by the peers and by this grid's block 0, which never waits for
// another block before setting them), and only the last block waits in B, so the waits cannot deadlock whatever
// the residency.  A wait never traps: like a collective it simply waits for a late peer (validation, a checkpoint
// write, a DataLoader stall).  After `timeout_ns` (default 30 minutes, 0 = no limit) it gives up, raises the error word
// in the local flag buffer (explainn_p2p_error reads it) and lets the kernel finish so that the host can report it.
#include "common.cuh"
#include <string.h>
#include "../../include/explainn_b200.h"

namespace {

constexpr int P2P_MAX_WORLD = 8;

struct P2pArgs {
    float *p[P2P_MAX_WORLD];                       // parameter buffer of every rank (p[rank] is local)
    float *m, *v;
    const float *g[P2P_MAX_WORLD];                 // gradient buffer of every rank (g[rank] is local)
    unsigned long long *flags[P2P_MAX_WORLD];      // flag array of every rank: [2][P2P_MAX_WORLD] u64, ticket, error
    unsigned int *ticket;                          // local block counter
    unsigned int *error;                           // local error word
    const float *step;                             // Adam step count (already incremented)
    long long n;                                   // floats of the range this call updates
    long long off;                                 // its first float inside the flat buffers
    int world, rank;
    unsigned long long seq;                        // step sequence number, identical on all ranks, > 0, increasing
    unsigned long long timeout_ns;
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
__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

__device__ void wait_flags(const unsigned long long *mine, int phase, int world, unsigned long long seq, unsigned long long timeout_ns,
                           unsigned int *error)
{
    unsigned long long t0 = 0;
    for (int q = 0; q < world; ++q) {
        const unsigned long long *f = mine + phase * P2P_MAX_WORLD + q;
        unsigned spins = 0;
        while (ld_sys(f) < seq) {
            __nanosleep(spins < 64 ? 32 : 1000);
            if ((++spins & 1023u) == 0 && timeout_ns) {
                const unsigned long long now = global_ns();
                if (t0 == 0) t0 = now;
                else if (now - t0 > timeout_ns) { atomicExch(error, 1u + (unsigned)q); return; }
            }
        }
    }
    __threadfence_system();
}

__global__ void __launch_bounds__(256) p2p_adam_kernel(const P2pArgs a)
{
    // ---- barrier A; meanwhile a thread of another warp computes the bias corrections (double-precision pow, once per
    //      CTA: the step counter is local, it does not depend on the peers) ----
    __shared__ float s_step_size, s_bc2s;
    if (threadIdx.x == 0) {
        if (blockIdx.x == 0) {
            __threadfence_system();
            for (int q = 0; q < a.world; ++q) st_sys(a.flags[q] + 0 * P2P_MAX_WORLD + a.rank, a.seq);
        }
        wait_flags(a.flags[a.rank], 0, a.world, a.seq, a.timeout_ns, a.error);
    } else if (threadIdx.x == 32) {
        const double t = (double)*a.step;
        const double bc1 = 1.0 - pow((double)a.b1, t), bc2 = 1.0 - pow((double)a.b2, t);
        s_step_size = (float)((double)a.lr / bc1);
        s_bc2s = (float)sqrt(bc2);
    }
    __syncthreads();
    const float step_size = s_step_size, bc2s = s_bc2s;
    const float w1 = 1.0f - a.b1, w2 = 1.0f - a.b2;
    const float scale = 1.0f / (float)a.world;
    const long long n4 = a.n >> 2, o4 = a.off >> 2;      // the flat buffers are 16-byte aligned and padded to multiples of 4
    const long long lo = o4 + n4 * a.rank / a.world, hi = o4 + n4 * (a.rank + 1) / a.world;      // my slice, in float4 units
    float4 *pl = reinterpret_cast<float4 *>(a.p[a.rank]);
    for (long long i = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < hi; i += (long long)gridDim.x * blockDim.x) {
        float4 gs = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int q = 0; q < P2P_MAX_WORLD; ++q) {
            if (q < a.world) {                 // rank order: a fixed summation order
                const float4 gq = __ldcv(reinterpret_cast<const float4 *>(a.g[q]) + i);
                gs.x += gq.x; gs.y += gq.y; gs.z += gq.z; gs.w += gq.w;
            }
        }
        float4 pi = pl[i], mi = reinterpret_cast<float4 *>(a.m)[i], vi = reinterpret_cast<float4 *>(a.v)[i];
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
        reinterpret_cast<float4 *>(a.m)[i] = mi;
        reinterpret_cast<float4 *>(a.v)[i] = vi;
#pragma unroll
        for (int q = 0; q < P2P_MAX_WORLD; ++q)
            if (q < a.world) reinterpret_cast<float4 *>(a.p[q])[i] = pi;      // my slice, stored into every replica
    }

    // ---- barrier B: the last block of this rank tells everybody that its slice is in place, then waits for the peers ----
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int done = atomicAdd(a.ticket, 1u);
        if (done == gridDim.x - 1) {
            *a.ticket = 0u;
            __threadfence_system();
            for (int q = 0; q < a.world; ++q) st_sys(a.flags[q] + 1 * P2P_MAX_WORLD + a.rank, a.seq);
            wait_flags(a.flags[a.rank], 1, a.world, a.seq, a.timeout_ns, a.error);
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

// Error word of a flag buffer (device pointer of THIS rank's buffer): 0 = healthy, 1 + q = the wait for rank q timed
// out in some step.  Synchronous 4-byte device-to-host copy; call it where the host synchronises anyway.
extern "C" int explainn_p2p_error(const unsigned long long *flags, int *out)
{
    XB_REQUIRE(flags && out, "p2p_error: null pointer");
    unsigned int w = 0;
    const cudaError_t e = cudaMemcpy(&w, reinterpret_cast<const unsigned int *>(flags + 2 * P2P_MAX_WORLD) + 2, 4, cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) {
        xb_set_error("p2p_error: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    *out = (int)w;
    return XB_OK;
}

extern "C" int explainn_p2p_adam_step(float *const *params, float *exp_avg, float *exp_avg_sq, long long offset, long long n,
                                      float *step, int tick, const float *const *grads, unsigned long long *const *flags,
                                      int world, int rank, unsigned long long seq, float lr, float beta1, float beta2, float eps,
                                      double timeout_s, cudaStream_t stream)
{
    XB_REQUIRE(params && exp_avg && exp_avg_sq && step && grads && flags && n >= 0 && offset >= 0, "p2p_adam_step: null pointer");
    XB_REQUIRE((offset & 3) == 0, "p2p_adam_step: offset must be a multiple of 4 elements");
    XB_REQUIRE(world >= 1 && world <= P2P_MAX_WORLD && rank >= 0 && rank < world && seq > 0, "p2p_adam_step: world %d rank %d seq %llu",
               world, rank, seq);
    XB_REQUIRE((n & 3) == 0 && (((uintptr_t)exp_avg | (uintptr_t)exp_avg_sq) & 15) == 0,
               "p2p_adam_step: buffers must be 16-byte aligned and padded to a multiple of 4 elements");
    P2pArgs a{};
    a.m = exp_avg; a.v = exp_avg_sq; a.n = n; a.off = offset; a.step = step; a.world = world; a.rank = rank; a.seq = seq;
    a.lr = lr; a.b1 = beta1; a.b2 = beta2; a.eps = eps;
    a.timeout_ns = timeout_s > 0 ? (unsigned long long)(timeout_s * 1e9) : 0ull;
    for (int q = 0; q < world; ++q) {
        XB_REQUIRE(grads[q] && flags[q] && params[q] && (((uintptr_t)grads[q] | (uintptr_t)params[q]) & 15) == 0,
                   "p2p_adam_step: bad peer pointer for rank %d", q);
        a.g[q] = grads[q];
        a.p[q] = params[q];
        a.flags[q] = flags[q];
    }
    // the ticket counter and the error word live behind the two flag rows of this rank's own flag buffer
    a.ticket = reinterpret_cast<unsigned int *>(flags[rank] + 2 * P2P_MAX_WORLD);
    a.error = a.ticket + 2;
    if (tick) p2p_tick_kernel<<<1, 1, 0, stream>>>(step);
    long long blocks = (n / 4 / world + 255) / 256;
    const long long cap = 2LL * xb_num_sms();
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    p2p_adam_kernel<<<(unsigned)blocks, 256, 0, stream>>>(a);
    return xb_check_launch("p2p_adam_step");
}
