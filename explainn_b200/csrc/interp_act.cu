// interp_act.cu -- threshold / count stage of the PWM extraction on activations the CALLER already holds.
//
// What it replaces in the reference (paths relative to /root/reference):
//   get_pwms_explainn(activations, sequences, filter_size) .. explainn/interpretation/interpretation.py:176-227
//     :187      activation_threshold = 0.5 * np.amax(activations, axis=(0, 2))      (in the array's own dtype)
//     :203-214  for every (n, u, s) with activations[n,u,s] > threshold[u]: add sequences[n, :, s:s+F]
// The activations are the (N,U,S) array get_explainn_unit_activations returned (float16 in the reference flow,
// float32 accepted too); nothing is recomputed here, so given the same array the counts are the reference's
// integers bit for bit.  The fused path (conv.cu: interp_unit_max / interp_pwm_count) never builds that array;
// this file exists so that `interpretation.get_pwms_explainn` keeps the reference signature.
//
// Both kernels are plain streaming reads of the activation array (2*U*S bytes per sequence): HBM-bound.
#include "common.cuh"
#include "../../include/explainn_b200.h"

namespace {

constexpr int ACT_WARPS = 8;

__device__ __forceinline__ float act_load(const __half *p) { return __half2float(*p); }
__device__ __forceinline__ float act_load(const float *p) { return *p; }

// order-preserving map float -> uint32 (larger float <=> larger integer); 0 is below every real value
__device__ __forceinline__ uint32_t ord_enc(float f)
{
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord_dec(uint32_t e)
{
    if (e == 0u) return -INFINITY;
    return __uint_as_float((e & 0x80000000u) ? (e & 0x7FFFFFFFu) : ~e);
}

// state[u] = ordered-int maximum of the non-NaN values of unit u, state[U + u] != 0 if a NaN was seen (np.amax
// propagates NaN).  One warp per (n, u) row; a CTA owns a contiguous range of rows and keeps its maxima in shared
// memory until the end.
template <typename T>
__global__ void __launch_bounds__(ACT_WARPS * 32) act_unit_max_kernel(const T *__restrict__ act, long long rows, int U, int S,
                                                                     uint32_t *__restrict__ state)
{
    extern __shared__ uint32_t smax[];       // (2U)
    for (int i = threadIdx.x; i < 2 * U; i += blockDim.x) smax[i] = 0u;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long r0 = rows * blockIdx.x / gridDim.x, r1 = rows * (blockIdx.x + 1) / gridDim.x;
    for (long long r = r0 + warp; r < r1; r += ACT_WARPS) {
        const T *row = act + r * S;
        float m = -INFINITY;
        bool nan = false;
        for (int s = lane; s < S; s += 32) {
            const float v = act_load(row + s);
            nan |= v != v;
            m = fmaxf(m, v);                 // fmaxf drops NaN; the flag carries it
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        nan = __any_sync(0xffffffffu, nan);
        if (lane == 0) {
            const int u = (int)(r % U);
            atomicMax(&smax[u], ord_enc(m));
            if (nan) smax[U + u] = 1u;
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < U; i += blockDim.x) {
        if (smax[i]) atomicMax(&state[i], smax[i]);
        if (smax[U + i]) state[U + i] = 1u;
    }
}

// amax / thr in the activation dtype: thr = 0.5 * amax evaluated in that dtype (numpy: python float * float16 array
// stays float16)
template <typename T>
__global__ void act_thresholds_kernel(const uint32_t *__restrict__ state, int U, T *__restrict__ amax, T *__restrict__ thr)
{
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= U) return;
    float m = ord_dec(state[u]);
    if (state[U + u]) m = __uint_as_float(0x7FC00000u);
    if constexpr (sizeof(T) == 2) {
        const __half h = __float2half_rn(m);          // m came from a float16: exact
        amax[u] = h;
        thr[u] = __hmul(__float2half_rn(0.5f), h);
    } else {
        amax[u] = m;
        thr[u] = __fmul_rn(0.5f, m);
    }
}

// Count stage.  CTA = (tile of ACT_WARPS units, split of the sequences); warp w owns unit u0 + w and walks the
// sequences of the split, so its 4F+1 counters live in shared memory and one hit costs one pass of the warp over
// the F taps (lane = tap).
template <typename T>
__global__ void __launch_bounds__(ACT_WARPS * 32) act_pwm_count_kernel(const T *__restrict__ act, const uint32_t *__restrict__ codes,
                                                                      const uint32_t *__restrict__ nmask, long long N, int L, int U,
                                                                      int S, int F, const T *__restrict__ thr,
                                                                      unsigned long long *__restrict__ counts,
                                                                      unsigned long long *__restrict__ hits,
                                                                      unsigned char *__restrict__ seq_hit)
{
    extern __shared__ uint32_t hist[];       // (ACT_WARPS, 4F+1)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row_len = 4 * F + 1;
    uint32_t *h = hist + warp * row_len;
    for (int i = lane; i < row_len; i += 32) h[i] = 0u;
    __syncwarp();
    const int u = blockIdx.x * ACT_WARPS + warp;
    const long long n0 = N * blockIdx.y / gridDim.y, n1 = N * (blockIdx.y + 1) / gridDim.y;
    const int wc = xb_words_codes(L), wm = xb_words_mask(L);
    if (u < U) {
        const float t = act_load(thr + u);
        for (long long n = n0; n < n1; ++n) {
            const T *row = act + (n * U + u) * S;
            const uint32_t *cw = codes + n * wc, *mw = nmask + n * wm;
            bool any = false;
            for (int s0 = 0; s0 < S; s0 += 32) {
                const int s = s0 + lane;
                const bool hit = s < S && act_load(row + s) > t;          // NaN compares false, as in numpy
                unsigned ball = __ballot_sync(0xffffffffu, hit);
                if (!ball) continue;
                any = true;
                if (lane == 0) h[4 * F] += __popc(ball);
                while (ball) {
                    const int sb = s0 + __ffs(ball) - 1;
                    ball &= ball - 1;
                    for (int k = lane; k < F; k += 32) {
                        const int pos = sb + k;
                        if (pos < L && !((mw[pos >> 5] >> (pos & 31)) & 1u)) {
                            const int c = (cw[pos >> 4] >> (2 * (pos & 15))) & 3;
                            h[c * F + k] += 1u;                             // (c, k) differs per lane: no conflict
                        }
                    }
                    __syncwarp();
                }
            }
            if (seq_hit && lane == 0) seq_hit[n * U + u] = any ? 1 : 0;
        }
        __syncwarp();
        for (int i = lane; i < row_len; i += 32) {
            const uint32_t v = h[i];
            if (!v) continue;
            if (i < 4 * F) atomicAdd(&counts[(long long)u * 4 * F + i], (unsigned long long)v);
            else atomicAdd(&hits[u], (unsigned long long)v);
        }
    }
}

int act_dtype_ok(int dt)
{
    if (dt == EXPLAINN_DTYPE_F16 || dt == EXPLAINN_DTYPE_F32) return 1;
    xb_set_error("activation dtype %d not supported (float16 or float32)", dt);
    return 0;
}

}  // namespace

extern "C" int explainn_act_unit_max(const void *act, int act_dtype, long long N, int U, int S, uint32_t *state,
                                     cudaStream_t stream)
{
    XB_REQUIRE(act && state && N >= 0 && U >= 1 && S >= 1, "act_unit_max: bad argument");
    if (!act_dtype_ok(act_dtype)) return XB_ERR_UNSUPPORTED;
    if (N == 0) return XB_OK;
    const size_t smem = (size_t)2 * U * 4;
    if (smem > 200 * 1024) {
        xb_set_error("act_unit_max: %d units exceed the shared-memory table (max %d)", U, 200 * 1024 / 8);
        return XB_ERR_LIMIT;
    }
    const long long rows = N * U;
    long long grid = 4LL * xb_num_sms();
    if (grid > (rows + ACT_WARPS - 1) / ACT_WARPS) grid = (rows + ACT_WARPS - 1) / ACT_WARPS;
    cudaError_t e;
    if (act_dtype == EXPLAINN_DTYPE_F16) {
        e = cudaFuncSetAttribute(act_unit_max_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) act_unit_max_kernel<__half><<<(unsigned)grid, ACT_WARPS * 32, smem, stream>>>((const __half *)act, rows, U, S, state);
    } else {
        e = cudaFuncSetAttribute(act_unit_max_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) act_unit_max_kernel<float><<<(unsigned)grid, ACT_WARPS * 32, smem, stream>>>((const float *)act, rows, U, S, state);
    }
    if (e != cudaSuccess) {
        xb_set_error("act_unit_max: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    return xb_check_launch("act_unit_max");
}

extern "C" int explainn_act_thresholds(const uint32_t *state, int U, int act_dtype, void *amax, void *thr, cudaStream_t stream)
{
    XB_REQUIRE(state && amax && thr && U >= 1, "act_thresholds: bad argument");
    if (!act_dtype_ok(act_dtype)) return XB_ERR_UNSUPPORTED;
    if (act_dtype == EXPLAINN_DTYPE_F16) act_thresholds_kernel<__half><<<(U + 127) / 128, 128, 0, stream>>>(state, U, (__half *)amax, (__half *)thr);
    else act_thresholds_kernel<float><<<(U + 127) / 128, 128, 0, stream>>>(state, U, (float *)amax, (float *)thr);
    return xb_check_launch("act_thresholds");
}

extern "C" int explainn_act_pwm_count(const void *act, int act_dtype, const uint32_t *codes, const uint32_t *nmask, long long N,
                                      int L, int U, int F, const void *thr, unsigned long long *counts,
                                      unsigned long long *hits, unsigned char *seq_hit, cudaStream_t stream)
{
    XB_REQUIRE(act && codes && nmask && thr && counts && hits && N >= 0 && U >= 1 && F >= 1 && L >= F, "act_pwm_count: bad argument");
    if (!act_dtype_ok(act_dtype)) return XB_ERR_UNSUPPORTED;
    if (N == 0) return XB_OK;
    const int S = L - F + 1;
    const size_t smem = (size_t)ACT_WARPS * (4 * F + 1) * 4;
    if (smem > 200 * 1024) {
        xb_set_error("act_pwm_count: filter_size %d exceeds the shared-memory counters", F);
        return XB_ERR_LIMIT;
    }
    const int tiles = (U + ACT_WARPS - 1) / ACT_WARPS;
    long long split = (4LL * xb_num_sms() + tiles - 1) / tiles;
    if (split > N) split = N;
    if (split > 65535) split = 65535;
    if (split < 1) split = 1;
    dim3 grid((unsigned)tiles, (unsigned)split);
    cudaError_t e;
    if (act_dtype == EXPLAINN_DTYPE_F16) {
        e = cudaFuncSetAttribute(act_pwm_count_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess)
            act_pwm_count_kernel<__half><<<grid, ACT_WARPS * 32, smem, stream>>>((const __half *)act, codes, nmask, N, L, U, S, F, (const __half *)thr,
                                                                                 counts, hits, seq_hit);
    } else {
        e = cudaFuncSetAttribute(act_pwm_count_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess)
            act_pwm_count_kernel<float><<<grid, ACT_WARPS * 32, smem, stream>>>((const float *)act, codes, nmask, N, L, U, S, F, (const float *)thr,
                                                                                counts, hits, seq_hit);
    }
    if (e != cudaSuccess) {
        xb_set_error("act_pwm_count: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    return xb_check_launch("act_pwm_count");
}
