// prep.cu -- per-launch preparation that depends only on the weights:
//   * BatchNorm folded into the conv filter and k-mer lookup tables for conv.cu
//   * float16 threshold bracket for the PWM count pass
// Reference semantics: BatchNorm1d eval affine (explainn/models/networks.py:100, eps 1e-5) and
// `0.5 * amax(activations)` on float16 activations (explainn/interpretation/interpretation.py:187).
#include "common.cuh"
#include "../../include/explainn_b200.h"

// grid (tiles, G), 256 threads (one per k-mer).
__global__ void __launch_bounds__(256) fold_tables_kernel(const float *__restrict__ w, const float *__restrict__ bias,
                                                          const float *__restrict__ gamma, const float *__restrict__ beta,
                                                          const float *__restrict__ mean, const float *__restrict__ var,
                                                          float eps, int U, int F, int G, float *__restrict__ tables,
                                                          float *__restrict__ wq_t, float *__restrict__ shift,
                                                          float *__restrict__ scale)
{
    extern __shared__ float wq[];   // [lane][4*F] folded weights of this tile
    __shared__ float sc[XB_UT];
    const int tile = blockIdx.x, g = blockIdx.y, u0 = tile * XB_UT;
    if (threadIdx.x < XB_UT) {
        const int u = u0 + threadIdx.x;
        float s = 0.0f;
        if (u < U) {
            s = __fdiv_rn(gamma[u], __fsqrt_rn(__fadd_rn(var[u], eps)));
            if (g == 0) {
                shift[u] = __fadd_rn(__fmul_rn(__fsub_rn(bias[u], mean[u]), s), beta[u]);
                if (scale) scale[u] = s;
            }
        }
        sc[threadIdx.x] = s;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < XB_UT * 4 * F; i += blockDim.x) {
        const int l = i / (4 * F), r = i % (4 * F);
        const int u = u0 + l;
        wq[i] = u < U ? __fmul_rn(w[(size_t)u * 4 * F + r], sc[l]) : 0.0f;
    }
    __syncthreads();
    if (g == 0) {
        // lane-minor copy for the per-tap path: wq_t[tile][c][k][lane]
        for (int i = threadIdx.x; i < XB_UT * 4 * F; i += blockDim.x) {
            const int l = i & 31, r = i >> 5;   // r = c*F + k
            wq_t[(size_t)tile * 4 * F * XB_UT + i] = wq[l * 4 * F + r];
        }
    }
    const int m = threadIdx.x;   // k-mer: base j at bits [2j, 2j+2)
    float *dst = tables + (((size_t)tile * G + g) * 256 + m) * XB_UT;
    for (int l = 0; l < XB_UT; ++l) {
        float t = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int k = 4 * g + j, c = (m >> (2 * j)) & 3;
            const float a = k < F ? wq[l * 4 * F + c * F + k] : 0.0f;
            t = j == 0 ? a : __fadd_rn(t, a);
        }
        dst[l] = t;
    }
}

extern "C" long long explainn_conv_tables_floats(int U, int F)
{
    const long long tiles = (U + XB_UT - 1) / XB_UT, G = (F + 3) / 4;
    return tiles * G * 256 * XB_UT;
}
extern "C" long long explainn_conv_wq_floats(int U, int F)
{
    const long long tiles = (U + XB_UT - 1) / XB_UT;
    return tiles * 4 * F * XB_UT;
}

extern "C" int explainn_conv_prepare(const float *w, const float *bias, const float *gamma, const float *beta,
                                     const float *mean, const float *var, float eps, int U, int F, float *tables,
                                     float *wq_t, float *shift, float *scale, cudaStream_t stream)
{
    XB_REQUIRE(w && bias && gamma && beta && mean && var && tables && wq_t && shift, "conv_prepare: null pointer");
    XB_REQUIRE(U >= 1 && F >= 1, "conv_prepare: bad sizes U=%d F=%d", U, F);
    if (F > 4 * XB_MAX_G) {
        xb_set_error("conv_prepare: filter_size %d not supported (max %d)", F, 4 * XB_MAX_G);
        return XB_ERR_UNSUPPORTED;
    }
    const int tiles = (U + XB_UT - 1) / XB_UT, G = (F + 3) / 4;
    fold_tables_kernel<<<dim3(tiles, G), 256, XB_UT * 4 * F * sizeof(float), stream>>>(w, bias, gamma, beta, mean, var, eps, U,
                                                                                     F, G, tables, wq_t, shift, scale);
    return xb_check_launch("conv_prepare");
}

// ---- float16 threshold bracket ---------------------------------------------------------------------
// ordered-integer view of floats: monotone map float -> int32
__device__ __forceinline__ int f2ord(float f)
{
    int i = __float_as_int(f);
    return i >= 0 ? i : (int)(0x80000000u - (unsigned)i);
}
__device__ __forceinline__ float ord2f(int o)
{
    return __int_as_float(o >= 0 ? o : (int)(0x80000000u - (unsigned)o));
}

__global__ void interp_thresholds_kernel(const float *__restrict__ zmax, int U, float *__restrict__ thr,
                                         float *__restrict__ zlo, float *__restrict__ zhi, float *__restrict__ amax)
{
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= U) return;
    const float zm = zmax[u];
    // max over (n,s) of half(exp(z)) = half(exp(max z)): both maps are monotone
    const float a16 = fp16_round(det_expf(zm));
    const float t16 = fp16_round(__fmul_rn(0.5f, a16));   // float16 product, as numpy computes it
    amax[u] = a16;
    thr[u] = t16;
    float lo = INFINITY, hi = INFINITY;
    if (t16 == t16 && t16 < INFINITY && zm > -INFINITY) {
        // smallest z with half(exp(z)) > t16, by bisection on the ordered-int representation.
        // pred(-104.5) is false (exp -> 0, and t16 >= 0), pred(89) is true (exp -> inf > finite t16).
        long long a = f2ord(-104.5f), b = f2ord(89.0f);
        while (b - a > 1) {
            const int mid = (int)(a + ((b - a) >> 1));
            const bool pr = fp16_round(det_expf(ord2f(mid))) > t16;
            if (pr) b = mid; else a = mid;
        }
        lo = ord2f((int)b - 64);
        hi = ord2f((int)b + 64);
    }
    zlo[u] = lo;
    zhi[u] = hi;
}

extern "C" int explainn_interp_thresholds(const float *zmax, int U, float *thr, float *zlo, float *zhi, float *amax,
                                          cudaStream_t stream)
{
    XB_REQUIRE(zmax && thr && zlo && zhi && amax && U >= 1, "interp_thresholds: bad argument");
    interp_thresholds_kernel<<<(U + 63) / 64, 64, 0, stream>>>(zmax, U, thr, zlo, zhi, amax);
    return xb_check_launch("interp_thresholds");
}
