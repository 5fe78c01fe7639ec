// common.cuh -- shared device helpers for the explainn_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#define XB_HIDDEN 100          // hidden width per unit, hard-coded by the reference (networks.py:106)
#define XB_POOL 7              // MaxPool1d(7, 7): the reference default (networks.py:48,102)
#define XB_UT 32               // units per tile (one lane per unit)
#define XB_MAX_G 6             // k-mer groups supported by the table kernel (filter_size <= 24)

// ---- error plumbing for the C ABI (no exceptions cross the boundary) ------------------------
enum {
    XB_OK = 0,
    XB_ERR_ARG = 1,        // bad argument (null pointer, negative size, unsupported shape)
    XB_ERR_CUDA = 2,       // CUDA runtime error (message in explainn_last_error)
    XB_ERR_UNSUPPORTED = 3,
    XB_ERR_LIMIT = 4       // size beyond a compiled-in maximum
};
void xb_set_error(const char *fmt, ...);
int xb_check_launch(const char *what);
int xb_num_sms();
// conv.cu, shared with interp_tc.cu
int xb_hit_count(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F, const uint32_t *bits, uint32_t *bmask,
                 const unsigned char *flags, unsigned char *seq_hit_out, unsigned long long *counts, unsigned long long *hits,
                 cudaStream_t stream);
long long xb_pwm_count_layout(long long B, int L, int U, int F, long long *off_bmask, long long *off_bits);
// conv_bwd_tc.cu, used by train.cu: slices of `partial` the tensor-core scatter writes (0 = shape not handled); launch
int xb_conv_bwd_tc_slices(long long B, int L, int U, int F, int nsplit, const void *gpool, const void *pooled, const void *xhat);
int xb_conv_bwd_tc(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F, const float *gpool, const float *pooled,
                   const float *xhat, const unsigned char *idx, float *partial, int nsplit, cudaStream_t stream);

#define XB_REQUIRE(cond, ...)                         \
    do {                                              \
        if (!(cond)) {                                \
            xb_set_error(__VA_ARGS__);                \
            return XB_ERR_ARG;                        \
        }                                             \
    } while (0)

// ---- packed sequence format ------------------------------------------------------------------
// codes: 2 bits per base, 16 bases per uint32, base i of a sequence at bits [2*(i%16), +2) of
//        word i/16.  A,C,G,T = 0,1,2,3 (tools.py:160-163).  N / all-zero columns carry code 0.
// nmask: 1 bit per base, 32 per uint32; bit set = all-zero column (tools.py:169-175).
// Bits at positions >= L are zero in both arrays.
__host__ __device__ inline int xb_words_codes(int L) { return (L + 15) >> 4; }
__host__ __device__ inline int xb_words_mask(int L) { return (L + 31) >> 5; }

// ---- deterministic expf ------------------------------------------------------------------------
// Same operation sequence as oracle/explainn_oracle.c::orc_expf, every step IEEE-754 RN, so that
// the GPU and the CPU oracle agree bit for bit (libm / CUDA expf / SLEEF all differ in the last
// ulp, which would leak into the float16 activations the PWM thresholds are computed on).
__device__ __forceinline__ float det_expf(float x)
{
    if (x != x) return x;
    if (x > 88.7228394f) return __int_as_float(0x7f800000);
    if (x < -104.0f) return 0.0f;
    float n = rintf(__fmul_rn(x, 1.44269502f));
    float r = __fmaf_rn(n, -0.693145751953125f, x);
    r = __fmaf_rn(n, -1.42860676533018e-06f, r);
    float p = 1.9875691500e-4f;
    p = __fmaf_rn(p, r, 1.3981999507e-3f);
    p = __fmaf_rn(p, r, 8.3334519073e-3f);
    p = __fmaf_rn(p, r, 4.1665795894e-2f);
    p = __fmaf_rn(p, r, 1.6666665459e-1f);
    p = __fmaf_rn(p, r, 5.0000001201e-1f);
    float rr = __fmul_rn(r, r);
    p = __fmaf_rn(p, rr, r);
    p = __fadd_rn(p, 1.0f);
    int ni = (int)n;
    int n1 = ni >> 1;
    int n2 = ni - n1;
    float s1 = __int_as_float((n1 + 127) << 23);
    float s2 = __int_as_float((n2 + 127) << 23);
    return __fmul_rn(__fmul_rn(p, s1), s2);
}

__device__ __forceinline__ float fp16_round(float x) { return __half2float(__float2half_rn(x)); }

// ---- dropout ---------------------------------------------------------------------------------------
// counter-based dropout mask.  One 32-bit multiply-xorshift hash per QUAD of consecutive elements (keyed with both
// halves of the seed) plus one cheap multiply-xorshift step per further element; element e is kept iff its 32-bit
// value >= p * 2^32.  32-bit arithmetic and the quad sharing are on purpose: the mask generation sits in the inner
// loop of the FC2 training kernel.  Returns the keep bits of elements 4*quad .. 4*quad+3 in bits 0..3.
__device__ __forceinline__ uint32_t drop_keep4(unsigned long long seed, unsigned long long quad, unsigned thresh)
{
    uint32_t x = (uint32_t)quad ^ (uint32_t)seed ^ ((uint32_t)(quad >> 32) * 0x9E3779B9u);
    x ^= x >> 16;
    x *= 0x7FEB352Du;
    x ^= x >> 15;
    x *= 0x846CA68Bu;
    x ^= x >> 16;
    x += (uint32_t)(seed >> 32);
    x *= 0x9E3779B1u;
    x ^= x >> 15;
    uint32_t bits = x >= thresh ? 1u : 0u;
#pragma unroll
    for (int i = 1; i < 4; ++i) {
        x = x * 0x2C1B3C6Du + 0x7F4A7C15u;
        x ^= x >> 15;
        bits |= (x >= thresh ? 1u : 0u) << i;
    }
    return bits;
}

__device__ __forceinline__ float drop_scale(unsigned long long seed, unsigned long long elem, unsigned thresh, float keep_scale)
{
    return ((drop_keep4(seed, elem >> 2, thresh) >> (elem & 3ull)) & 1u) ? keep_scale : 0.0f;
}

// ---- small warp helpers ------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// ---- mbarrier + TMA 1-D bulk copy (global -> shared) ---------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init()
{
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
