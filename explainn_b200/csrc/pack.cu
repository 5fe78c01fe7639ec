// pack.cu -- one-hot (B,4,L) float -> 2-bit codes + N mask.
// Replaces the U-fold channel tiling of the reference forward (networks.py:153): the tiled input
// is never built; every kernel downstream reads 2 bits per base.
#include "common.cuh"
#include "../../include/explainn_b200.h"

// One warp per 32 consecutive positions of one sequence.
template <typename T>
__global__ void __launch_bounds__(256) pack_onehot_kernel(const T *__restrict__ x, long long B, int L,
                                                          long long bstride, uint32_t *__restrict__ codes,
                                                          uint32_t *__restrict__ nmask, int *__restrict__ n_invalid)
{
    const int wc = xb_words_codes(L), wm = xb_words_mask(L);
    const int lane = threadIdx.x & 31;
    const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const long long b = warp / wm;
    const int chunk = (int)(warp % wm);
    if (b >= B) return;
    const int i = chunk * 32 + lane;
    int code = 0, isn = 0, bad = 0;
    if (i < L) {
        const T *xb = x + b * bstride + i;
        int ones = 0, zeros = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float v = (float)xb[(long long)c * L];
            if (v == 1.0f) { ones++; code = c; }
            else if (v == 0.0f) zeros++;
        }
        if (!(ones == 1 && zeros == 3)) {
            isn = 1;
            code = 0;
            bad = !(ones == 0 && zeros == 4);
        }
    }
    const unsigned m = __ballot_sync(0xffffffffu, isn);
    const unsigned v = (unsigned)code << (2 * (lane & 15));
    const unsigned lo = __reduce_or_sync(0xffffffffu, lane < 16 ? v : 0u);
    const unsigned hi = __reduce_or_sync(0xffffffffu, lane >= 16 ? v : 0u);
    const unsigned anybad = __ballot_sync(0xffffffffu, bad);
    if (lane == 0) {
        nmask[b * wm + chunk] = m;
        codes[b * wc + 2 * chunk] = lo;
        if (2 * chunk + 1 < wc) codes[b * wc + 2 * chunk + 1] = hi;
        if (anybad) atomicAdd(n_invalid, __popc(anybad));
    }
}

extern "C" int explainn_pack_onehot(const void *x, int x_dtype, long long B, int L, long long batch_stride,
                                    uint32_t *codes, uint32_t *nmask, int *n_invalid, cudaStream_t stream)
{
    XB_REQUIRE(x && codes && nmask && n_invalid, "pack_onehot: null pointer");
    XB_REQUIRE(B >= 0 && L > 0 && batch_stride >= 4LL * L, "pack_onehot: bad shape B=%lld L=%d stride=%lld", B, L,
               batch_stride);
    if (B == 0) return XB_OK;
    const long long warps = B * xb_words_mask(L);
    const unsigned grid = (unsigned)((warps + 7) / 8);
    if (x_dtype == EXPLAINN_DTYPE_F32)
        pack_onehot_kernel<float><<<grid, 256, 0, stream>>>((const float *)x, B, L, batch_stride, codes, nmask, n_invalid);
    else if (x_dtype == EXPLAINN_DTYPE_F16)
        pack_onehot_kernel<__half><<<grid, 256, 0, stream>>>((const __half *)x, B, L, batch_stride, codes, nmask, n_invalid);
    else {
        xb_set_error("pack_onehot: unsupported dtype %d", x_dtype);
        return XB_ERR_UNSUPPORTED;
    }
    return xb_check_launch("pack_onehot");
}
