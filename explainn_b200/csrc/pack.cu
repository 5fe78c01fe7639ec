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


// Reverse complement of packed sequences: position i of the output is the complement (A<->T, C<->G, i.e. 3 - code)
// of position L-1-i of the input; N stays N.  This is `s[::-1, ::-1]` on the one-hot matrix
// (scripts/fasta2predictions.py:160).  One thread per output code word.
__global__ void __launch_bounds__(256) revcomp_packed_kernel(const uint32_t *__restrict__ codes, const uint32_t *__restrict__ nmask,
                                                             long long B, int L, uint32_t *__restrict__ codes_rc,
                                                             uint32_t *__restrict__ nmask_rc)
{
    const int wc = xb_words_codes(L), wm = xb_words_mask(L);
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= B * wc) return;
    const long long b = t / wc;
    const int w = (int)(t - b * wc);
    const uint32_t *cb = codes + b * wc, *mb = nmask + b * wm;
    uint32_t cw = 0, mbits = 0;
#pragma unroll 4
    for (int k = 0; k < 16; ++k) {
        const int i = w * 16 + k;
        if (i >= L) break;
        const int src = L - 1 - i;
        const uint32_t isn = (mb[src >> 5] >> (src & 31)) & 1u;
        const uint32_t c = (cb[src >> 4] >> (2 * (src & 15))) & 3u;
        if (isn) mbits |= 1u << k;
        else cw |= (3u - c) << (2 * k);
    }
    codes_rc[b * wc + w] = cw;
    // two code words make one mask word
    // mask halves are written with a 16-bit store each (word w covers bits 16*(w&1) .. +15 of mask word w/2)
    reinterpret_cast<unsigned short *>(nmask_rc + b * wm)[w] = (unsigned short)mbits;
}

extern "C" int explainn_revcomp_packed(const uint32_t *codes, const uint32_t *nmask, long long B, int L, uint32_t *codes_rc,
                                       uint32_t *nmask_rc, cudaStream_t stream)
{
    XB_REQUIRE(codes && nmask && codes_rc && nmask_rc && B >= 0 && L > 0, "revcomp_packed: bad argument");
    if (B == 0) return XB_OK;
    // odd number of code words: the upper half of the last mask word is never written by the kernel
    cudaError_t e = cudaMemsetAsync(nmask_rc, 0, (size_t)B * xb_words_mask(L) * 4, stream);
    if (e != cudaSuccess) { xb_set_error("cudaMemsetAsync: %s", cudaGetErrorString(e)); return XB_ERR_CUDA; }
    const long long n = B * xb_words_codes(L);
    revcomp_packed_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(codes, nmask, B, L, codes_rc, nmask_rc);
    return xb_check_launch("revcomp_packed");
}


// ------------------------------------------------------------------------------------------------
// DNA text -> packed codes, on the device (replaces the per-character Python loop of
// explainn/utils/tools.py:136-181, `dna_one_hot`, and the 16-bytes-per-base fp32 host-to-device copy).
// Same semantics as the reference: case-insensitive A/C/G/T -> 0..3; because the reference substitutes the letters
// by the digits '0'..'3' before indexing, those digits in the input mean A..T as well; every other character is an
// all-zero column (N).  A sequence longer than L is centre-cropped (start (len-L)//2), a shorter one is placed at
// column (L-len)//2 and the rest is N.  One warp per 32 consecutive positions of one sequence.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) encode_dna_kernel(const unsigned char *__restrict__ text, const long long *__restrict__ offsets,
                                                         long long B, int L, uint32_t *__restrict__ codes, uint32_t *__restrict__ nmask)
{
    const int wc = xb_words_codes(L), wm = xb_words_mask(L);
    const int lane = threadIdx.x & 31;
    const long long warp = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const long long b = warp / wm;
    const int chunk = (int)(warp % wm);
    if (b >= B) return;
    const long long beg = offsets[b], len = offsets[b + 1] - beg;
    const long long shift = len >= L ? (len - L) / 2 : -(((long long)L - len) / 2);      // source index = i + shift
    const int i = chunk * 32 + lane;
    int code = 0, isn = 0;
    if (i < L) {
        const long long src = i + shift;
        isn = 1;
        if (src >= 0 && src < len) {
            unsigned c = text[beg + src];
            if (c >= 'a' && c <= 'z') c -= 32;
            if (c == 'A' || c == '0') { code = 0; isn = 0; }
            else if (c == 'C' || c == '1') { code = 1; isn = 0; }
            else if (c == 'G' || c == '2') { code = 2; isn = 0; }
            else if (c == 'T' || c == '3') { code = 3; isn = 0; }
        }
    }
    const unsigned m = __ballot_sync(0xffffffffu, isn);
    const unsigned v = (unsigned)code << (2 * (lane & 15));
    const unsigned lo = __reduce_or_sync(0xffffffffu, lane < 16 ? v : 0u);
    const unsigned hi = __reduce_or_sync(0xffffffffu, lane >= 16 ? v : 0u);
    if (lane == 0) {
        nmask[b * wm + chunk] = m;
        codes[b * wc + 2 * chunk] = lo;
        if (2 * chunk + 1 < wc) codes[b * wc + 2 * chunk + 1] = hi;
    }
}

extern "C" int explainn_encode_dna(const unsigned char *text, const long long *offsets, long long B, int L, uint32_t *codes,
                                   uint32_t *nmask, cudaStream_t stream)
{
    XB_REQUIRE(text && offsets && codes && nmask && B >= 0 && L > 0, "encode_dna: bad argument");
    if (B == 0) return XB_OK;
    const long long warps = B * xb_words_mask(L);
    encode_dna_kernel<<<(unsigned)((warps + 7) / 8), 256, 0, stream>>>(text, offsets, B, L, codes, nmask);
    return xb_check_launch("encode_dna");
}

// Packed codes -> one-hot (B,4,L), fp32 or fp16 (the dtype `dna_one_hot` returns); N columns are all zero.
template <typename T>
__global__ void __launch_bounds__(256) unpack_onehot_kernel(const uint32_t *__restrict__ codes, const uint32_t *__restrict__ nmask,
                                                            long long B, int L, T *__restrict__ x)
{
    const int wc = xb_words_codes(L), wm = xb_words_mask(L);
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= B * 4 * L) return;
    const long long b = t / (4LL * L);
    const int rem = (int)(t - b * 4LL * L), c = rem / L, i = rem - c * L;
    const uint32_t isn = (nmask[b * wm + (i >> 5)] >> (i & 31)) & 1u;
    const uint32_t code = (codes[b * wc + (i >> 4)] >> (2 * (i & 15))) & 3u;
    x[t] = (T)((!isn && code == (uint32_t)c) ? 1.0f : 0.0f);
}

extern "C" int explainn_unpack_onehot(const uint32_t *codes, const uint32_t *nmask, long long B, int L, void *x, int x_dtype,
                                      cudaStream_t stream)
{
    XB_REQUIRE(codes && nmask && x && B >= 0 && L > 0, "unpack_onehot: bad argument");
    if (B == 0) return XB_OK;
    const long long n = B * 4 * L;
    const unsigned grid = (unsigned)((n + 255) / 256);
    if (x_dtype == EXPLAINN_DTYPE_F32)
        unpack_onehot_kernel<float><<<grid, 256, 0, stream>>>(codes, nmask, B, L, (float *)x);
    else if (x_dtype == EXPLAINN_DTYPE_F16)
        unpack_onehot_kernel<__half><<<grid, 256, 0, stream>>>(codes, nmask, B, L, (__half *)x);
    else {
        xb_set_error("unpack_onehot: unsupported dtype %d", x_dtype);
        return XB_ERR_UNSUPPORTED;
    }
    return xb_check_launch("unpack_onehot");
}
