// fc.cu -- the per-unit fully-connected stack and the final linear layer, forward.
//
// Replaces (paths relative to /root/reference, explainn/models/networks.py):
//   linears[4..13]: Flatten, Unsqueeze, grouped 1x1 Conv1d(U*P -> 100U), BatchNorm1d(100U), ReLU,
//                   Dropout(0.3), grouped 1x1 Conv1d(100U -> U), BatchNorm1d(U), ReLU, Flatten   (:103-116)
//   final = Linear(U, C)                                                                        (:147,161)
//
// Layout.  The pooled activations arrive as pooled[b][u][p] (the reference's (B,U,P)); a CTA works on the tile
// (one unit x 128 sequences), i.e. 128 rows of P floats U*P apart.  The hidden activations are unit-major,
// h1[u][b][100], so that their tile is one contiguous block.  A CTA keeps W1[u] (100 x P) transposed in shared memory; each thread owns two
// sequences x 25 hidden channels, so every W1 value is fetched by a warp-uniform (broadcast) LDS.128
// and the pooled values by conflict-free scalar loads.
#include "common.cuh"
#include <stdlib.h>
#include "../../include/explainn_b200.h"
#include <algorithm>

// tensor-core variants (fc_tc.cu)
int xb_fc1_fwd_train_tc(const float *pooled, long long B, int U, int P, const float *W1, const float *b1, float *h1, float *part2,
                        cudaStream_t stream);
int xb_fc_stack_fwd_eval_tc(const float *pooled, long long B, int U, int P, const float *W1, const float *b1, const float *g2,
                            const float *be2, const float *rm2, const float *rv2, const float *W2, const float *b2, const float *g3,
                            const float *be3, const float *rm3, const float *rv3, float eps, float *unit_out, cudaStream_t stream);
bool xb_fc_tc_enabled();
int xb_linear_tc(const float *x, long long B, int U, int C, const float *W, const float *bias, float *out, cudaStream_t stream);
int xb_bwd_fc1_tc(const float *h1, const float *gh2, const float *pooled, long long B, int U, int P, const float *g2, const float *be2,
                  const float *mean2, const float *invstd2, const float *W1, const float *W2, float pdrop, unsigned long long seed,
                  const unsigned long long *seed_dev, const uint32_t *dropmask, const float *pb2, float *gpool, float *dW1p, float *dW2,
                  float *db1, float *db2, float *dg2, float *dbe2, cudaStream_t stream);

namespace {

constexpr int H = XB_HIDDEN;      // 100
constexpr int JQ = 4;             // hidden channels split in 4 quarters of 25
constexpr int JW = 25;
constexpr int JP = 28;            // padded quarter (float4 aligned)
constexpr int BM = 128;           // sequences per tile
constexpr int FC_THREADS = 256;

// shared-memory carve-up common to the FC kernels
struct FcSmem {
    float *Ws;   // [P][JQ*JP]   W1[u] transposed, quarter-padded
    float *As;   // [BM][Pp]     pooled tile, row stride Pp (odd)
    int Pp;
    __device__ FcSmem(unsigned char *smem, int P)
    {
        Pp = P | 1;
        Ws = reinterpret_cast<float *>(smem);
        As = Ws + P * JQ * JP;
    }
    static __host__ __device__ size_t bytes(int P) { return ((size_t)P * JQ * JP + (size_t)BM * (P | 1)) * 4; }
};

__device__ __forceinline__ void load_w1(const FcSmem &s, const float *__restrict__ W1u, int P)
{
    // W1u: (100, P) row-major  ->  Ws[p][q*28 + jj]
    for (int i = threadIdx.x; i < H * P; i += blockDim.x) {
        const int j = i / P, p = i - j * P;
        s.Ws[p * (JQ * JP) + (j / JW) * JP + (j % JW)] = W1u[i];
    }
    for (int i = threadIdx.x; i < P * JQ * (JP - JW); i += blockDim.x) {
        const int p = i / (JQ * (JP - JW)), r = i % (JQ * (JP - JW));
        s.Ws[p * (JQ * JP) + (r / (JP - JW)) * JP + JW + (r % (JP - JW))] = 0.0f;
    }
}

__device__ __forceinline__ void load_a(const FcSmem &s, const float *__restrict__ tile, size_t row_stride, int rows, int P)
{
    // tile: `rows` rows of P floats, row_stride elements apart (pooled is (B,U,P): stride U*P)
    for (int i = threadIdx.x; i < BM * P; i += blockDim.x) {
        const int r = i / P, p = i - r * P;
        s.As[r * s.Pp + p] = r < rows ? tile[(size_t)r * row_stride + p] : 0.0f;
    }
}

// acc[r][jj] = sum_p A[row_r][p] * W1[q*25+jj][p]   for the thread's two rows and its quarter q
__device__ __forceinline__ void fc1_tile(const FcSmem &s, int P, int q, int row0, float (&acc)[2][JP])
{
#pragma unroll
    for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int j = 0; j < JP; ++j) acc[r][j] = 0.0f;
    const float *a0 = s.As + row0 * s.Pp;
    const float *a1 = s.As + (row0 + 64) * s.Pp;
    const float4 *w = reinterpret_cast<const float4 *>(s.Ws + q * JP);
#pragma unroll 2
    for (int p = 0; p < P; ++p) {
        const float x0 = a0[p], x1 = a1[p];
#pragma unroll
        for (int v = 0; v < JP / 4; ++v) {
            const float4 ww = w[p * (JQ * JP / 4) + v];
            acc[0][4 * v + 0] = fmaf(x0, ww.x, acc[0][4 * v + 0]);
            acc[0][4 * v + 1] = fmaf(x0, ww.y, acc[0][4 * v + 1]);
            acc[0][4 * v + 2] = fmaf(x0, ww.z, acc[0][4 * v + 2]);
            acc[0][4 * v + 3] = fmaf(x0, ww.w, acc[0][4 * v + 3]);
            acc[1][4 * v + 0] = fmaf(x1, ww.x, acc[1][4 * v + 0]);
            acc[1][4 * v + 1] = fmaf(x1, ww.y, acc[1][4 * v + 1]);
            acc[1][4 * v + 2] = fmaf(x1, ww.z, acc[1][4 * v + 2]);
            acc[1][4 * v + 3] = fmaf(x1, ww.w, acc[1][4 * v + 3]);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Eval-mode fused stack: pooled -> FC1 -> BN2(running) -> ReLU -> FC2 -> BN3(running) -> ReLU
// grid (U, ceil(B/128)), 256 threads.  unit_out: (B,U).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(FC_THREADS) fc_stack_eval_kernel(
    const float *__restrict__ pooled, long long B, int U, int P, const float *__restrict__ W1, const float *__restrict__ b1,
    const float *__restrict__ g2, const float *__restrict__ be2, const float *__restrict__ rm2, const float *__restrict__ rv2,
    const float *__restrict__ W2, const float *__restrict__ b2, const float *__restrict__ g3, const float *__restrict__ be3,
    const float *__restrict__ rm3, const float *__restrict__ rv3, float eps, float *__restrict__ unit_out)
{
    extern __shared__ __align__(16) unsigned char smem[];
    FcSmem s(smem, P);
    __shared__ float sc2[JQ * JP], sh2[JQ * JP], w2s[JQ * JP];
    __shared__ float red[JQ][BM];
    const int u = blockIdx.x;
    const long long b0 = (long long)blockIdx.y * BM;
    const int rows = (int)min((long long)BM, B - b0);
    load_w1(s, W1 + (size_t)u * H * P, P);
    load_a(s, pooled + ((size_t)b0 * U + u) * P, (size_t)U * P, rows, P);
    if (threadIdx.x < JQ * JP) {
        const int q = threadIdx.x / JP, jj = threadIdx.x % JP;
        float sc = 0.f, sh = 0.f, w2 = 0.f;
        if (jj < JW) {
            const int j = u * H + q * JW + jj;
            sc = g2[j] / sqrtf(rv2[j] + eps);
            sh = (b1[j] - rm2[j]) * sc + be2[j];
            w2 = W2[j];
        }
        sc2[threadIdx.x] = sc; sh2[threadIdx.x] = sh; w2s[threadIdx.x] = w2;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, row0 = (warp >> 2) * 32 + lane;
    float acc[2][JP];
    fc1_tile(s, P, q, row0, acc);
    float d0 = 0.f, d1 = 0.f;
#pragma unroll
    for (int j = 0; j < JW; ++j) {
        const float sc = sc2[q * JP + j], sh = sh2[q * JP + j], w2 = w2s[q * JP + j];
        d0 = fmaf(fmaxf(fmaf(acc[0][j], sc, sh), 0.f), w2, d0);
        d1 = fmaf(fmaxf(fmaf(acc[1][j], sc, sh), 0.f), w2, d1);
    }
    red[q][row0] = d0;
    red[q][row0 + 64] = d1;
    __syncthreads();
    if (threadIdx.x < rows) {
        const int r = threadIdx.x;
        float h2 = ((red[0][r] + red[1][r]) + (red[2][r] + red[3][r])) + b2[u];
        const float sc = g3[u] / sqrtf(rv3[u] + eps);
        const float z = (h2 - rm3[u]) * sc + be3[u];
        unit_out[(b0 + r) * U + u] = fmaxf(z, 0.f);
    }
}

// ------------------------------------------------------------------------------------------------
// out[b][c] = sum_u x[b][u] * W[c][u] + bias[c]      (nn.Linear, networks.py:147,161)
// grid ceil(B/32); 256 threads; thread = (row = tid%32, c = tid/32 + 8*i)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) linear_fwd_kernel(const float *__restrict__ x, long long B, int U, int C,
                                                         const float *__restrict__ W, const float *__restrict__ bias,
                                                         float *__restrict__ out)
{
    extern __shared__ float xs[];   // [32][U+1]
    const long long b0 = (long long)blockIdx.x * 32;
    const int rows = (int)min(32LL, B - b0);
    const int Up = U + 1;
    for (int i = threadIdx.x; i < 32 * U; i += blockDim.x) {
        const int r = i / U, c = i - r * U;
        xs[r * Up + c] = r < rows ? x[(b0 + r) * U + c] : 0.f;
    }
    __syncthreads();
    const int r = threadIdx.x & 31;
    // four classes at a time per thread: four independent FMA chains share every x value
    for (int c = threadIdx.x >> 5; c < C; c += 32) {
        const float *w0 = W + (size_t)c * U;
        const bool v1 = c + 8 < C, v2 = c + 16 < C, v3 = c + 24 < C;
        const float *w1 = v1 ? w0 + (size_t)8 * U : w0, *w2 = v2 ? w0 + (size_t)16 * U : w0, *w3 = v3 ? w0 + (size_t)24 * U : w0;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll 4
        for (int k = 0; k < U; ++k) {
            const float xv = xs[r * Up + k];
            a0 = fmaf(xv, __ldg(w0 + k), a0);
            a1 = fmaf(xv, __ldg(w1 + k), a1);
            a2 = fmaf(xv, __ldg(w2 + k), a2);
            a3 = fmaf(xv, __ldg(w3 + k), a3);
        }
        if (r < rows) {
            float *o = out + (b0 + r) * C;
            o[c] = a0 + bias[c];
            if (v1) o[c + 8] = a1 + bias[c + 8];
            if (v2) o[c + 16] = a2 + bias[c + 16];
            if (v3) o[c + 24] = a3 + bias[c + 24];
        }
    }
}


// ================================================================================================
// Training-mode forward (BatchNorm batch statistics, dropout)
// ================================================================================================
constexpr int HS = H + 1;   // row stride of the hidden tile in shared memory (odd: conflict-free columns)

// h1[u][b][:] = W1[u] . pooled[u][b][:] + b1[u]  and per-tile column sums for BatchNorm-2
__global__ void __launch_bounds__(FC_THREADS) fc1_train_kernel(const float *__restrict__ pooled, long long B, int U, int P,
                                                               const float *__restrict__ W1, const float *__restrict__ b1,
                                                               float *__restrict__ h1, float *__restrict__ part2)
{
    extern __shared__ __align__(16) unsigned char smem[];
    FcSmem s(smem, P);
    float *Hs = s.As + BM * s.Pp;
    __shared__ float b1s[H];
    const int u = blockIdx.x, chunk = blockIdx.y, nchunk = gridDim.y;
    const long long b0 = (long long)chunk * BM;
    const int rows = (int)min((long long)BM, B - b0);
    load_w1(s, W1 + (size_t)u * H * P, P);
    load_a(s, pooled + ((size_t)b0 * U + u) * P, (size_t)U * P, rows, P);
    if (threadIdx.x < H) b1s[threadIdx.x] = b1[u * H + threadIdx.x];
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = warp & 3, row0 = (warp >> 2) * 32 + lane;
    float acc[2][JP];
    fc1_tile(s, P, q, row0, acc);
#pragma unroll
    for (int j = 0; j < JW; ++j) {
        Hs[row0 * HS + q * JW + j] = acc[0][j] + b1s[q * JW + j];
        Hs[(row0 + 64) * HS + q * JW + j] = acc[1][j] + b1s[q * JW + j];
    }
    __syncthreads();
    float *dst = h1 + ((size_t)u * B + b0) * H;
    for (int i = threadIdx.x; i < rows * H; i += blockDim.x) {
        const int r = i / H, j = i - r * H;
        dst[i] = Hs[r * HS + j];
    }
    if (threadIdx.x < H) {
        float sm = 0.f, sq = 0.f;
        for (int r = 0; r < rows; ++r) {
            const float v = Hs[r * HS + threadIdx.x];
            sm += v;
            sq = fmaf(v, v, sq);
        }
        float *pp = part2 + ((size_t)u * nchunk + chunk) * 2 * H;
        pp[threadIdx.x] = sm;
        pp[H + threadIdx.x] = sq;
    }
}

// BatchNorm-2 (batch stats) -> ReLU -> dropout -> FC2 (+b2) ; per-tile sums for BatchNorm-3
constexpr int F2_THREADS = 128;        // fc2_train_kernel: one thread per sequence of a 128-row tile
__global__ void __launch_bounds__(F2_THREADS) fc2_train_kernel(
    const float *__restrict__ h1, long long B, int U, const float *__restrict__ part2, int nchunk,
    const float *__restrict__ g2, const float *__restrict__ be2, float *__restrict__ rm2, float *__restrict__ rv2,
    float momentum, float eps, const float *__restrict__ W2, const float *__restrict__ b2, float *__restrict__ mean2,
    float *__restrict__ invstd2, float pdrop, unsigned long long seed0, const unsigned long long *__restrict__ seed_dev,
    uint32_t *__restrict__ dropmask, float *__restrict__ h2, float *__restrict__ part3, double inv_batch)
{
    __shared__ __align__(16) float sc[128], sh[128], w2s[128];
    __shared__ float red[F2_THREADS / 32][2];
    const unsigned long long seed = seed0 + (seed_dev ? *seed_dev * 0xD1B54A32D192ED03ull : 0ull);
    const int u = blockIdx.x, chunk = blockIdx.y;
    const long long b0 = (long long)chunk * BM;
    const int rows = (int)min((long long)BM, B - b0);
    {
        const int j = threadIdx.x;
        float a = 0.f, c = 0.f, w = 0.f;
        if (j < H) {
            double sm = 0.0, sq = 0.0;
            const double inv_B = (double)inv_batch;
            // statistics count = 1 / inv_batch: the local batch, or the whole data-parallel batch when the partial sums were
            // all-reduced (synchronised BatchNorm)
            const double nstat = 1.0 / inv_B;
            const float unbias = (float)(nstat / (nstat > 1.5 ? nstat - 1.0 : 1.0));
            for (int k = 0; k < nchunk; ++k) {
                const float *pp = part2 + ((size_t)u * nchunk + k) * 2 * H;
                sm += (double)pp[j];
                sq += (double)pp[H + j];
            }
            // only the cancellation-prone difference E[x^2] - mean^2 is formed in double (multiplications by the
            // reciprocal batch size, no double division or square root in this per-CTA prologue)
            const double mean = sm * inv_B;
            double var = sq * inv_B - mean * mean;
            if (var < 0.0) var = 0.0;
            const float varf = (float)var;
            const float inv = 1.0f / sqrtf(varf + eps);
            const int ch = u * H + j;
            if (chunk == 0) {
                mean2[ch] = (float)mean;
                invstd2[ch] = inv;
                if (rm2) {
                    const float unb = B > 1 ? varf * unbias : varf;
                    rm2[ch] = (1.0f - momentum) * rm2[ch] + momentum * (float)mean;
                    rv2[ch] = (1.0f - momentum) * rv2[ch] + momentum * unb;
                }
            }
            a = g2[ch] * inv;
            c = be2[ch] - (float)mean * a;
            w = W2[ch];
        }
        sc[j] = a; sh[j] = c; w2s[j] = w;
    }
    __syncthreads();
    // thread = one sequence of the tile: its 100 hidden values are read as 25 quads (pairs of quads = whole 32-byte
    // sectors), BatchNorm-2 -> ReLU -> dropout -> dot with W2 accumulated in a register.  No shuffles, no idle lanes; the
    // per-channel coefficients come from shared memory as broadcasts.
    static_assert(H % 4 == 0 && F2_THREADS == BM, "one thread per row, quads of channels");
    const unsigned thresh = pdrop > 0.f ? (unsigned)min(4294967295.0, (double)pdrop * 4294967296.0) : 0u;
    const float keep_scale = pdrop > 0.f ? 1.0f / (1.0f - pdrop) : 1.0f;
    const int r = threadIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool valid = r < rows;
    const size_t rowi = (size_t)u * B + b0 + (valid ? r : 0);
    const float4 *hp = reinterpret_cast<const float4 *>(h1 + rowi * H);
    float dot = 0.f;
    uint32_t mw[4] = {0u, 0u, 0u, 0u};
    auto quad = [&](int q, const float4 hc) {
        const float4 a = *reinterpret_cast<const float4 *>(&sc[4 * q]), c = *reinterpret_cast<const float4 *>(&sh[4 * q]);
        const float4 w = *reinterpret_cast<const float4 *>(&w2s[4 * q]);
        uint32_t nib = 15u;
        if (pdrop > 0.f) nib = drop_keep4(seed, rowi * (H / 4) + q, thresh);
        const float v0 = (nib & 1u) ? fmaxf(fmaf(hc.x, a.x, c.x), 0.f) * keep_scale : 0.f;
        const float v1 = (nib & 2u) ? fmaxf(fmaf(hc.y, a.y, c.y), 0.f) * keep_scale : 0.f;
        const float v2 = (nib & 4u) ? fmaxf(fmaf(hc.z, a.z, c.z), 0.f) * keep_scale : 0.f;
        const float v3 = (nib & 8u) ? fmaxf(fmaf(hc.w, a.w, c.w), 0.f) * keep_scale : 0.f;
        dot += fmaf(v3, w.w, fmaf(v2, w.z, fmaf(v1, w.y, v0 * w.x)));
        mw[q >> 3] |= nib << (4 * (q & 7));
    };
    if (valid) {
        // three pairs of quads in flight ahead of the arithmetic
        constexpr int NQ = H / 4, NP = NQ / 2;
        float4 buf[6];
#pragma unroll
        for (int i = 0; i < 6; ++i) buf[i] = hp[i];
#pragma unroll
        for (int pq = 0; pq < NP; ++pq) {
            const float4 x0 = buf[(2 * pq) % 6], x1 = buf[(2 * pq + 1) % 6];
            if (2 * pq + 6 < NQ) buf[(2 * pq) % 6] = hp[2 * pq + 6];
            if (2 * pq + 7 < NQ) buf[(2 * pq + 1) % 6] = hp[2 * pq + 7];
            quad(2 * pq, x0);
            quad(2 * pq + 1, x1);
        }
        if (NQ & 1) quad(NQ - 1, buf[(NQ - 1) % 6]);
    }
    float s3 = 0.f, q3 = 0.f;
    if (valid) {
        const float o = dot + b2[u];
        h2[rowi] = o;
        s3 = o;
        q3 = o * o;
        // the keep bits are saved for the backward: word w of the row holds channels 32w .. 32w+31 (8 quads' nibbles)
        if (pdrop > 0.f && dropmask) *reinterpret_cast<uint4 *>(dropmask + rowi * 4) = make_uint4(mw[0], mw[1], mw[2], mw[3]);
    }
    s3 = warp_sum(s3);
    q3 = warp_sum(q3);
    if (lane == 0) { red[warp][0] = s3; red[warp][1] = q3; }
    __syncthreads();
    if (threadIdx.x < 2) {
        float v = 0.f;
        for (int w = 0; w < F2_THREADS / 32; ++w) v += red[w][threadIdx.x];
        part3[((size_t)u * nchunk + chunk) * 2 + threadIdx.x] = v;
    }
}

// BatchNorm-3 (batch stats) -> ReLU -> unit outputs (B,U) -> final Linear -> logits (B,C)
// grid ceil(B/8): 8 sequences per CTA; warp w owns row w, lanes run over the classes.
constexpr int OT_ROWS = 8;
__global__ void __launch_bounds__(256) out_train_kernel(const float *__restrict__ h2, long long B, int U, int C,
                                                        const float *__restrict__ part3, int nchunk,
                                                        const float *__restrict__ g3, const float *__restrict__ be3,
                                                        float *__restrict__ rm3, float *__restrict__ rv3, float momentum,
                                                        float eps, float *__restrict__ mean3, float *__restrict__ invstd3,
                                                        const float *__restrict__ Wf, const float *__restrict__ bf,
                                                        float *__restrict__ unit, float *__restrict__ logits, int wf_in_smem, double inv_batch)
{
    extern __shared__ float xs[];   // [8][U] then sc3[U], sh3[U], then (when it fits) Wf as [C][U+1]
    float *sc3 = xs + OT_ROWS * U, *sh3 = sc3 + U;
    float *wfs = sh3 + U;
    // blockIdx.y selects a slice of the classes, so that more SMs share the work and each CTA stages less of Wf
    const int cper = (C + gridDim.y - 1) / gridDim.y;
    const int c_lo = blockIdx.y * cper, c_hi = min(C, c_lo + cper);
    if (wf_in_smem)
        for (int c = c_lo + (threadIdx.x >> 5); c < c_hi; c += blockDim.x >> 5)
            for (int k = threadIdx.x & 31; k < U; k += 32) wfs[(c - c_lo) * (U + 1) + k] = Wf[(size_t)c * U + k];
    const long long b0 = (long long)blockIdx.x * OT_ROWS;
    const int rows = (int)min((long long)OT_ROWS, B - b0);
    for (int u = threadIdx.x; u < U; u += blockDim.x) {
        double sm = 0.0, sq = 0.0;
        for (int k = 0; k < nchunk; ++k) {
            sm += (double)part3[((size_t)u * nchunk + k) * 2];
            sq += (double)part3[((size_t)u * nchunk + k) * 2 + 1];
        }
        const double mean = sm * inv_batch;
        double var = sq * inv_batch - mean * mean;       // the only double arithmetic: a cancellation-prone difference
        if (var < 0.0) var = 0.0;
        const float varf = (float)var;
        const float inv = 1.0f / sqrtf(varf + eps);
        if (blockIdx.x == 0 && blockIdx.y == 0) {
            mean3[u] = (float)mean;
            invstd3[u] = inv;
            if (rm3) {
                const double nstat = 1.0 / inv_batch;
                const float unb = nstat > 1.5 ? varf * (float)(nstat / (nstat - 1.0)) : varf;
                rm3[u] = (1.0f - momentum) * rm3[u] + momentum * (float)mean;
                rv3[u] = (1.0f - momentum) * rv3[u] + momentum * unb;
            }
        }
        const float a = g3[u] * inv;
        sc3[u] = a;
        sh3[u] = be3[u] - (float)mean * a;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp < rows) {
        for (int u = lane; u < U; u += 32) {
            const float v = fmaxf(fmaf(h2[(size_t)u * B + b0 + warp], sc3[u], sh3[u]), 0.f);
            if (blockIdx.y == 0) unit[(b0 + warp) * U + u] = v;
            xs[warp * U + u] = v;
        }
    }
    __syncthreads();
    if (warp < rows) {
        const float *xr = xs + warp * U;
        for (int c = c_lo + lane; c < c_hi; c += 32) {
            // lanes run over the classes: in shared memory the row stride U+1 keeps them on different banks
            const float *w = wf_in_smem ? wfs + (size_t)(c - c_lo) * (U + 1) : Wf + (size_t)c * U;
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
            int k = 0;
            for (; k + 3 < U; k += 4) {
                a0 = fmaf(xr[k], w[k], a0);
                a1 = fmaf(xr[k + 1], w[k + 1], a1);
                a2 = fmaf(xr[k + 2], w[k + 2], a2);
                a3 = fmaf(xr[k + 3], w[k + 3], a3);
            }
            for (; k < U; ++k) a0 = fmaf(xr[k], w[k], a0);
            logits[(b0 + warp) * C + c] = ((a0 + a1) + (a2 + a3)) + bf[c];
        }
    }
}

// ================================================================================================
// Backward
// ================================================================================================

// gz3[b][u] = (unit[b][u] > 0) * sum_c gl[b][c] * Wf[c][u]     (ReLU-3 mask applied; (B,U) layout)
// grid (ceil(B/8), ceil(U/64)); warp = one sequence, each lane two units of the 64-unit slice whose Wf columns
// sit in shared memory.  Wf == NULL: gl already is d/d unit.
constexpr int GU_UT = 64;
__global__ void __launch_bounds__(256) bwd_gunit_kernel(const float *__restrict__ gl, const float *__restrict__ Wf,
                                                        const float *__restrict__ unit, long long B, int U, int C,
                                                        float *__restrict__ gz)
{
    extern __shared__ float gsm[];   // gls [8][C], then wfs [C][64]
    float *gls = gsm, *wfs = gsm + 8 * C;
    const long long b0 = (long long)blockIdx.x * 8;
    const int u0 = blockIdx.y * GU_UT;
    const int rows = (int)min(8LL, B - b0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (Wf != nullptr) {
        for (int i = threadIdx.x; i < rows * C; i += blockDim.x) gls[i] = gl[b0 * C + i];
        // eight loads in flight per thread (issued before the stores: one exposed latency per batch, not per element)
        for (int i0 = threadIdx.x; i0 < C * GU_UT; i0 += 8 * 256) {
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int i = i0 + 256 * j, c = i / GU_UT, uu = i - c * GU_UT;
                v[j] = (i < C * GU_UT && u0 + uu < U) ? Wf[(size_t)c * U + u0 + uu] : 0.f;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j)
                if (i0 + 256 * j < C * GU_UT) wfs[i0 + 256 * j] = v[j];
        }
        __syncthreads();
    }
    if (warp >= rows) return;
    const long long b = b0 + warp;
    const int ua = u0 + lane, ub = ua + 32;
    float ga = 0.f, gb = 0.f;
    if (Wf != nullptr) {
        const float *gr = gls + warp * C;
        float a0 = 0.f, a1 = 0.f, c0 = 0.f, c1 = 0.f;
        int c = 0;
        for (; c + 1 < C; c += 2) {
            const float g0 = gr[c], g1 = gr[c + 1];
            a0 = fmaf(g0, wfs[c * GU_UT + lane], a0);
            c0 = fmaf(g0, wfs[c * GU_UT + lane + 32], c0);
            a1 = fmaf(g1, wfs[(c + 1) * GU_UT + lane], a1);
            c1 = fmaf(g1, wfs[(c + 1) * GU_UT + lane + 32], c1);
        }
        if (c < C) {
            a0 = fmaf(gr[c], wfs[c * GU_UT + lane], a0);
            c0 = fmaf(gr[c], wfs[c * GU_UT + lane + 32], c0);
        }
        ga = a0 + a1;
        gb = c0 + c1;
    } else {
        if (ua < U) ga = gl[b * U + ua];
        if (ub < U) gb = gl[b * U + ub];
    }
    if (ua < U) gz[b * U + ua] = unit[b * U + ua] > 0.f ? ga : 0.f;
    if (ub < U) gz[b * U + ub] = unit[b * U + ub] > 0.f ? gb : 0.f;
}

// Register-tiled dWf / dbf:  dWf[c][u] += sum_b gl[b][c] * unit[b][u].
// grid (ceil(U/128), ceil(C/8), nsplit): a CTA owns 128 units x 8 classes and one slice of the batch; its 4 warps take every
// fourth sequence of the slice (lane l: units u0+l+32j), keep 32 accumulators each, and are combined through shared
// memory before ONE atomicAdd per output and CTA.  The gl tile of the slice is staged class-minor in shared memory.
constexpr int FL_SLICE = 64;
__global__ void __launch_bounds__(128) bwd_final_linear2_kernel(const float *__restrict__ gl, const float *__restrict__ unit,
                                                                long long B, int U, int C, float *__restrict__ dWf,
                                                                float *__restrict__ dbf)
{
    __shared__ __align__(16) float gls[FL_SLICE][8];
    __shared__ float red[4][8][128];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int u0 = blockIdx.x * 128, c0 = blockIdx.y * 8;
    const long long per = (B + gridDim.z - 1) / gridDim.z;
    const long long bs = per * blockIdx.z, be = min(B, bs + per);
    float acc[8][4];
#pragma unroll
    for (int c = 0; c < 8; ++c)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[c][j] = 0.f;
    float sg = 0.f;
    for (long long t0 = bs; t0 < be; t0 += FL_SLICE) {
        const int nb = (int)min((long long)FL_SLICE, be - t0);
        __syncthreads();
        for (int i = threadIdx.x; i < nb * 8; i += 128) {
            const int bb = i >> 3, c = i & 7;
            gls[bb][c] = c0 + c < C ? gl[(t0 + bb) * C + c0 + c] : 0.f;
        }
        __syncthreads();
        if (threadIdx.x < 8)
            for (int bb = 0; bb < nb; ++bb) sg += gls[bb][threadIdx.x];
        for (int bb = warp; bb < nb; bb += 4) {
            const float *ur = unit + (t0 + bb) * U + u0 + lane;
            float x[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) x[j] = u0 + lane + 32 * j < U ? ur[32 * j] : 0.f;
            const float4 ga = *reinterpret_cast<const float4 *>(&gls[bb][0]), gb = *reinterpret_cast<const float4 *>(&gls[bb][4]);
            const float g[8] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w};
#pragma unroll
            for (int c = 0; c < 8; ++c)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[c][j] = fmaf(g[c], x[j], acc[c][j]);
        }
    }
#pragma unroll
    for (int c = 0; c < 8; ++c)
#pragma unroll
        for (int j = 0; j < 4; ++j) red[warp][c][lane + 32 * j] = acc[c][j];
    __syncthreads();
    if (dWf) {
        const int uu = threadIdx.x, u = u0 + uu;
        if (u < U)
            for (int c = 0; c < 8 && c0 + c < C; ++c)
                atomicAdd(&dWf[(size_t)(c0 + c) * U + u], (red[0][c][uu] + red[1][c][uu]) + (red[2][c][uu] + red[3][c][uu]));
    }
    if (dbf && blockIdx.x == 0 && threadIdx.x < 8 && c0 + threadIdx.x < C) atomicAdd(&dbf[c0 + threadIdx.x], sg);
}

// BatchNorm-3 backward per unit: sums over the batch, then gh2 (U,B), dgamma3, dbeta3.  One warp per unit.
// mode 0: both halves; mode 1: only the sums, written to sums[2u], sums[2u+1]; mode 2: only the second half, with the
// sums read back from there (synchronised BatchNorm: the caller averages them over the ranks in between).
__global__ void __launch_bounds__(128) bwd_bn3_kernel(const float *__restrict__ gz, const float *__restrict__ h2, long long B,
                                                      int U, const float *__restrict__ g3, const float *__restrict__ mean3,
                                                      const float *__restrict__ invstd3, float *__restrict__ gh2,
                                                      float *__restrict__ dg3, float *__restrict__ db3, float *__restrict__ sums,
                                                      int mode)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int u = blockIdx.x * 4 + warp;
    if (u >= U) return;
    const float mean = mean3[u], inv = invstd3[u];
    const float *h2u = h2 + (size_t)u * B;
    float s1 = 0.f, s2 = 0.f;
    if (mode != 2) {
        // eight rows per trip, all loads of a trip issued before the arithmetic (the warp is alone on its unit: without
        // this the loop is one memory latency per 32 rows)
        for (long long b0 = 0; b0 < B; b0 += 256) {
            float g[8], x[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const long long b = b0 + 32 * k + lane;
                g[k] = b < B ? gz[b * U + u] : 0.f;
                x[k] = b < B ? h2u[b] : mean;
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const float xh = (x[k] - mean) * inv;
                s1 += g[k];
                s2 = fmaf(g[k], xh, s2);
            }
        }
        s1 = warp_sum(s1);
        s2 = warp_sum(s2);
        if (mode == 1) {
            if (lane == 0) { sums[2 * u] = s1; sums[2 * u + 1] = s2; }
            return;
        }
    } else {
        s1 = sums[2 * u];
        s2 = sums[2 * u + 1];
    }
    if (lane == 0) {
        if (db3) db3[u] = s1;
        if (dg3) dg3[u] = s2;
    }
    const float m1 = s1 / (float)B, m2 = s2 / (float)B, k = g3[u] * inv;
    for (long long b0 = 0; b0 < B; b0 += 256) {
        float g[8], x[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const long long b = b0 + 32 * i + lane;
            g[i] = b < B ? gz[b * U + u] : 0.f;
            x[i] = b < B ? h2u[b] : mean;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const long long b = b0 + 32 * i + lane;
            const float xh = (x[i] - mean) * inv;
            if (b < B) gh2[(size_t)u * B + b] = k * (g[i] - m1 - xh * m2);
        }
    }
}

// FC2 backward + ReLU/dropout mask + BatchNorm-2 backward sums, per tile
__global__ void __launch_bounds__(FC_THREADS) bwd_fc2_kernel(const float *__restrict__ h1, const float *__restrict__ gh2,
                                                             long long B, int U, const float *__restrict__ g2,
                                                             const float *__restrict__ be2, const float *__restrict__ mean2,
                                                             const float *__restrict__ invstd2, const float *__restrict__ W2,
                                                             float pdrop, unsigned long long seed0,
                                                             const unsigned long long *__restrict__ seed_dev,
                                                             const uint32_t *__restrict__ dropmask, float *__restrict__ pb2)
{
    __shared__ float mu[128], inv[128], ga[128], be[128], w2s[128];
    __shared__ float red[8][3][128];
    const unsigned long long seed = seed0 + (seed_dev ? *seed_dev * 0xD1B54A32D192ED03ull : 0ull);
    const int u = blockIdx.x, chunk = blockIdx.y, nchunk = gridDim.y;
    const long long b0 = (long long)chunk * BM;
    const int rows = (int)min((long long)BM, B - b0);
    if (threadIdx.x < 128) {
        const int j = threadIdx.x, ch = u * H + j;
        const bool v = j < H;
        mu[j] = v ? mean2[ch] : 0.f; inv[j] = v ? invstd2[ch] : 0.f; ga[j] = v ? g2[ch] : 0.f;
        be[j] = v ? be2[ch] : 0.f; w2s[j] = v ? W2[ch] : 0.f;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned thresh = pdrop > 0.f ? (unsigned)min(4294967295.0, (double)pdrop * 4294967296.0) : 0u;
    const float keep_scale = pdrop > 0.f ? 1.0f / (1.0f - pdrop) : 1.0f;
    // lane l < 25 owns hidden channels 4l .. 4l+3 (one 16-byte load per row); coefficients stay in registers
    const bool act = lane < H / 4;
    const int j0 = act ? 4 * lane : 0;
    float muv[4], inv4[4], gav[4], bev[4], w2v[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        muv[i] = act ? mu[j0 + i] : 0.f; inv4[i] = act ? inv[j0 + i] : 0.f; gav[i] = act ? ga[j0 + i] : 0.f;
        bev[i] = act ? be[j0 + i] : 0.f; w2v[i] = act ? w2s[j0 + i] : 0.f;
    }
    float dW2[4] = {0, 0, 0, 0}, sg[4] = {0, 0, 0, 0}, sgx[4] = {0, 0, 0, 0};
    const bool use_mask = pdrop > 0.f && dropmask;
    // two register sets (X, Y), each written in place by its loads and never copied: a register copy of a load
    // destination waits for the load (such a copy was 46 % of this kernel's stall samples)
    float4 hX = make_float4(0.f, 0.f, 0.f, 0.f), hY = hX;
    float gX = 0.f, gY = 0.f;
    uint32_t wX = 0u, wY = 0u;
    auto fetch = [&](int r, float4 &hh, float &gg, uint32_t &ww) {
        if (r < rows) {
            const size_t rowi = (size_t)u * B + b0 + r;
            if (act) hh = *reinterpret_cast<const float4 *>(h1 + rowi * H + j0);
            gg = gh2[rowi];
            if (use_mask) ww = dropmask[rowi * 4 + (lane >> 3)];
        }
    };
    auto row = [&](int r, const float4 &hh, float g, uint32_t ww) {
        const size_t rowi = (size_t)u * B + b0 + r;
        const float hc[4] = {hh.x, hh.y, hh.z, hh.w};
        uint32_t nib = 15u;
        if (pdrop > 0.f) nib = use_mask ? (ww >> (4 * (lane & 7))) & 15u : drop_keep4(seed, rowi * (H / 4) + lane, thresh);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float xh = (hc[i] - muv[i]) * inv4[i];
            const float zb = fmaf(gav[i], xh, bev[i]);
            const float ks = ((nib >> i) & 1u) ? keep_scale : 0.f;
            const float a2 = fmaxf(zb, 0.f) * ks;
            dW2[i] = fmaf(g, a2, dW2[i]);
            const float gz2 = zb > 0.f ? g * w2v[i] * ks : 0.f;
            sg[i] += gz2;
            sgx[i] = fmaf(gz2, xh, sgx[i]);
        }
    };
    fetch(warp, hX, gX, wX);
    fetch(warp + 8, hY, gY, wY);
    for (int r = warp; r < rows; r += 16) {      // rows r, r+8, ... of this warp; each is requested two rows ahead
        row(r, hX, gX, wX);
        fetch(r + 16, hX, gX, wX);
        if (r + 8 < rows) row(r + 8, hY, gY, wY);
        fetch(r + 24, hY, gY, wY);
    }
    if (act) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            red[warp][0][j0 + i] = dW2[i];
            red[warp][1][j0 + i] = sg[i];
            red[warp][2][j0 + i] = sgx[i];
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 3 * H; i += blockDim.x) {
        const int t = i / H, j = i - t * H;
        float v = 0.f;
        for (int w = 0; w < 8; ++w) v += red[w][t][j];
        pb2[(((size_t)u * nchunk + chunk) * 3 + t) * H + j] = v;
    }
}

// BatchNorm-2 backward -> gh1 tile -> d(pooled) = gh1 . W1 and dW1 partial = gh1^T . pooled
template <int PH>
__global__ void __launch_bounds__(FC_THREADS) bwd_fc1_kernel(
    const float *__restrict__ h1, const float *__restrict__ gh2, const float *__restrict__ pooled, long long B, int U, int P,
    const float *__restrict__ g2, const float *__restrict__ be2, const float *__restrict__ mean2, const float *__restrict__ invstd2,
    const float *__restrict__ W1, const float *__restrict__ W2, float pdrop, unsigned long long seed0,
    const unsigned long long *__restrict__ seed_dev, const float *__restrict__ pb2, float *__restrict__ gpool, float *__restrict__ dW1p, float *__restrict__ dW2,
    float *__restrict__ db1, float *__restrict__ db2, float *__restrict__ dg2, float *__restrict__ dbe2)
{
    constexpr int PQ = 2 * PH;
    extern __shared__ __align__(16) unsigned char smem[];
    float *Gs = reinterpret_cast<float *>(smem);   // [BM][HS]
    float *As = Gs + BM * HS;                       // [BM][PQ]   (BM*HS = 12928 floats: 16-byte aligned)
    float *Wr = As + BM * PQ;                       // [H][PQ]
    __shared__ float mu[128], inv[128], ga[128], be[128], w2s[128], m1s[128], m2s[128];
    const unsigned long long seed = seed0 + (seed_dev ? *seed_dev * 0xD1B54A32D192ED03ull : 0ull);
    const int u = blockIdx.x, chunk = blockIdx.y, nchunk = gridDim.y;
    const long long b0 = (long long)chunk * BM;
    const int rows = (int)min((long long)BM, B - b0);
    if (threadIdx.x < 128) {
        const int j = threadIdx.x, ch = u * H + j;
        const bool v = j < H;
        float a = 0.f, s1 = 0.f, s2 = 0.f;
        if (v) {
            for (int k = 0; k < nchunk; ++k) {
                const float *pp = pb2 + ((size_t)u * nchunk + k) * 3 * H;
                a += pp[j]; s1 += pp[H + j]; s2 += pp[2 * H + j];
            }
            if (chunk == 0) {
                if (dW2) dW2[ch] = a;
                if (dbe2) dbe2[ch] = s1;
                if (dg2) dg2[ch] = s2;
                if (db1) db1[ch] = 0.f;          // bias followed by train-mode BatchNorm: gradient is exactly zero
                if (db2 && j == 0) db2[u] = 0.f;
            }
        }
        mu[j] = v ? mean2[ch] : 0.f; inv[j] = v ? invstd2[ch] : 0.f; ga[j] = v ? g2[ch] : 0.f;
        be[j] = v ? be2[ch] : 0.f; w2s[j] = v ? W2[ch] : 0.f;
        m1s[j] = s1 / (float)B; m2s[j] = s2 / (float)B;
    }
    // pooled tile and W1[u], both zero-padded to PQ columns
    for (int i = threadIdx.x; i < BM * PQ; i += blockDim.x) {
        const int r = i / PQ, p = i - r * PQ;
        As[i] = (r < rows && p < P) ? pooled[((size_t)(b0 + r) * U + u) * P + p] : 0.f;
    }
    for (int i = threadIdx.x; i < H * PQ; i += blockDim.x) {
        const int j = i / PQ, p = i - j * PQ;
        Wr[i] = p < P ? W1[((size_t)u * H + j) * P + p] : 0.f;
    }
    __syncthreads();
    const unsigned thresh = pdrop > 0.f ? (unsigned)min(4294967295.0, (double)pdrop * 4294967296.0) : 0u;
    const float keep_scale = pdrop > 0.f ? 1.0f / (1.0f - pdrop) : 1.0f;
    for (int i = threadIdx.x; i < BM * H; i += blockDim.x) {
        const int r = i / H, j = i - r * H;
        float v = 0.f;
        if (r < rows) {
            const size_t rowi = (size_t)u * B + b0 + r;
            const float xh = (h1[rowi * H + j] - mu[j]) * inv[j];
            const float zb = fmaf(ga[j], xh, be[j]);
            float gz2 = 0.f;
            if (zb > 0.f) {
                const float ks = pdrop > 0.f ? drop_scale(seed, rowi * H + j, thresh, keep_scale) : 1.0f;
                gz2 = gh2[rowi] * w2s[j] * ks;
            }
            v = ga[j] * inv[j] * (gz2 - m1s[j] - xh * m2s[j]);
        }
        Gs[r * HS + j] = v;
    }
    __syncthreads();
    const int t = threadIdx.x & 127, ph = threadIdx.x >> 7;
    float acc[PH];
    // (a) d(pooled)[b][p] = sum_j G[b][j] * W1[j][p]
#pragma unroll
    for (int i = 0; i < PH; ++i) acc[i] = 0.f;
    for (int j = 0; j < H; ++j) {
        const float g = Gs[t * HS + j];
        const float4 *w = reinterpret_cast<const float4 *>(Wr + j * PQ + ph * PH);
#pragma unroll
        for (int v = 0; v < PH / 4; ++v) {
            const float4 ww = w[v];
            acc[4 * v + 0] = fmaf(g, ww.x, acc[4 * v + 0]);
            acc[4 * v + 1] = fmaf(g, ww.y, acc[4 * v + 1]);
            acc[4 * v + 2] = fmaf(g, ww.z, acc[4 * v + 2]);
            acc[4 * v + 3] = fmaf(g, ww.w, acc[4 * v + 3]);
        }
    }
    if (t < rows) {
        float *dst = gpool + ((size_t)(b0 + t) * U + u) * P;
#pragma unroll
        for (int i = 0; i < PH; ++i)
            if (ph * PH + i < P) dst[ph * PH + i] = acc[i];
    }
    // (b) dW1[j][p] partial = sum_b G[b][j] * pooled[b][p]
#pragma unroll
    for (int i = 0; i < PH; ++i) acc[i] = 0.f;
    if (t < H) {
        for (int b = 0; b < BM; ++b) {
            const float g = Gs[b * HS + t];
            const float4 *a = reinterpret_cast<const float4 *>(As + b * PQ + ph * PH);
#pragma unroll
            for (int v = 0; v < PH / 4; ++v) {
                const float4 aa = a[v];
                acc[4 * v + 0] = fmaf(g, aa.x, acc[4 * v + 0]);
                acc[4 * v + 1] = fmaf(g, aa.y, acc[4 * v + 1]);
                acc[4 * v + 2] = fmaf(g, aa.z, acc[4 * v + 2]);
                acc[4 * v + 3] = fmaf(g, aa.w, acc[4 * v + 3]);
            }
        }
        float *dst = dW1p + (((size_t)chunk * U + u) * H + t) * P;
#pragma unroll
        for (int i = 0; i < PH; ++i)
            if (ph * PH + i < P) dst[ph * PH + i] = acc[i];
    }
}

template <int PH>
size_t bwd_fc1_smem() { return ((size_t)BM * HS + (size_t)BM * 2 * PH + (size_t)H * 2 * PH) * 4; }

// ================================================================================================
// num_fc = 1 variant of the unit (explainn/models/networks.py:82-95): conv -> BN -> exp -> pool -> per-unit
// FC (P -> 1) -> BatchNorm -> ReLU.  h2[u][b] = b1[u] + sum_p W[u][p] * pooled[b][u][p]; everything after it
// (BatchNorm over the batch, ReLU, final Linear and their backward) reuses the kernels of the default stack.
// grid (U, nchunk), 128 threads: thread = one sequence of the 128-row tile.
// ================================================================================================
template <bool EVAL>
__global__ void __launch_bounds__(BM) fc1x1_fwd_kernel(const float *__restrict__ pooled, long long B, int U, int P,
                                                       const float *__restrict__ W, const float *__restrict__ b1,
                                                       const float *__restrict__ g, const float *__restrict__ be,
                                                       const float *__restrict__ rm, const float *__restrict__ rv, float eps,
                                                       float *__restrict__ h2, float *__restrict__ part3, float *__restrict__ unit_out)
{
    extern __shared__ float ws1[];      // W[u][0..P)
    __shared__ float red[2][4];
    const int u = blockIdx.x, chunk = blockIdx.y, nchunk = gridDim.y;
    for (int p = threadIdx.x; p < P; p += blockDim.x) ws1[p] = W[(size_t)u * P + p];
    __syncthreads();
    const long long b = (long long)chunk * BM + threadIdx.x;
    float h = 0.f;
    if (b < B) {
        const float *pr = pooled + ((size_t)b * U + u) * P;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        int p = 0;
        for (; p + 3 < P; p += 4) {
            a0 = fmaf(ws1[p], pr[p], a0);
            a1 = fmaf(ws1[p + 1], pr[p + 1], a1);
            a2 = fmaf(ws1[p + 2], pr[p + 2], a2);
            a3 = fmaf(ws1[p + 3], pr[p + 3], a3);
        }
        for (; p < P; ++p) a0 = fmaf(ws1[p], pr[p], a0);
        h = ((a0 + a1) + (a2 + a3)) + b1[u];
        if (EVAL) {
            const float sc = g[u] / sqrtf(rv[u] + eps);
            unit_out[b * U + u] = fmaxf((h - rm[u]) * sc + be[u], 0.f);
        } else {
            h2[(size_t)u * B + b] = h;
        }
    }
    if (!EVAL) {
        const float hv = b < B ? h : 0.f;
        const float s = warp_sum(hv), q = warp_sum(hv * hv);
        const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
        if (lane == 0) { red[0][warp] = s; red[1][warp] = q; }
        __syncthreads();
        if (threadIdx.x < 2) {
            const float *r = red[threadIdx.x];
            part3[((size_t)u * nchunk + chunk) * 2 + threadIdx.x] = (r[0] + r[1]) + (r[2] + r[3]);
        }
    }
}

// backward of the P -> 1 FC: gpool[b][u][p] = gh2[u][b] * W[u][p]; dW partial[chunk][u][p] = sum_b gh2[u][b] * pooled[b][u][p]
__global__ void __launch_bounds__(BM) fc1x1_bwd_kernel(const float *__restrict__ pooled, const float *__restrict__ gh2, long long B,
                                                       int U, int P, const float *__restrict__ W, float *__restrict__ gpool,
                                                       float *__restrict__ dWp)
{
    extern __shared__ float sm1[];      // W[u][0..P), then gh[128], then the pooled tile [128][P+1]
    float *ws1 = sm1, *ghs = sm1 + P, *tile = ghs + BM;
    const int u = blockIdx.x, chunk = blockIdx.y;
    const long long b0 = (long long)chunk * BM;
    const int rows = (int)min((long long)BM, B - b0);
    for (int p = threadIdx.x; p < P; p += blockDim.x) ws1[p] = W[(size_t)u * P + p];
    ghs[threadIdx.x] = threadIdx.x < rows ? gh2[(size_t)u * B + b0 + threadIdx.x] : 0.f;
    __syncthreads();
    // rows of P floats: consecutive threads on consecutive addresses
    const int n = rows * P;
    for (int i = threadIdx.x; i < n; i += BM) {
        const int r = i / P, p = i - r * P;
        const size_t o = ((size_t)(b0 + r) * U + u) * P + p;
        tile[r * (P + 1) + p] = pooled[o];
        gpool[o] = ghs[r] * ws1[p];
    }
    __syncthreads();
    for (int p = threadIdx.x; p < P; p += BM) {
        float a = 0.f;
        for (int r = 0; r < rows; ++r) a = fmaf(ghs[r], tile[r * (P + 1) + p], a);
        dWp[((size_t)chunk * U + u) * P + p] = a;
    }
}

}  // namespace

extern "C" int explainn_fc_stack_fwd_eval(const float *pooled, long long B, int U, int P, const float *W1, const float *b1,
                                          const float *bn2_w, const float *bn2_b, const float *bn2_rm, const float *bn2_rv,
                                          const float *W2, const float *b2, const float *bn3_w, const float *bn3_b,
                                          const float *bn3_rm, const float *bn3_rv, float eps, float *unit_out,
                                          cudaStream_t stream)
{
    XB_REQUIRE(pooled && W1 && b1 && bn2_w && bn2_b && bn2_rm && bn2_rv && W2 && b2 && bn3_w && bn3_b && bn3_rm && bn3_rv && unit_out,
               "fc_stack_fwd_eval: null pointer");
    XB_REQUIRE(B >= 0 && U >= 1 && P >= 1, "fc_stack_fwd_eval: bad sizes");
    if (B == 0) return XB_OK;
    if (xb_fc_tc_enabled())      // tcgen05 path (XB_FC_TC=0 selects the CUDA-core kernel below)
        return xb_fc_stack_fwd_eval_tc(pooled, B, U, P, W1, b1, bn2_w, bn2_b, bn2_rm, bn2_rv, W2, b2, bn3_w, bn3_b, bn3_rm, bn3_rv, eps,
                                       unit_out, stream);
    const size_t smem = FcSmem::bytes(P);
    if (smem > 200 * 1024) {
        xb_set_error("fc_stack_fwd_eval: pooled length %d needs %zu bytes of shared memory", P, smem);
        return XB_ERR_UNSUPPORTED;
    }
    cudaFuncSetAttribute(fc_stack_eval_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    dim3 grid(U, (unsigned)((B + BM - 1) / BM));
    fc_stack_eval_kernel<<<grid, FC_THREADS, smem, stream>>>(pooled, B, U, P, W1, b1, bn2_w, bn2_b, bn2_rm, bn2_rv, W2, b2, bn3_w,
                                                            bn3_b, bn3_rm, bn3_rv, eps, unit_out);
    return xb_check_launch("fc_stack_fwd_eval");
}

extern "C" int explainn_linear_fwd(const float *x, long long B, int U, int C, const float *W, const float *bias, float *out,
                                   cudaStream_t stream)
{
    XB_REQUIRE(x && W && bias && out && B >= 0 && U >= 1 && C >= 1, "linear_fwd: bad argument");
    if (B == 0) return XB_OK;
    if (xb_fc_tc_enabled()) {    // tcgen05 path (XB_FC_TC=0 selects the CUDA-core kernel below)
        const int rc = xb_linear_tc(x, B, U, C, W, bias, out, stream);
        if (rc != XB_ERR_UNSUPPORTED) return rc;
    }
    const size_t smem = (size_t)32 * (U + 1) * 4;
    if (smem > 200 * 1024) {
        xb_set_error("linear_fwd: %d input features need %zu bytes of shared memory", U, smem);
        return XB_ERR_UNSUPPORTED;
    }
    cudaFuncSetAttribute(linear_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    linear_fwd_kernel<<<(unsigned)((B + 31) / 32), 256, smem, stream>>>(x, B, U, C, W, bias, out);
    return xb_check_launch("linear_fwd");
}

// ---- training entry points --------------------------------------------------------------------------

extern "C" int explainn_fc_nchunk(long long B) { return (int)((B + BM - 1) / BM); }

// phases: 1 = FC1 (h1, part2), 2 = BatchNorm-2 .. FC2 (h2, part3), 4 = BatchNorm-3 .. logits; batch_total = number of
// samples behind the partial sums (B, or the whole data-parallel batch once the caller has all-reduced part2 / part3)
static int fc_stack_fwd_train_impl(int phases, long long batch_total, const float *pooled, long long B, int U, int P, int C,
                                   const float *W1, const float *b1, const float *bn2_w, const float *bn2_b, float *bn2_rm,
                                   float *bn2_rv, const float *W2, const float *b2, const float *bn3_w, const float *bn3_b,
                                   float *bn3_rm, float *bn3_rv, float momentum, float eps, float pdrop, unsigned long long seed,
                                   const unsigned long long *seed_dev, uint32_t *dropmask, const float *Wf, const float *bf,
                                   float *h1, float *part2, float *mean2, float *invstd2, float *h2, float *part3, float *mean3,
                                   float *invstd3, float *unit, float *logits, cudaStream_t stream)
{
    XB_REQUIRE(pooled && W1 && b1 && bn2_w && bn2_b && W2 && b2 && bn3_w && bn3_b && Wf && bf && h1 && part2 && mean2 && invstd2 &&
                   h2 && part3 && mean3 && invstd3 && unit && logits,
               "fc_stack_fwd_train: null pointer");
    XB_REQUIRE(B >= 2 && U >= 1 && P >= 1 && C >= 1, "fc_stack_fwd_train: bad sizes (train-mode BatchNorm needs B >= 2)");
    XB_REQUIRE(pdrop >= 0.f && pdrop < 1.f, "fc_stack_fwd_train: dropout probability %f", pdrop);
    XB_REQUIRE(((uintptr_t)h1 & 15) == 0, "fc_stack_fwd_train: h1 must be 16-byte aligned");
    const int nchunk = (int)((B + BM - 1) / BM);
    const size_t smem1 = FcSmem::bytes(P) + (size_t)BM * HS * 4;
    size_t smem3 = ((size_t)OT_ROWS * U + 2 * (size_t)U) * 4;
    const int ot_ny = C >= 48 ? 3 : 1;           // class slices per row group
    const size_t smem3_wf = smem3 + (size_t)((C + ot_ny - 1) / ot_ny) * (U + 1) * 4;
    const int wf_in_smem = smem3_wf <= 200 * 1024;
    if (wf_in_smem) smem3 = smem3_wf;
    if (smem1 > 220 * 1024 || smem3 > 220 * 1024) {
        xb_set_error("fc_stack_fwd_train: shared memory need (%zu / %zu bytes) exceeds the SM", smem1, smem3);
        return XB_ERR_UNSUPPORTED;
    }
    XB_REQUIRE(batch_total >= B && phases >= 1 && phases <= 7, "fc_stack_fwd_train: bad phase arguments");
    const double inv_batch = 1.0 / (double)batch_total;
    if (phases & 1) {
        if (xb_fc_tc_enabled()) {    // tcgen05 path (XB_FC_TC=0 selects the CUDA-core kernel)
            int rc = xb_fc1_fwd_train_tc(pooled, B, U, P, W1, b1, h1, part2, stream);
            if (rc != XB_OK) return rc;
        } else {
            cudaFuncSetAttribute(fc1_train_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1);
            fc1_train_kernel<<<dim3(U, nchunk), FC_THREADS, smem1, stream>>>(pooled, B, U, P, W1, b1, h1, part2);
        }
    }
    if (phases & 2)
        fc2_train_kernel<<<dim3(U, nchunk), F2_THREADS, 0, stream>>>(h1, B, U, part2, nchunk, bn2_w, bn2_b, bn2_rm, bn2_rv, momentum, eps,
                                                                     W2, b2, mean2, invstd2, pdrop, seed, seed_dev, dropmask, h2, part3, inv_batch);
    if (phases & 4) {
        cudaFuncSetAttribute(out_train_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem3);
        out_train_kernel<<<dim3((unsigned)((B + OT_ROWS - 1) / OT_ROWS), ot_ny), 256, smem3, stream>>>(h2, B, U, C, part3, nchunk, bn3_w, bn3_b, bn3_rm, bn3_rv,
                                                                            momentum, eps, mean3, invstd3, Wf, bf, unit, logits, wf_in_smem, inv_batch);
    }
    return xb_check_launch("fc_stack_fwd_train");
}

extern "C" int explainn_fc_stack_fwd_train(const float *pooled, long long B, int U, int P, int C, const float *W1, const float *b1,
                                           const float *bn2_w, const float *bn2_b, float *bn2_rm, float *bn2_rv, const float *W2,
                                           const float *b2, const float *bn3_w, const float *bn3_b, float *bn3_rm, float *bn3_rv,
                                           float momentum, float eps, float pdrop, unsigned long long seed,
                                           const unsigned long long *seed_dev, uint32_t *dropmask, const float *Wf, const float *bf, float *h1, float *part2, float *mean2, float *invstd2, float *h2,
                                           float *part3, float *mean3, float *invstd3, float *unit, float *logits,
                                           cudaStream_t stream)
{
    return fc_stack_fwd_train_impl(7, B, pooled, B, U, P, C, W1, b1, bn2_w, bn2_b, bn2_rm, bn2_rv, W2, b2, bn3_w, bn3_b, bn3_rm, bn3_rv,
                                   momentum, eps, pdrop, seed, seed_dev, dropmask, Wf, bf, h1, part2, mean2, invstd2, h2, part3, mean3,
                                   invstd3, unit, logits, stream);
}

// The same forward in separately callable phases (bitmask: 1 FC1, 2 BatchNorm-2 .. FC2, 4 BatchNorm-3 .. logits), for
// data-parallel training with synchronised BatchNorm statistics: the caller all-reduces (SUM) part2 between phases 1
// and 2 and part3 between 2 and 4 and passes batch_total = the number of samples of ALL ranks.
extern "C" int explainn_fc_stack_fwd_train_phase(int phases, long long batch_total, const float *pooled, long long B, int U, int P,
                                                 int C, const float *W1, const float *b1, const float *bn2_w, const float *bn2_b,
                                                 float *bn2_rm, float *bn2_rv, const float *W2, const float *b2, const float *bn3_w,
                                                 const float *bn3_b, float *bn3_rm, float *bn3_rv, float momentum, float eps,
                                                 float pdrop, unsigned long long seed, const unsigned long long *seed_dev,
                                                 uint32_t *dropmask, const float *Wf, const float *bf, float *h1, float *part2,
                                                 float *mean2, float *invstd2, float *h2, float *part3, float *mean3, float *invstd3,
                                                 float *unit, float *logits, cudaStream_t stream)
{
    return fc_stack_fwd_train_impl(phases, batch_total, pooled, B, U, P, C, W1, b1, bn2_w, bn2_b, bn2_rm, bn2_rv, W2, b2, bn3_w, bn3_b,
                                   bn3_rm, bn3_rv, momentum, eps, pdrop, seed, seed_dev, dropmask, Wf, bf, h1, part2, mean2, invstd2, h2,
                                   part3, mean3, invstd3, unit, logits, stream);
}

// Backward of the final Linear alone (networks.py:147,161), for the units without an FC stack (num_fc = 0):
// dWf (C,U) = glogits^T unit, dbf (C) = column sums of glogits, gunit (B,U) = glogits Wf masked where unit <= 0
// (unit outputs are post-ReLU or post-exp, so the mask only ever hides entries whose gradient is dead anyway).
extern "C" int explainn_linear_bwd(const float *glogits, const float *unit, const float *Wf, long long B, int U, int C, float *gunit,
                                   float *dWf, float *dbf, cudaStream_t stream)
{
    XB_REQUIRE(glogits && unit && Wf && gunit && B >= 1 && U >= 1 && C >= 1, "linear_bwd: bad argument");
    if (dWf) cudaMemsetAsync(dWf, 0, (size_t)C * U * sizeof(float), stream);
    if (dbf) cudaMemsetAsync(dbf, 0, (size_t)C * sizeof(float), stream);
    if (dWf || dbf) {
        const int nsb = (int)std::min<long long>(8, (B + 63) / 64);
        bwd_final_linear2_kernel<<<dim3((U + 127) / 128, (C + 7) / 8, nsb), 128, 0, stream>>>(glogits, unit, B, U, C, dWf, dbf);
    }
    const size_t smem_gu = ((size_t)8 * C + (size_t)C * GU_UT) * 4;
    if (smem_gu > 200 * 1024) {
        xb_set_error("linear_bwd: %d classes need %zu bytes of shared memory", C, smem_gu);
        return XB_ERR_LIMIT;
    }
    cudaFuncSetAttribute(bwd_gunit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_gu);
    bwd_gunit_kernel<<<dim3((unsigned)((B + 7) / 8), (U + GU_UT - 1) / GU_UT), 256, smem_gu, stream>>>(glogits, Wf, unit, B, U, C, gunit);
    return xb_check_launch("linear_bwd");
}

// One side stream per device for kernels that are independent of the main chain of a call (created on first use, kept
// until the process ends: the library's only long-lived CUDA objects besides the static scratch of the no-alloc entry).
struct SideStream { cudaStream_t stream; cudaEvent_t fork, join; };
static SideStream *side_stream_for_current_device()
{
    static SideStream table[64];
    static bool ready[64];
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    if (!__atomic_load_n(&ready[dev], __ATOMIC_ACQUIRE)) {
        SideStream t;
        if (cudaStreamCreateWithFlags(&t.stream, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&t.fork, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        if (cudaEventCreateWithFlags(&t.join, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        table[dev] = t;          // two racing threads would each create a set; the later store wins, the other set leaks once
        __atomic_store_n(&ready[dev], true, __ATOMIC_RELEASE);
    }
    return &table[dev];
}

// phases: 1 = final-Linear gradients, d/d unit, BatchNorm-3 sums; 2 = BatchNorm-3 backward, FC2 backward + BatchNorm-2 sums
// (pb2); 4 = FC1 backward + dW1 reduction.  phases == 7 runs the fused BatchNorm-3 kernel; otherwise the BatchNorm-3 sums
// travel through the first 2U floats of pb2 between phases 1 and 2 (the caller may average them over ranks there).
static int fc_stack_bwd_impl(int phases, const float *glogits, long long B, int U, int P, int C, const float *pooled, const float *h1,
                             const float *h2, const float *unit, const float *W1, const float *bn2_w, const float *bn2_b,
                             const float *mean2, const float *invstd2, const float *W2, const float *bn3_w, const float *mean3,
                             const float *invstd3, const float *Wf, float pdrop, unsigned long long seed,
                             const unsigned long long *seed_dev, const uint32_t *dropmask, float *gh2, float *pb2,
                             float *dW1_partials, float *gpool, float *dW1, float *db1, float *dbn2_w, float *dbn2_b, float *dW2,
                             float *db2, float *dbn3_w, float *dbn3_b, float *dWf, float *dbf, cudaStream_t stream)
{
    XB_REQUIRE(phases >= 1 && phases <= 7, "fc_stack_bwd: bad phase mask");
    XB_REQUIRE(glogits && pooled && h1 && h2 && unit && W1 && bn2_w && bn2_b && mean2 && invstd2 && W2 && bn3_w && mean3 && invstd3 &&
                   gh2 && pb2 && dW1_partials && gpool && dW1,
               "fc_stack_bwd: null pointer");
    XB_REQUIRE(B >= 2 && U >= 1 && P >= 1 && C >= 1, "fc_stack_bwd: bad sizes");
    XB_REQUIRE(((uintptr_t)h1 & 15) == 0, "fc_stack_bwd: h1 must be 16-byte aligned");
    const int nchunk = (int)((B + BM - 1) / BM);
    // gz3 (B,U) lives in the head of the gpool buffer (U*B*P floats >= B*U); it is consumed by bwd_bn3_kernel
    // before bwd_fc1_kernel overwrites gpool
    float *gz3 = gpool;
    // The final-Linear weight gradient depends on nothing the rest of this backward produces: in the unphased call it runs
    // on a side stream (fork here, join at the end; the same pattern is legal inside a CUDA-graph capture).
    SideStream *side = (phases == 7 && Wf && (dWf || dbf)) ? side_stream_for_current_device() : nullptr;
    if (phases & 1) {
        // Wf == NULL: glogits is the gradient w.r.t. the unit outputs (B,U) and the final layer is skipped
        if (Wf && (dWf || dbf)) {
            cudaStream_t s2 = stream;
            if (side) {
                cudaEventRecord(side->fork, stream);
                cudaStreamWaitEvent(side->stream, side->fork, 0);
                s2 = side->stream;
            }
            if (dWf) cudaMemsetAsync(dWf, 0, (size_t)C * U * sizeof(float), s2);
            if (dbf) cudaMemsetAsync(dbf, 0, (size_t)C * sizeof(float), s2);
            const int nsb = (int)std::min<long long>(8, (B + 63) / 64);
            bwd_final_linear2_kernel<<<dim3((U + 127) / 128, (C + 7) / 8, nsb), 128, 0, s2>>>(glogits, unit, B, U, C, dWf, dbf);
            if (side) cudaEventRecord(side->join, side->stream);
        }
        const size_t smem_gu = ((size_t)8 * C + (size_t)C * GU_UT) * 4;
        if (smem_gu > 200 * 1024) {
            xb_set_error("fc_stack_bwd: %d classes need %zu bytes of shared memory", C, smem_gu);
            return XB_ERR_LIMIT;
        }
        cudaFuncSetAttribute(bwd_gunit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_gu);
        bwd_gunit_kernel<<<dim3((unsigned)((B + 7) / 8), (U + GU_UT - 1) / GU_UT), 256, smem_gu, stream>>>(glogits, Wf, unit, B, U, C, gz3);
        if (phases != 7) bwd_bn3_kernel<<<(U + 3) / 4, 128, 0, stream>>>(gz3, h2, B, U, bn3_w, mean3, invstd3, gh2, dbn3_w, dbn3_b, pb2, 1);
    }
    if (phases & 2) {
        bwd_bn3_kernel<<<(U + 3) / 4, 128, 0, stream>>>(gz3, h2, B, U, bn3_w, mean3, invstd3, gh2, dbn3_w, dbn3_b, pb2, phases == 7 ? 0 : 2);
        bwd_fc2_kernel<<<dim3(U, nchunk), FC_THREADS, 0, stream>>>(h1, gh2, B, U, bn2_w, bn2_b, mean2, invstd2, W2, pdrop, seed, seed_dev, dropmask, pb2);
    }
    if (!(phases & 4)) return xb_check_launch("fc_stack_bwd");
#define XB_LAUNCH_BWD_FC1(PHV)                                                                                                      \
    do {                                                                                                                            \
        const size_t sm = bwd_fc1_smem<PHV>();                                                                                      \
        cudaFuncSetAttribute(bwd_fc1_kernel<PHV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);                            \
        bwd_fc1_kernel<PHV><<<dim3(U, nchunk), FC_THREADS, sm, stream>>>(h1, gh2, pooled, B, U, P, bn2_w, bn2_b, mean2, invstd2, W1, \
                                                                          W2, pdrop, seed, seed_dev, pb2, gpool, dW1_partials, dW2, db1, db2,  \
                                                                          dbn2_w, dbn2_b);                                          \
    } while (0)
    int rc_tc = XB_ERR_UNSUPPORTED;
    if (xb_fc_tc_enabled())
        rc_tc = xb_bwd_fc1_tc(h1, gh2, pooled, B, U, P, bn2_w, bn2_b, mean2, invstd2, W1, W2, pdrop, seed, seed_dev, dropmask, pb2, gpool, dW1_partials,
                              dW2, db1, db2, dbn2_w, dbn2_b, stream);
    if (rc_tc == XB_OK) {
    } else if (rc_tc != XB_ERR_UNSUPPORTED) return rc_tc;
    else if (P <= 32) XB_LAUNCH_BWD_FC1(16);
    else if (P <= 40) XB_LAUNCH_BWD_FC1(20);
    else if (P <= 72) XB_LAUNCH_BWD_FC1(36);
    else if (P <= 144) XB_LAUNCH_BWD_FC1(72);
    else {
        xb_set_error("fc_stack_bwd: pooled length %d > 144 not supported", P);
        return XB_ERR_UNSUPPORTED;
    }
#undef XB_LAUNCH_BWD_FC1
    if (side) cudaStreamWaitEvent(stream, side->join, 0);
    int rc = xb_check_launch("fc_stack_bwd");
    if (rc != XB_OK) return rc;
    // dW1 = sum over batch chunks of the per-tile partials
    return explainn_reduce_partials(dW1, dW1_partials, (long long)U * H * P, nchunk, stream);
}

extern "C" int explainn_fc_stack_bwd(const float *glogits, long long B, int U, int P, int C, const float *pooled, const float *h1,
                                     const float *h2, const float *unit, const float *W1, const float *bn2_w, const float *bn2_b,
                                     const float *mean2, const float *invstd2, const float *W2, const float *bn3_w,
                                     const float *mean3, const float *invstd3, const float *Wf, float pdrop,
                                     unsigned long long seed, const unsigned long long *seed_dev, const uint32_t *dropmask, float *gh2, float *pb2, float *dW1_partials, float *gpool,
                                     float *dW1, float *db1, float *dbn2_w, float *dbn2_b, float *dW2, float *db2, float *dbn3_w,
                                     float *dbn3_b, float *dWf, float *dbf, cudaStream_t stream)
{
    return fc_stack_bwd_impl(7, glogits, B, U, P, C, pooled, h1, h2, unit, W1, bn2_w, bn2_b, mean2, invstd2, W2, bn3_w, mean3, invstd3, Wf,
                             pdrop, seed, seed_dev, dropmask, gh2, pb2, dW1_partials, gpool, dW1, db1, dbn2_w, dbn2_b, dW2, db2, dbn3_w,
                             dbn3_b, dWf, dbf, stream);
}

// The same backward in separately callable phases (bitmask 1 / 2 / 4, see fc_stack_bwd_impl), for synchronised BatchNorm
// under data parallel: between phases 1 and 2 the caller averages the first 2U floats of pb2 (BatchNorm-3 sums) over the
// ranks, between 2 and 4 all of pb2 (BatchNorm-2 sums).
extern "C" int explainn_fc_stack_bwd_phase(int phases, const float *glogits, long long B, int U, int P, int C, const float *pooled,
                                           const float *h1, const float *h2, const float *unit, const float *W1, const float *bn2_w,
                                           const float *bn2_b, const float *mean2, const float *invstd2, const float *W2,
                                           const float *bn3_w, const float *mean3, const float *invstd3, const float *Wf, float pdrop,
                                           unsigned long long seed, const unsigned long long *seed_dev, const uint32_t *dropmask,
                                           float *gh2, float *pb2, float *dW1_partials, float *gpool, float *dW1, float *db1,
                                           float *dbn2_w, float *dbn2_b, float *dW2, float *db2, float *dbn3_w, float *dbn3_b,
                                           float *dWf, float *dbf, cudaStream_t stream)
{
    return fc_stack_bwd_impl(phases, glogits, B, U, P, C, pooled, h1, h2, unit, W1, bn2_w, bn2_b, mean2, invstd2, W2, bn3_w, mean3, invstd3,
                             Wf, pdrop, seed, seed_dev, dropmask, gh2, pb2, dW1_partials, gpool, dW1, db1, dbn2_w, dbn2_b, dW2, db2,
                             dbn3_w, dbn3_b, dWf, dbf, stream);
}


// ---- num_fc = 1 entry points (explainn/models/networks.py:82-95) ---------------------------------------
extern "C" int explainn_fc1_stack_fwd_eval(const float *pooled, long long B, int U, int P, const float *W, const float *b1,
                                           const float *bn_w, const float *bn_b, const float *bn_rm, const float *bn_rv, float eps,
                                           float *unit_out, cudaStream_t stream)
{
    XB_REQUIRE(pooled && W && b1 && bn_w && bn_b && bn_rm && bn_rv && unit_out, "fc1_stack_fwd_eval: null pointer");
    XB_REQUIRE(B >= 0 && U >= 1 && P >= 1 && P <= 8192, "fc1_stack_fwd_eval: bad sizes");
    if (B == 0) return XB_OK;
    fc1x1_fwd_kernel<true><<<dim3(U, (unsigned)((B + BM - 1) / BM)), BM, (size_t)P * 4, stream>>>(pooled, B, U, P, W, b1, bn_w, bn_b, bn_rm, bn_rv,
                                                                                               eps, nullptr, nullptr, unit_out);
    return xb_check_launch("fc1_stack_fwd_eval");
}

extern "C" int explainn_fc1_stack_fwd_train(const float *pooled, long long B, int U, int P, int C, const float *W, const float *b1,
                                            const float *bn_w, const float *bn_b, float *bn_rm, float *bn_rv, float momentum,
                                            float eps, const float *Wf, const float *bf, float *h2, float *part3, float *mean3,
                                            float *invstd3, float *unit, float *logits, cudaStream_t stream)
{
    XB_REQUIRE(pooled && W && b1 && bn_w && bn_b && Wf && bf && h2 && part3 && mean3 && invstd3 && unit && logits,
               "fc1_stack_fwd_train: null pointer");
    XB_REQUIRE(B >= 2 && U >= 1 && P >= 1 && P <= 8192 && C >= 1, "fc1_stack_fwd_train: bad sizes (train-mode BatchNorm needs B >= 2)");
    const int nchunk = (int)((B + BM - 1) / BM);
    fc1x1_fwd_kernel<false><<<dim3(U, nchunk), BM, (size_t)P * 4, stream>>>(pooled, B, U, P, W, b1, nullptr, nullptr, nullptr, nullptr, eps, h2,
                                                                        part3, nullptr);
    size_t smem3 = ((size_t)OT_ROWS * U + 2 * (size_t)U) * 4;
    const int ot_ny = C >= 48 ? 3 : 1;
    const size_t smem3_wf = smem3 + (size_t)((C + ot_ny - 1) / ot_ny) * (U + 1) * 4;
    const int wf_in_smem = smem3_wf <= 200 * 1024;
    if (wf_in_smem) smem3 = smem3_wf;
    if (smem3 > 220 * 1024) {
        xb_set_error("fc1_stack_fwd_train: %d units need %zu bytes of shared memory", U, smem3);
        return XB_ERR_UNSUPPORTED;
    }
    cudaFuncSetAttribute(out_train_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem3);
    out_train_kernel<<<dim3((unsigned)((B + OT_ROWS - 1) / OT_ROWS), ot_ny), 256, smem3, stream>>>(h2, B, U, C, part3, nchunk, bn_w, bn_b, bn_rm, bn_rv,
                                                                                                 momentum, eps, mean3, invstd3, Wf, bf, unit, logits,
                                                                                                 wf_in_smem, 1.0 / (double)B);
    return xb_check_launch("fc1_stack_fwd_train");
}

extern "C" int explainn_fc1_stack_bwd(const float *glogits, long long B, int U, int P, int C, const float *pooled, const float *h2,
                                      const float *unit, const float *W, const float *bn_w, const float *mean3, const float *invstd3,
                                      const float *Wf, float *gh2, float *dW_partials, float *gpool, float *dW, float *db1,
                                      float *dbn_w, float *dbn_b, float *dWf, float *dbf, cudaStream_t stream)
{
    XB_REQUIRE(glogits && pooled && h2 && unit && W && bn_w && mean3 && invstd3 && gh2 && dW_partials && gpool && dW,
               "fc1_stack_bwd: null pointer");
    XB_REQUIRE(B >= 2 && U >= 1 && P >= 1 && C >= 1, "fc1_stack_bwd: bad sizes");
    const int nchunk = (int)((B + BM - 1) / BM);
    if (Wf) {
        if (dWf) cudaMemsetAsync(dWf, 0, (size_t)C * U * sizeof(float), stream);
        if (dbf) cudaMemsetAsync(dbf, 0, (size_t)C * sizeof(float), stream);
        const int nsb = (int)std::min<long long>(8, (B + 63) / 64);
        bwd_final_linear2_kernel<<<dim3((U + 127) / 128, (C + 7) / 8, nsb), 128, 0, stream>>>(glogits, unit, B, U, C, dWf, dbf);
    }
    float *gz3 = gpool;      // (B,U) in the head of the (B,U,P) buffer, consumed before fc1x1_bwd_kernel overwrites it
    const size_t smem_gu = ((size_t)8 * C + (size_t)C * GU_UT) * 4;
    if (smem_gu > 200 * 1024) {
        xb_set_error("fc1_stack_bwd: %d classes need %zu bytes of shared memory", C, smem_gu);
        return XB_ERR_LIMIT;
    }
    cudaFuncSetAttribute(bwd_gunit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_gu);
    bwd_gunit_kernel<<<dim3((unsigned)((B + 7) / 8), (U + GU_UT - 1) / GU_UT), 256, smem_gu, stream>>>(glogits, Wf, unit, B, U, C, gz3);
    bwd_bn3_kernel<<<(U + 3) / 4, 128, 0, stream>>>(gz3, h2, B, U, bn_w, mean3, invstd3, gh2, dbn_w, dbn_b, nullptr, 0);
    const size_t smem = ((size_t)P + BM + (size_t)BM * (P + 1)) * 4;
    if (smem > 200 * 1024) {
        xb_set_error("fc1_stack_bwd: pooled length %d needs %zu bytes of shared memory", P, smem);
        return XB_ERR_UNSUPPORTED;
    }
    cudaFuncSetAttribute(fc1x1_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    fc1x1_bwd_kernel<<<dim3(U, nchunk), BM, smem, stream>>>(pooled, gh2, B, U, P, W, gpool, dW_partials);
    if (db1) cudaMemsetAsync(db1, 0, (size_t)U * sizeof(float), stream);      // bias in front of a train-mode BatchNorm: exactly zero
    int rc = xb_check_launch("fc1_stack_bwd");
    if (rc != XB_OK) return rc;
    return explainn_reduce_partials(dW, dW_partials, (long long)U * P, nchunk, stream);
}
