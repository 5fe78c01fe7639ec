// fc.cu -- the per-unit fully-connected stack and the final linear layer, forward.
//
// Replaces (paths relative to /root/reference, explainn/models/networks.py):
//   linears[4..13]: Flatten, Unsqueeze, grouped 1x1 Conv1d(U*P -> 100U), BatchNorm1d(100U), ReLU,
//                   Dropout(0.3), grouped 1x1 Conv1d(100U -> U), BatchNorm1d(U), ReLU, Flatten   (:103-116)
//   final = Linear(U, C)                                                                        (:147,161)
//
// Layout.  The pooled activations arrive unit-major, pooled[u][b][p], so that the tile a CTA works on
// (one unit x 128 sequences) is one contiguous block.  The hidden activations h1[u][b][100] use the
// same convention.  A CTA keeps W1[u] (100 x P) transposed in shared memory; each thread owns two
// sequences x 25 hidden channels, so every W1 value is fetched by a warp-uniform (broadcast) LDS.128
// and the pooled values by conflict-free scalar loads.
#include "common.cuh"
#include "../../include/explainn_b200.h"

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

__device__ __forceinline__ void load_a(const FcSmem &s, const float *__restrict__ tile, int rows, int P)
{
    // tile: rows x P contiguous
    for (int i = threadIdx.x; i < BM * P; i += blockDim.x) {
        const int r = i / P, p = i - r * P;
        s.As[r * s.Pp + p] = r < rows ? tile[i] : 0.0f;
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
    load_a(s, pooled + ((size_t)u * B + b0) * P, rows, P);
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
    for (int c = threadIdx.x >> 5; c < C; c += 8) {
        const float *w = W + (size_t)c * U;
        float a0 = 0.f, a1 = 0.f;
        int k = 0;
        for (; k + 1 < U; k += 2) {
            a0 = fmaf(xs[r * Up + k], __ldg(w + k), a0);
            a1 = fmaf(xs[r * Up + k + 1], __ldg(w + k + 1), a1);
        }
        if (k < U) a0 = fmaf(xs[r * Up + k], __ldg(w + k), a0);
        if (r < rows) out[(b0 + r) * C + c] = (a0 + a1) + bias[c];
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
    const size_t smem = (size_t)32 * (U + 1) * 4;
    if (smem > 200 * 1024) {
        xb_set_error("linear_fwd: %d input features need %zu bytes of shared memory", U, smem);
        return XB_ERR_UNSUPPORTED;
    }
    cudaFuncSetAttribute(linear_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    linear_fwd_kernel<<<(unsigned)((B + 31) / 32), 256, smem, stream>>>(x, B, U, C, W, bias, out);
    return xb_check_launch("linear_fwd");
}
