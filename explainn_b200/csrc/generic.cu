// generic.cu -- the model variants beside the default unit, built from small general kernels:
//   * MaxPool1d with any (pool_size, pool_stride), overlapping or skipping windows included, and the global max-pool of
//     the num_fc = 0 unit in training mode ............................ explainn/models/networks.py:48-49,73-80,102
//   * the extra hidden layers of num_fc > 2: Dropout -> grouped Conv1d(100U -> 100U, k=1) -> BatchNorm1d -> ReLU
//     ................................................................. explainn/models/networks.py:117-145
// These paths materialise what the default path keeps on chip -- the (B,U,S) pre-activation map, (U,B,100) hidden
// tensors -- and run on the CUDA cores: they serve the reference's non-default constructor arguments, not the
// benchmarked configurations.
#include "common.cuh"
#include "../../include/explainn_b200.h"

namespace {

// ------------------------------------------------------------------------------------------------
// General max-pool over the pre-activation map z (B,U,S) (explainn_unit_activations, EXPLAINN_DTYPE_F32_PREACT).
// One thread per (b, u, p): window [p*stride, p*stride + size).  Eval: z already carries the BatchNorm affine map,
// pooled = exp(max z).  Train: the map holds x_hat; z = gamma * x_hat + beta, first maximum (MaxPool1d backward routes
// the gradient there), outputs pooled = exp(z*), x_hat* and the offset of the maximum inside the window.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pool_general_kernel(const float *__restrict__ z, long long BU, int U, int S, int P, int psize,
                                                           int pstride, const float *__restrict__ gamma, const float *__restrict__ beta,
                                                           float *__restrict__ pooled, float *__restrict__ xhat,
                                                           unsigned short *__restrict__ idx)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= BU * P) return;
    const long long r = i / P;
    const int p = (int)(i - r * P);
    const float *row = z + r * S + (long long)p * pstride;
    float g = 1.0f, be = 0.0f;
    if (gamma) {
        const int u = (int)(r % U);
        g = gamma[u];
        be = beta[u];
    }
    float xb = row[0];
    float zb = gamma ? __fmaf_rn(g, xb, be) : xb;
    int ib = 0;
    for (int j = 1; j < psize; ++j) {
        const float x = row[j];
        const float zz = gamma ? __fmaf_rn(g, x, be) : x;
        if (zz > zb) { zb = zz; xb = x; ib = j; }
    }
    pooled[i] = det_expf(zb);
    if (xhat) xhat[i] = xb;
    if (idx) idx[i] = (unsigned short)ib;
}

// ------------------------------------------------------------------------------------------------
// Strided batched GEMM on the CUDA cores: C[g](m,n) = sum_k A[g](m,k) * B[g](k,n) (+ bias[g](n)), every operand with
// its own element strides, so the same kernel multiplies (B,U,P)-, (U,B,H)- and (U,H,P)-shaped views without copies.
// 64 x 64 tile per CTA, 16 x 16 threads, 4 x 4 outputs per thread, K in steps of 16.  fp32 FMA accumulation.
// ------------------------------------------------------------------------------------------------
struct GemmArgs {
    const float *A, *B, *bias;
    float *C;
    int M, N, K;
    long long sAg, sAm, sAk, sBg, sBk, sBn, sCg, sCm, sCn, sbg;
};

__global__ void __launch_bounds__(256) grouped_gemm_kernel(const GemmArgs a)
{
    __shared__ float sA[16][64 + 1], sB[16][64 + 1];
    const int g = blockIdx.z, m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const float *A = a.A + (long long)g * a.sAg, *B = a.B + (long long)g * a.sBg;
    float acc[4][4] = {};
    for (int k0 = 0; k0 < a.K; k0 += 16) {
        for (int i = threadIdx.x; i < 16 * 64; i += 256) {
            const int kk = i & 15, mm = i >> 4;             // consecutive threads walk k: good when k is the contiguous axis
            const int m = m0 + mm, k = k0 + kk;
            sA[kk][mm] = (m < a.M && k < a.K) ? A[m * a.sAm + k * a.sAk] : 0.0f;
            const int n = n0 + mm;
            sB[kk][mm] = (n < a.N && k < a.K) ? B[k * a.sBk + n * a.sBn] : 0.0f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < 16; ++kk) {
            float av[4], bv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { av[i] = sA[kk][ty * 4 + i]; bv[i] = sB[kk][tx * 4 + i]; }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
    float *C = a.C + (long long)g * a.sCg;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        if (m >= a.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tx * 4 + j;
            if (n >= a.N) continue;
            float v = acc[i][j];
            if (a.bias) v += a.bias[(long long)g * a.sbg + n];
            C[m * a.sCm + n * a.sCn] = v;
        }
    }
}

int launch_gemm(const GemmArgs &a, int groups, cudaStream_t stream)
{
    if (a.M <= 0 || a.N <= 0 || groups <= 0) return XB_OK;
    dim3 grid((a.N + 63) / 64, (a.M + 63) / 64, groups);
    grouped_gemm_kernel<<<grid, 256, 0, stream>>>(a);
    return xb_check_launch("grouped_gemm");
}

}  // namespace

extern "C" int explainn_pool_general(const float *z, long long B, int U, int S, int pool_size, int pool_stride, const float *gamma,
                                     const float *beta, float *pooled, float *xhat, unsigned short *idx, cudaStream_t stream)
{
    XB_REQUIRE(z && pooled && B >= 0 && U >= 1 && pool_size >= 1 && pool_stride >= 1 && S >= pool_size, "pool_general: bad argument");
    XB_REQUIRE((gamma == nullptr) == (beta == nullptr), "pool_general: gamma and beta go together");
    if (pool_size > 65535) {
        xb_set_error("pool_general: pool_size %d exceeds the 16-bit window offset", pool_size);
        return XB_ERR_LIMIT;
    }
    if (B == 0) return XB_OK;
    const int P = (S - pool_size) / pool_stride + 1;
    const long long n = B * U * P;
    pool_general_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(z, B * U, U, S, P, pool_size, pool_stride, gamma, beta, pooled, xhat, idx);
    return xb_check_launch("pool_general");
}

// ================================================================================================
// Hidden block of the num_fc > 2 unit: in (B,U,Pin) -> grouped FC (Pin -> 100) -> BatchNorm1d(100U) -> ReLU -> Dropout
// -> out (B,U,100).  (networks.py:117-138: the first hidden layer lives in `linears`, the dropout in front of the next
// layer belongs to `linears_bg[i]` / `last_linear`; grouping it behind the ReLU gives blocks of one shape.)
// Y (U,B,100) is the pre-BatchNorm tensor kept for the backward.
// ================================================================================================
namespace {

constexpr int GH = XB_HIDDEN;

// batch statistics of Y per channel (u, j): grid U, 512 threads = 4 row groups x 128 channel slots
__global__ void __launch_bounds__(512) hidden_stats_kernel(const float *__restrict__ Y, long long B, int U, float momentum, float eps,
                                                           float *__restrict__ run_mean, float *__restrict__ run_var,
                                                           float *__restrict__ mean, float *__restrict__ invstd)
{
    __shared__ double s1[4][128], s2[4][128];
    const int u = blockIdx.x, j = threadIdx.x & 127, rg = threadIdx.x >> 7;
    double a = 0.0, q = 0.0;
    if (j < GH) {
        const float *y = Y + (long long)u * B * GH + j;
        for (long long b = rg; b < B; b += 4) {
            const double v = (double)y[b * GH];
            a += v;
            q += v * v;
        }
    }
    s1[rg][j] = a; s2[rg][j] = q;
    __syncthreads();
    if (rg == 0 && j < GH) {
        a = s1[0][j] + s1[1][j] + s1[2][j] + s1[3][j];
        q = s2[0][j] + s2[1][j] + s2[2][j] + s2[3][j];
        const double m = a / (double)B;
        double var = q / (double)B - m * m;
        if (var < 0.0) var = 0.0;
        const int ch = u * GH + j;
        mean[ch] = (float)m;
        invstd[ch] = (float)(1.0 / sqrt(var + (double)eps));
        if (run_mean) {
            const double unb = B > 1 ? var * (double)B / (double)(B - 1) : var;
            run_mean[ch] = (1.0f - momentum) * run_mean[ch] + momentum * (float)m;
            run_var[ch] = (1.0f - momentum) * run_var[ch] + momentum * (float)unb;
        }
    }
}

// out[b][u][j] = dropout(relu(gamma * (y - mean) * invstd + beta)); one thread per 4 consecutive channels (one keep nibble)
__global__ void __launch_bounds__(256) hidden_apply_kernel(const float *__restrict__ Y, long long B, int U, const float *__restrict__ gamma,
                                                           const float *__restrict__ beta, const float *__restrict__ mean,
                                                           const float *__restrict__ invstd, float pdrop, unsigned long long seed0,
                                                           const unsigned long long *__restrict__ seed_dev, float *__restrict__ out)
{
    const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;      // quad index over (u, b, j/4)
    const long long nq = (long long)U * B * (GH / 4);
    if (q >= nq) return;
    const unsigned long long seed = seed0 + (seed_dev ? *seed_dev * 0xD1B54A32D192ED03ull : 0ull);
    const int jq = (int)(q % (GH / 4));
    const long long ub = q / (GH / 4);
    const long long b = ub % B;
    const int u = (int)(ub / B);
    const unsigned thresh = pdrop > 0.f ? (unsigned)min(4294967295.0, (double)pdrop * 4294967296.0) : 0u;
    const float keep_scale = pdrop > 0.f ? 1.0f / (1.0f - pdrop) : 1.0f;
    const unsigned nib = pdrop > 0.f ? drop_keep4(seed, (unsigned long long)q, thresh) : 15u;
    const float4 y = *reinterpret_cast<const float4 *>(Y + q * 4);
    const float yv[4] = {y.x, y.y, y.z, y.w};
    float o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int ch = u * GH + jq * 4 + i;
        const float zz = fmaf(gamma[ch] * invstd[ch], yv[i] - mean[ch], beta[ch]);
        o[i] = (zz > 0.f && ((nib >> i) & 1u)) ? zz * keep_scale : 0.0f;
    }
    *reinterpret_cast<float4 *>(out + ((b * U + u) * GH + jq * 4)) = make_float4(o[0], o[1], o[2], o[3]);
}

// BatchNorm backward, pass 1: per channel s1 = sum_b ghat, s2 = sum_b ghat * yhat with
// ghat = gout * keep * scale * [z > 0].  grid U, 512 threads.  Also writes dgamma = s2, dbeta = s1.
__global__ void __launch_bounds__(512) hidden_bwd_sums_kernel(const float *__restrict__ gout, const float *__restrict__ Y, long long B, int U,
                                                              const float *__restrict__ gamma, const float *__restrict__ beta,
                                                              const float *__restrict__ mean, const float *__restrict__ invstd,
                                                              float pdrop, unsigned long long seed0,
                                                              const unsigned long long *__restrict__ seed_dev, float *__restrict__ sums,
                                                              float *__restrict__ dgamma, float *__restrict__ dbeta)
{
    __shared__ double s1[4][128], s2[4][128];
    const unsigned long long seed = seed0 + (seed_dev ? *seed_dev * 0xD1B54A32D192ED03ull : 0ull);
    const int u = blockIdx.x, j = threadIdx.x & 127, rg = threadIdx.x >> 7;
    const unsigned thresh = pdrop > 0.f ? (unsigned)min(4294967295.0, (double)pdrop * 4294967296.0) : 0u;
    const float keep_scale = pdrop > 0.f ? 1.0f / (1.0f - pdrop) : 1.0f;
    double a = 0.0, q = 0.0;
    if (j < GH) {
        const int ch = u * GH + j;
        const float g = gamma[ch], be = beta[ch], mu = mean[ch], is = invstd[ch];
        for (long long b = rg; b < B; b += 4) {
            const long long e = ((long long)u * B + b) * GH + j;
            const float yh = (Y[e] - mu) * is;
            const float zz = fmaf(g, yh, be);
            float gh = gout[(b * U + u) * GH + j];
            const bool keep = pdrop > 0.f ? ((drop_keep4(seed, (unsigned long long)(e >> 2), thresh) >> (e & 3)) & 1u) : true;
            gh = (zz > 0.f && keep) ? gh * keep_scale : 0.0f;
            a += (double)gh;
            q += (double)gh * (double)yh;
        }
    }
    s1[rg][j] = a; s2[rg][j] = q;
    __syncthreads();
    if (rg == 0 && j < GH) {
        a = s1[0][j] + s1[1][j] + s1[2][j] + s1[3][j];
        q = s2[0][j] + s2[1][j] + s2[2][j] + s2[3][j];
        const int ch = u * GH + j;
        sums[2 * ch] = (float)(a / (double)B);
        sums[2 * ch + 1] = (float)(q / (double)B);
        if (dgamma) dgamma[ch] = (float)q;
        if (dbeta) dbeta[ch] = (float)a;
    }
}

// pass 2: G[u][b][j] = gamma * invstd * (ghat - mean(ghat) - yhat * mean(ghat * yhat))
__global__ void __launch_bounds__(256) hidden_bwd_apply_kernel(const float *__restrict__ gout, const float *__restrict__ Y, long long B, int U,
                                                               const float *__restrict__ gamma, const float *__restrict__ beta,
                                                               const float *__restrict__ mean, const float *__restrict__ invstd,
                                                               const float *__restrict__ sums, float pdrop, unsigned long long seed0,
                                                               const unsigned long long *__restrict__ seed_dev, float *__restrict__ G)
{
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long)U * B * GH) return;
    const unsigned long long seed = seed0 + (seed_dev ? *seed_dev * 0xD1B54A32D192ED03ull : 0ull);
    const int j = (int)(e % GH);
    const long long ub = e / GH;
    const long long b = ub % B;
    const int u = (int)(ub / B);
    const int ch = u * GH + j;
    const unsigned thresh = pdrop > 0.f ? (unsigned)min(4294967295.0, (double)pdrop * 4294967296.0) : 0u;
    const float keep_scale = pdrop > 0.f ? 1.0f / (1.0f - pdrop) : 1.0f;
    const float is = invstd[ch], g = gamma[ch];
    const float yh = (Y[e] - mean[ch]) * is;
    const float zz = fmaf(g, yh, beta[ch]);
    float gh = gout[(b * U + u) * GH + j];
    const bool keep = pdrop > 0.f ? ((drop_keep4(seed, (unsigned long long)(e >> 2), thresh) >> (e & 3)) & 1u) : true;
    gh = (zz > 0.f && keep) ? gh * keep_scale : 0.0f;
    G[e] = g * is * (gh - sums[2 * ch] - yh * sums[2 * ch + 1]);
}

int hidden_gemm_fwd(const float *X, long long B, int U, int Pin, const float *W, const float *b, float *Y, cudaStream_t stream)
{
    GemmArgs a{};
    a.A = X; a.B = W; a.bias = b; a.C = Y; a.M = (int)B; a.N = GH; a.K = Pin;
    a.sAg = Pin; a.sAm = (long long)U * Pin; a.sAk = 1;
    a.sBg = (long long)GH * Pin; a.sBk = 1; a.sBn = Pin;
    a.sCg = B * GH; a.sCm = GH; a.sCn = 1; a.sbg = GH;
    return launch_gemm(a, U, stream);
}

}  // namespace

extern "C" int explainn_hidden_fwd_train(const float *X, long long B, int U, int Pin, const float *W, const float *b,
                                         const float *bn_w, const float *bn_b, float *bn_rm, float *bn_rv, float momentum,
                                         float eps, float pdrop, unsigned long long seed, const unsigned long long *seed_dev,
                                         float *Y, float *mean, float *invstd, float *out, cudaStream_t stream)
{
    XB_REQUIRE(X && W && b && bn_w && bn_b && Y && mean && invstd && out, "hidden_fwd_train: null pointer");
    XB_REQUIRE(B >= 2 && U >= 1 && Pin >= 1 && pdrop >= 0.f && pdrop < 1.f, "hidden_fwd_train: bad sizes (train-mode BatchNorm needs B >= 2)");
    int rc = hidden_gemm_fwd(X, B, U, Pin, W, b, Y, stream);
    if (rc != XB_OK) return rc;
    hidden_stats_kernel<<<U, 512, 0, stream>>>(Y, B, U, momentum, eps, bn_rm, bn_rv, mean, invstd);
    const long long nq = (long long)U * B * (GH / 4);
    hidden_apply_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, stream>>>(Y, B, U, bn_w, bn_b, mean, invstd, pdrop, seed, seed_dev, out);
    return xb_check_launch("hidden_fwd_train");
}

// eval: running statistics, no dropout.  Y (U,B,100) and stat (2*100U: mean, invstd) are scratch.
extern "C" int explainn_hidden_fwd_eval(const float *X, long long B, int U, int Pin, const float *W, const float *b,
                                        const float *bn_w, const float *bn_b, const float *bn_rm, const float *bn_rv, float eps,
                                        float *Y, float *stat, float *out, cudaStream_t stream);

namespace {
__global__ void hidden_eval_stat_kernel(const float *__restrict__ rm, const float *__restrict__ rv, float eps, int n, float *__restrict__ stat)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    stat[i] = rm[i];
    stat[n + i] = __fdiv_rn(1.0f, __fsqrt_rn(__fadd_rn(rv[i], eps)));
}
}  // namespace

extern "C" int explainn_hidden_fwd_eval(const float *X, long long B, int U, int Pin, const float *W, const float *b,
                                        const float *bn_w, const float *bn_b, const float *bn_rm, const float *bn_rv, float eps,
                                        float *Y, float *stat, float *out, cudaStream_t stream)
{
    XB_REQUIRE(X && W && b && bn_w && bn_b && bn_rm && bn_rv && Y && stat && out, "hidden_fwd_eval: null pointer");
    XB_REQUIRE(B >= 0 && U >= 1 && Pin >= 1, "hidden_fwd_eval: bad sizes");
    if (B == 0) return XB_OK;
    int rc = hidden_gemm_fwd(X, B, U, Pin, W, b, Y, stream);
    if (rc != XB_OK) return rc;
    const int n = U * GH;
    hidden_eval_stat_kernel<<<(n + 255) / 256, 256, 0, stream>>>(bn_rm, bn_rv, eps, n, stat);
    const long long nq = (long long)U * B * (GH / 4);
    hidden_apply_kernel<<<(unsigned)((nq + 255) / 256), 256, 0, stream>>>(Y, B, U, bn_w, bn_b, stat, stat + n, 0.0f, 0ull, nullptr, out);
    return xb_check_launch("hidden_fwd_eval");
}

// Backward of the block.  gout (B,U,100) = d loss / d out.  Scratch: sums (2*100U), G (U,B,100).  Outputs: gin (B,U,Pin)
// (may be NULL), dW (100U,Pin), db (100U, exact zeros: the bias feeds a train-mode BatchNorm), dbn_w / dbn_b (100U).
extern "C" int explainn_hidden_bwd(const float *gout, long long B, int U, int Pin, const float *X, const float *Y, const float *W,
                                   const float *bn_w, const float *bn_b, const float *mean, const float *invstd, float pdrop,
                                   unsigned long long seed, const unsigned long long *seed_dev, float *sums, float *G, float *gin,
                                   float *dW, float *db, float *dbn_w, float *dbn_b, cudaStream_t stream)
{
    XB_REQUIRE(gout && X && Y && W && bn_w && bn_b && mean && invstd && sums && G && dW, "hidden_bwd: null pointer");
    XB_REQUIRE(B >= 2 && U >= 1 && Pin >= 1, "hidden_bwd: bad sizes");
    hidden_bwd_sums_kernel<<<U, 512, 0, stream>>>(gout, Y, B, U, bn_w, bn_b, mean, invstd, pdrop, seed, seed_dev, sums, dbn_w, dbn_b);
    const long long ne = (long long)U * B * GH;
    hidden_bwd_apply_kernel<<<(unsigned)((ne + 255) / 256), 256, 0, stream>>>(gout, Y, B, U, bn_w, bn_b, mean, invstd, sums, pdrop, seed, seed_dev, G);
    int rc = XB_OK;
    if (gin) {      // gin[b][u][p] = sum_j G[u][b][j] * W[u][j][p]
        GemmArgs a{};
        a.A = G; a.B = W; a.C = gin; a.M = (int)B; a.N = Pin; a.K = GH;
        a.sAg = B * GH; a.sAm = GH; a.sAk = 1;
        a.sBg = (long long)GH * Pin; a.sBk = Pin; a.sBn = 1;
        a.sCg = Pin; a.sCm = (long long)U * Pin; a.sCn = 1;
        rc = launch_gemm(a, U, stream);
        if (rc != XB_OK) return rc;
    }
    {               // dW[u][j][p] = sum_b G[u][b][j] * X[b][u][p]
        GemmArgs a{};
        a.A = G; a.B = X; a.C = dW; a.M = GH; a.N = Pin; a.K = (int)B;
        a.sAg = B * GH; a.sAm = 1; a.sAk = GH;
        a.sBg = Pin; a.sBk = (long long)U * Pin; a.sBn = 1;
        a.sCg = (long long)GH * Pin; a.sCm = Pin; a.sCn = 1;
        rc = launch_gemm(a, U, stream);
        if (rc != XB_OK) return rc;
    }
    if (db) cudaMemsetAsync(db, 0, (size_t)U * GH * sizeof(float), stream);
    return xb_check_launch("hidden_bwd");
}
