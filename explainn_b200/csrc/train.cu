// train.cu -- training-only kernels around the conv stage, the fused loss and the fused Adam step.
//
// Reference behaviour restated (paths relative to /root/reference):
//   BatchNorm1d(U) in train mode (batch statistics over (B,S), running-stat update with momentum 0.1 and
//   the unbiased variance) ........................................ explainn/models/networks.py:100
//   autograd backward of Conv1d -> BatchNorm1d -> exp -> MaxPool1d .... explainn/train/train.py:65
//   criterion (BCEWithLogitsLoss / MSELoss, mean reduction) ........... scripts/train.py:210-219, train.py:62
//   torch.optim.Adam defaults ......................................... scripts/train.py:246-247, train.py:67
//
// Because the conv input is one-hot, the batch statistics of the conv output and the dense part of the
// BatchNorm backward follow from input-only integer statistics (SURVEY.md 7.3 b):
//   M[i][j] = sum_{b,s} x[b,c,s+k] * x[b,c',s+k'],   i = c*F+k, j = c'*F+k'    (co-occurrence counts)
//   cnt[i]  = M[i][i]
//   mean(y_u) = bias_u + w_u.cnt / N ;  E[(y_u-bias_u)^2] = w_u^T M w_u / N ;  N = B*S
// so the (B,U,S) conv output is never needed twice, and the backward touches only the pool-argmax
// positions plus these tables.
#include "common.cuh"
#include "../../include/explainn_b200.h"
#include <algorithm>

namespace {

// ------------------------------------------------------------------------------------------------
// co-occurrence counts, bit-parallel.  Per sequence four indicator bitmasks I_c (bit s set iff base s is c; N sets
// none) are built in shared memory; the thread that owns the tap pair (k <= k') then gets, for all 16 base pairs,
//   M[(c,k),(c',k')] += popcount( (I_c >> k) & (I_c' >> k') & [s < S] )
// 32 positions per AND+POPC, counters in registers, one atomicAdd per counter and CTA at the end.
// ------------------------------------------------------------------------------------------------
// the 16 even-position bits of x packed into the low half (Morton decode)
__device__ __forceinline__ uint32_t even_bits(uint32_t x)
{
    x &= 0x55555555u;
    x = (x | (x >> 1)) & 0x33333333u;
    x = (x | (x >> 2)) & 0x0F0F0F0Fu;
    x = (x | (x >> 4)) & 0x00FF00FFu;
    x = (x | (x >> 8)) & 0x0000FFFFu;
    return x;
}

__global__ void __launch_bounds__(320) cooc_kernel(const uint32_t *__restrict__ codes, const uint32_t *__restrict__ nmask,
                                                   long long B, int L, int F, int wc, int wm, int *__restrict__ M)
{
    extern __shared__ uint32_t ind[];          // [4][wm + 1] indicator words, one zero word of padding per base
    const int NT = blockDim.x, tid = threadIdx.x;
    const int IW = wm + 1;
    const int npairs = F * (F + 1) / 2;
    int k = 0, k2 = 0;                         // pair index -> (k, k2) with k <= k2
    if (tid < npairs) {
        int r = tid;
        while (r >= F - k) { r -= F - k; ++k; }
        k2 = k + r;
    }
    int cnt[4][4];
#pragma unroll
    for (int c = 0; c < 4; ++c)
#pragma unroll
        for (int c2 = 0; c2 < 4; ++c2) cnt[c][c2] = 0;
    const int S = L - F + 1;
    for (long long b = blockIdx.x; b < B; b += gridDim.x) {
        __syncthreads();
        // indicator words: word w covers positions 32w .. 32w+31 = code words 2w, 2w+1
        for (int i = tid; i < 4 * IW; i += NT) {
            const int c = i / IW, w = i - c * IW;
            uint32_t bits = 0;
            if (w < wm) {
                const uint32_t c0 = codes[b * wc + 2 * w], c1 = 2 * w + 1 < wc ? codes[b * wc + 2 * w + 1] : 0u;
                // a position matches base c iff its 2-bit code equals c: xor with the replicated pattern, then test both bits
                const uint32_t pat = 0x55555555u * (uint32_t)c;
                const uint32_t x0 = c0 ^ pat, x1 = c1 ^ pat;
                const uint32_t z0 = ~(x0 | (x0 >> 1)) & 0x55555555u, z1 = ~(x1 | (x1 >> 1)) & 0x55555555u;   // bit 2j set iff code j == c
                bits = even_bits(z0) | (even_bits(z1) << 16);
                bits &= ~nmask[b * wm + w];
                if (32 * w + 32 > L) bits &= (L - 32 * w >= 32) ? 0xFFFFFFFFu : ((1u << (L - 32 * w)) - 1u);
            }
            ind[i] = bits;
        }
        __syncthreads();
        if (tid < npairs) {
            for (int w = 0; 32 * w < S; ++w) {
                const uint32_t valid = S - 32 * w >= 32 ? 0xFFFFFFFFu : ((1u << (S - 32 * w)) - 1u);
                uint32_t a[4], bb[4];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    a[c] = __funnelshift_r(ind[c * IW + w], ind[c * IW + w + 1], k) & valid;
                    bb[c] = __funnelshift_r(ind[c * IW + w], ind[c * IW + w + 1], k2);
                }
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int c2 = 0; c2 < 4; ++c2) cnt[c][c2] += __popc(a[c] & bb[c2]);
            }
        }
    }
    if (tid < npairs) {
        const int W = 4 * F;
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int c2 = 0; c2 < 4; ++c2) {
                const int v = cnt[c][c2];
                if (!v) continue;
                atomicAdd(&M[(c * F + k) * W + (c2 * F + k2)], v);
                if (k != k2) atomicAdd(&M[(c2 * F + k2) * W + (c * F + k)], v);
            }
    }
}

// ------------------------------------------------------------------------------------------------
// BatchNorm-1 batch statistics from M (one block per unit, fp64 accumulation)
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) bn1_stats_kernel(const int *__restrict__ M, const float *__restrict__ w,
                                                        const float *__restrict__ bias, int F, double N, float momentum,
                                                        float *__restrict__ run_mean, float *__restrict__ run_var,
                                                        long long *__restrict__ nbt, float *__restrict__ mean_out,
                                                        float *__restrict__ var_out, float *__restrict__ Mw_out)
{
    __shared__ double wsm[4 * 4 * XB_MAX_G];
    __shared__ double red_a[4], red_q[4];
    const int u = blockIdx.x, W = 4 * F;
    for (int i = threadIdx.x; i < W; i += blockDim.x) wsm[i] = (double)w[(size_t)u * W + i];
    __syncthreads();
    double a = 0.0, q = 0.0;
    for (int i = threadIdx.x; i < W; i += blockDim.x) {
        const int *row = M + (size_t)i * W;
        double d = 0.0;
        for (int j = 0; j < W; ++j) d += wsm[j] * (double)row[j];
        Mw_out[(size_t)u * W + i] = (float)d;
        q += wsm[i] * d;
        a += wsm[i] * (double)row[i];
    }
    a = warp_sum_d(a);
    q = warp_sum_d(q);
    if ((threadIdx.x & 31) == 0) { red_a[threadIdx.x >> 5] = a; red_q[threadIdx.x >> 5] = q; }
    __syncthreads();
    if (threadIdx.x == 0) {
        a = (red_a[0] + red_a[1]) + (red_a[2] + red_a[3]);
        q = (red_q[0] + red_q[1]) + (red_q[2] + red_q[3]);
        const double mc = a / N;
        double var = q / N - mc * mc;
        if (var < 0.0) var = 0.0;
        const float mean = (float)((double)bias[u] + mc);
        mean_out[u] = mean;
        var_out[u] = (float)var;
        if (run_mean) {
            const double unb = N > 1.0 ? var * N / (N - 1.0) : var;
            run_mean[u] = (1.0f - momentum) * run_mean[u] + momentum * mean;
            run_var[u] = (1.0f - momentum) * run_var[u] + momentum * (float)unb;
        }
        if (nbt && u == 0) *nbt += 1;
    }
}

// ------------------------------------------------------------------------------------------------
// Sparse part of the conv backward: only pool-argmax positions carry gradient.
// grid (U, nsplit); 4 warps; lanes walk the (b,p) entries of one unit, 32 at a time, with lane-private
// accumulators acc[k][c][lane] in shared memory (conflict-free, no atomics).  FT > 0: filter size known at
// compile time (taps fully unrolled, all loads of an entry issued before its stores); FT == 0: any F.
// ------------------------------------------------------------------------------------------------
constexpr int BW_WARPS = 4;

template <int FT>
__global__ void __launch_bounds__(BW_WARPS * 32) conv_bwd_sparse_kernel(
    const uint32_t *__restrict__ codes, const uint32_t *__restrict__ nmask, long long B, int L, int Frt, int P, int wc, int wm,
    const float *__restrict__ gpool, const float *__restrict__ pooled, const float *__restrict__ xhat,
    const unsigned char *__restrict__ idx, const unsigned short *__restrict__ idx16, int pstride,
    float *__restrict__ partial /* (nsplit, U, 4F+2) */, int U)
{
    extern __shared__ float acc[];   // [BW_WARPS][F][4][32] then [BW_WARPS][4F+2] for the block reduce
    const int F = FT > 0 ? FT : Frt;
    const int u = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31, W = 4 * F;
    float *my = acc + (size_t)warp * W * 32;
    for (int i = lane; i < W * 32; i += 32) my[i] = 0.0f;
    __syncwarp();
    const long long n = B * P;
    const long long per = (n + gridDim.y - 1) / gridDim.y;
    const long long e0 = per * blockIdx.y, e1 = min(n, e0 + per);
    float dbeta = 0.f, dgamma = 0.f;
    float *mine = my + lane;
    // Software pipeline over this lane's entries e = first, first + 128, ...:
    //   stage A (entry i+4): gradient, pooled value, xhat, argmax offset      stage B (entry i+2): the code / N-mask words
    //   at the argmax position (their address depends on stage A)             stage C (entry i): shared-memory updates
    // Entries past the end are clamped to the last one and weighted with zero, so every lane runs the same trip count.
    constexpr int STRIDE = BW_WARPS * 32;
    const int db = STRIDE / P, dp = STRIDE - db * P;
    const long long first = e0 + warp * 32 + lane, last = e1 - 1;
    const int niter = e1 > e0 ? (int)((e1 - e0 + STRIDE - 1) / STRIDE) : 0;
    struct StA { float gp, po, xh; int ix, p7, b; bool valid; };     // raw loaded values: nothing here may wait for a load
    struct StB { float gz, xh; uint32_t w0, w1, w2, m0, m1; int s; };
    long long ea = first;                      // entry index of the next stage-A fetch
    int ab, ap;                                // its (b, p), kept incrementally (host wrapper guarantees B*P < 2^31)
    {
        const long long ec = min(ea, last);
        ab = (int)(ec / P);
        ap = (int)(ec - (long long)ab * P);
    }
    const int lastb = (int)(last / P), lastp = (int)(last - (long long)lastb * P);
    auto fetchA = [&]() {
        StA r;
        r.valid = ea <= last;
        if (!r.valid) { ab = lastb; ap = lastp; }
        const size_t o = ((size_t)ab * U + u) * P + ap;   // (B,U,P) layout
        r.gp = gpool[o];
        r.po = pooled[o];
        r.xh = xhat[o];
        r.ix = idx16 ? (int)idx16[o] : (int)idx[o];      // offset of the maximum inside its pool window
        r.p7 = ap * pstride;
        r.b = ab;
        ea += STRIDE;
        ab += db; ap += dp;
        if (ap >= P) { ap -= P; ++ab; }
        return r;
    };
    auto fetchB = [&](const StA &x) {
        StB r;
        r.gz = x.valid ? x.gp * x.po : 0.f;
        r.xh = x.xh;
        r.s = x.p7 + x.ix;
        const uint32_t *cw = codes + (size_t)x.b * wc;
        const uint32_t *mw = nmask + (size_t)x.b * wm;
        const int wi = r.s >> 4, mi = r.s >> 5;
        r.w0 = cw[wi];
        r.w1 = wi + 1 < wc ? cw[wi + 1] : 0u;
        r.w2 = wi + 2 < wc ? cw[wi + 2] : 0u;
        r.m0 = mw[mi];
        r.m1 = mi + 1 < wm ? mw[mi + 1] : 0u;
        return r;
    };
    // shared-memory updates of one entry
    auto consume = [&](const StB &c) {
        const float gz = c.gz;
        dbeta += gz;
        dgamma = fmaf(gz, c.xh, dgamma);
        const int sh = 2 * (c.s & 15);
        const uint32_t A0 = __funnelshift_r(c.w0, c.w1, sh), A1 = __funnelshift_r(c.w1, c.w2, sh);
        const uint32_t MW = __funnelshift_r(c.m0, c.m1, c.s & 31) & (F >= 32 ? 0xFFFFFFFFu : ((1u << F) - 1u));
        if (FT > 0) {
            // byte offsets of the F (k, code) cells; k*512 is a compile-time part of the address
            float v[FT > 0 ? FT : 1];
            uint32_t off[FT > 0 ? FT : 1];
            char *base = reinterpret_cast<char *>(mine);
#pragma unroll
            for (int k = 0; k < FT; ++k) {
                const uint32_t A = k < 16 ? A0 : A1;
                const int shr = ((2 * k) & 31) - 7;             // code << 7  =  code * 32 floats * 4 bytes
                off[k] = (shr >= 0 ? (A >> shr) : (A << -shr)) & 0x180u;
            }
#pragma unroll
            for (int k = 0; k < FT; ++k) v[k] = *reinterpret_cast<float *>(base + k * 512 + off[k]);
            if (MW == 0u) {
#pragma unroll
                for (int k = 0; k < FT; ++k) *reinterpret_cast<float *>(base + k * 512 + off[k]) = v[k] + gz;
            } else {
#pragma unroll
                for (int k = 0; k < FT; ++k) *reinterpret_cast<float *>(base + k * 512 + off[k]) = v[k] + (((MW >> k) & 1u) ? 0.0f : gz);
            }
        } else {
            for (int k = 0; k < F; ++k) {
                const uint32_t cc = ((k < 16 ? A0 : A1) >> ((2 * k) & 31)) & 3u;
                if (!((MW >> k) & 1u)) mine[k * 128 + cc * 32] += gz;
            }
        }
    };
    // Unrolled by two with two register sets (X, Y) and NO register rotation: a copy of a stage-B register would wait
    // for its load (that copy was 46 % of the kernel's stall samples).  In the trip that consumes entries i and i+1 the
    // code words of entries i+2 / i+3 and the gradients of entries i+4 / i+5 are requested.
    StA saX{}, saY{};
    StB sbX{}, sbY{};
    if (niter > 0) {
        saX = fetchA();
        saY = fetchA();
        sbX = fetchB(saX);
        saX = fetchA();
        sbY = fetchB(saY);
        saY = fetchA();
    }
    for (int it = 0; it < niter; it += 2) {
        consume(sbX);
        sbX = fetchB(saX);
        saX = fetchA();
        if (it + 1 < niter) consume(sbY);
        sbY = fetchB(saY);
        saY = fetchA();
    }
    __syncwarp();
    // lane-transposed reduction: lane l sums cells l, l+32, ...  (cell r = k*4 + c  ->  output index c*F + k)
    float *out = acc + (size_t)BW_WARPS * W * 32 + (size_t)warp * (W + 2);
    for (int r = lane; r < W; r += 32) {
        float v = 0.f;
#pragma unroll 8
        for (int j = 0; j < 32; ++j) v += my[r * 32 + ((j + lane) & 31)];
        out[(r & 3) * F + (r >> 2)] = v;
    }
    dbeta = warp_sum(dbeta);
    dgamma = warp_sum(dgamma);
    if (lane == 0) { out[W] = dgamma; out[W + 1] = dbeta; }
    __syncthreads();
    for (int r = threadIdx.x; r < W + 2; r += blockDim.x) {
        float v = 0.f;
        for (int w = 0; w < BW_WARPS; ++w) v += acc[(size_t)BW_WARPS * W * 32 + (size_t)w * (W + 2) + r];
        partial[((size_t)blockIdx.y * U + u) * (W + 2) + r] = v;
    }
}

// dense correction + assembly of the conv/BN-1 parameter gradients (one block per unit).  The `nsplit` slices of
// `partial` (one per CTA of the scatter kernel: up to one per SM) are summed by four thread rows, slices c = r, r + 4, ...
// each, then across the rows in a fixed order.
constexpr int BF_COLS = 128;        // >= 4F + 2 (filter_size <= 24)
__global__ void __launch_bounds__(4 * BF_COLS) conv_bwd_final_kernel(const float *__restrict__ partial, int nsplit, int U, int F,
                                                                     const int *__restrict__ M, const float *__restrict__ Mw,
                                                                     const float *__restrict__ bias, const float *__restrict__ mean,
                                                                     const float *__restrict__ var, const float *__restrict__ gamma,
                                                                     float eps, double N, float *__restrict__ dW,
                                                                     float *__restrict__ dbias, float *__restrict__ dgamma,
                                                                     float *__restrict__ dbeta)
{
    const int u = blockIdx.x, W = 4 * F;
    const int i = threadIdx.x & (BF_COLS - 1), r = threadIdx.x / BF_COLS;
    __shared__ double part[4][BF_COLS];
    double acc = 0.0;
    if (i < W + 2) {
        const float *src = partial + (size_t)u * (W + 2) + i;
#pragma unroll 4
        for (int c = r; c < nsplit; c += 4) acc += (double)src[(size_t)c * U * (W + 2)];
    }
    part[r][i] = acc;
    __syncthreads();
    if (r != 0 || i >= W + 2) return;
    const double A = ((part[0][i] + part[1][i]) + part[2][i]) + part[3][i];
    const double dg = ((part[0][W] + part[1][W]) + part[2][W]) + part[3][W];
    const double db = ((part[0][W + 1] + part[1][W + 1]) + part[2][W + 1]) + part[3][W + 1];
    if (i == W) {
        if (dgamma) dgamma[u] = (float)dg;
        if (dbias) dbias[u] = 0.0f;      // exactly zero: the bias is followed by a train-mode BatchNorm
    }
    if (i == W + 1 && dbeta) dbeta[u] = (float)db;
    if (!dW || i >= W) return;
    const double inv = 1.0 / sqrt((double)var[u] + (double)eps);
    const double mc = (double)mean[u] - (double)bias[u];
    const double m1 = db / N, m2 = dg / N;
    const double g = (double)gamma[u] * inv;
    const double cnt = (double)M[(size_t)i * W + i];
    const double R = inv * ((double)Mw[(size_t)u * W + i] - mc * cnt);
    dW[(size_t)u * W + i] = (float)(g * (A - m1 * cnt - m2 * R));
}

// ------------------------------------------------------------------------------------------------
// fused loss (mean reduction) + d(loss)/d(logits).  Up to 64 blocks write partial sums; the last block to
// finish adds them in block order (deterministic) and resets the ticket counter for the next call.
// scratch: LOSS_MAX_BLOCKS doubles + one unsigned ticket (zero-initialised once by the caller).
// ------------------------------------------------------------------------------------------------
constexpr int LOSS_MAX_BLOCKS = 64;

__global__ void __launch_bounds__(256) loss_kernel(const float *__restrict__ logits, const float *__restrict__ target,
                                                   long long n, int kind, float grad_scale, float *__restrict__ loss,
                                                   float *__restrict__ dlogits, double *__restrict__ partial,
                                                   unsigned int *__restrict__ ticket)
{
    __shared__ double red[8];
    __shared__ bool last;
    double acc = 0.0;
    const float inv_n = 1.0f / (float)n;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float x = logits[i], y = target[i];
        float l, g;
        if (kind == EXPLAINN_LOSS_BCE_WITH_LOGITS) {
            // max(x,0) - x*y + log(1 + exp(-|x|)) ; gradient sigmoid(x) - y
            const float e = expf(-fabsf(x));
            l = fmaxf(x, 0.f) - x * y + log1pf(e);
            const float sg = x >= 0.f ? 1.f / (1.f + e) : e / (1.f + e);
            g = sg - y;
        } else {
            const float d = x - y;
            l = d * d;
            g = 2.f * d;
        }
        acc += (double)l;
        if (dlogits) dlogits[i] = g * inv_n * grad_scale;
    }
    acc = warp_sum_d(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double v = 0.0;
        for (int w = 0; w < 8; ++w) v += red[w];
        partial[blockIdx.x] = v;
        __threadfence();
        last = atomicAdd(ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (last && threadIdx.x < 64) {
        // the last block adds the partial sums: one per thread, then a fixed shuffle tree (deterministic)
        static_assert(LOSS_MAX_BLOCKS == 64, "two warps");
        __threadfence();
        double v = threadIdx.x < gridDim.x ? ((volatile double *)partial)[threadIdx.x] : 0.0;
        v = warp_sum_d(v);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
        asm volatile("bar.sync 1, 64;" ::: "memory");
        if (threadIdx.x == 0) {
            *loss = (float)((red[0] + red[1]) / (double)n);
            *ticket = 0u;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// fused Adam over one flat buffer (torch.optim.Adam, amsgrad=False, weight_decay=0, maximize=False)
//   m.lerp_(g, 1-b1); v = b2*v + (1-b2)*g*g; p -= step_size * m / (sqrt(v)/sqrt(bc2) + eps)
// `step` is a device scalar that this kernel's launch wrapper does not read: bias corrections are
// computed on the device from *step (float, already incremented by adam_tick_kernel).
// ------------------------------------------------------------------------------------------------
__global__ void adam_tick_kernel(float *step) { *step += 1.0f; }

__global__ void __launch_bounds__(256) adam_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m,
                                                   float *__restrict__ v, long long n, const float *__restrict__ step,
                                                   float lr, float b1, float b2, float eps, float grad_scale)
{
    // bias corrections: double-precision pow, once per CTA (every thread doing it cost more than the update itself)
    __shared__ float s_step_size, s_bc2s;
    if (threadIdx.x == 0) {
        const double t = (double)*step;
        const double bc1 = 1.0 - pow((double)b1, t), bc2 = 1.0 - pow((double)b2, t);
        s_step_size = (float)((double)lr / bc1);
        s_bc2s = (float)sqrt(bc2);
    }
    __syncthreads();
    const float step_size = s_step_size, bc2s = s_bc2s;
    const float w1 = 1.0f - b1, w2 = 1.0f - b2;
    auto upd = [&](float &pi, float gi_raw, float &mi, float &vi) {
        const float gi = gi_raw * grad_scale;
        mi = fmaf(gi - mi, w1, mi);
        vi = fmaf(gi * gi, w2, vi * b2);
        const float denom = sqrtf(vi) / bc2s + eps;
        pi = pi - step_size * (mi / denom);
    };
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x, nthr = (long long)gridDim.x * blockDim.x;
    const bool vec = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                       reinterpret_cast<uintptr_t>(v)) & 15u) == 0;
    const long long n4 = vec ? n >> 2 : 0;
    for (long long i = tid; i < n4; i += nthr) {
        float4 pi = reinterpret_cast<float4 *>(p)[i], mi = reinterpret_cast<float4 *>(m)[i], vi = reinterpret_cast<float4 *>(v)[i];
        const float4 gi = reinterpret_cast<const float4 *>(g)[i];
        upd(pi.x, gi.x, mi.x, vi.x);
        upd(pi.y, gi.y, mi.y, vi.y);
        upd(pi.z, gi.z, mi.z, vi.z);
        upd(pi.w, gi.w, mi.w, vi.w);
        reinterpret_cast<float4 *>(m)[i] = mi;
        reinterpret_cast<float4 *>(v)[i] = vi;
        reinterpret_cast<float4 *>(p)[i] = pi;
    }
    for (long long i = 4 * n4 + tid; i < n; i += nthr) {
        float pi = p[i], mi = m[i], vi = v[i];
        upd(pi, g[i], mi, vi);
        m[i] = mi;
        v[i] = vi;
        p[i] = pi;
    }
}

__global__ void reduce_partials_kernel(float *__restrict__ out, const float *__restrict__ part, long long n, int k)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float v = 0.f;
    for (int c = 0; c < k; ++c) v += part[(size_t)c * n + i];
    out[i] = v;
}

// the same four elements at a time (n a multiple of 4, 16-byte aligned buffers); same summation order per element
__global__ void reduce_partials4_kernel(float4 *__restrict__ out, const float4 *__restrict__ part, long long n4, int k)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int c = 0; c < k; ++c) {
        const float4 x = part[(size_t)c * n4 + i];
        v.x += x.x; v.y += x.y; v.z += x.z; v.w += x.w;
    }
    out[i] = v;
}

}  // namespace

extern "C" int explainn_input_cooc_stats(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int F, int *M,
                                         cudaStream_t stream)
{
    XB_REQUIRE(codes && nmask && M && B >= 0 && F >= 1 && F <= 4 * XB_MAX_G && L >= F, "input_cooc_stats: bad argument");
    const int W = 4 * F;
    cudaError_t e = cudaMemsetAsync(M, 0, (size_t)W * W * sizeof(int), stream);
    if (e != cudaSuccess) { xb_set_error("cudaMemsetAsync: %s", cudaGetErrorString(e)); return XB_ERR_CUDA; }
    if (B == 0) return XB_OK;
    const int npairs = F * (F + 1) / 2;
    const int NT = ((npairs + 31) / 32) * 32;
    const size_t smem = (size_t)4 * (xb_words_mask(L) + 1) * 4;
    int grid = 2 * xb_num_sms();         // the kernel is latency-bound (6 warps per CTA): two CTAs per SM
    if (grid > B) grid = (int)B;
    cudaFuncSetAttribute(cooc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cooc_kernel<<<grid, NT, smem, stream>>>(codes, nmask, B, L, F, xb_words_codes(L), xb_words_mask(L), M);
    return xb_check_launch("input_cooc_stats");
}

extern "C" int explainn_bn1_batch_stats(const int *M, const float *w, const float *bias, int U, int F, long long n_pos,
                                        float momentum, float *run_mean, float *run_var, long long *num_batches_tracked,
                                        float *mean, float *var, float *Mw, cudaStream_t stream)
{
    XB_REQUIRE(M && w && bias && mean && var && Mw && U >= 1 && F >= 1 && F <= 4 * XB_MAX_G && n_pos >= 1,
               "bn1_batch_stats: bad argument");
    bn1_stats_kernel<<<U, 128, 0, stream>>>(M, w, bias, F, (double)n_pos, momentum, run_mean, run_var, num_batches_tracked, mean,
                                           var, Mw);
    return xb_check_launch("bn1_batch_stats");
}

// slices of `partial` for conv_bwd_sparse_kernel
static int sparse_nsplit(int U)
{
    const int target = 5 * xb_num_sms();     // 5 CTAs of 4 warps fit on an SM (39 KB of accumulators each)
    int ns = (target + U - 1) / U;
    return ns < 1 ? 1 : ns;
}

// slices to allocate: one per SM selects the tensor-core scatter (conv_bwd_tc.cu) where the shape allows it; a caller
// may pass any smaller nsplit >= 1 (synchronised BatchNorm does, to keep the all-reduced buffer small) and gets the
// shared-memory scatter kernel
extern "C" int explainn_conv_bwd_nsplit(int U)
{
    return std::max(sparse_nsplit(U), xb_num_sms());
}

// phases: 1 = sparse scatter into `partial`, 2 = dense correction + assembly; n_pos_total = B*S of the batch the
// statistics (M, Mw, mean, var) describe (the whole data-parallel batch under synchronised BatchNorm)
static int conv_pool_bwd_impl(int phases, long long n_pos_total, const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F, int psize, int pstride,
                              const float *gpool, const float *pooled, const float *xhat, const unsigned char *idx,
                              const unsigned short *idx16, const int *M, const float *Mw, const float *bias, const float *mean,
                              const float *var, const float *gamma, float eps, float *partial, int nsplit, float *dW, float *dbias,
                              float *dgamma, float *dbeta, cudaStream_t stream)
{
    XB_REQUIRE(codes && nmask && gpool && pooled && xhat && (idx || idx16) && M && Mw && bias && mean && var && gamma && partial,
               "conv_pool_bwd: null pointer");
    XB_REQUIRE(B >= 1 && U >= 1 && F >= 1 && F <= 4 * XB_MAX_G && psize >= 1 && pstride >= 1 && L >= F + psize - 1 && nsplit >= 1,
               "conv_pool_bwd: bad sizes");
    const int S = L - F + 1, P = (S - psize) / pstride + 1, W = 4 * F;
    XB_REQUIRE(B * P < (1LL << 31), "conv_pool_bwd: B*P = %lld entries per unit exceed 2^31", B * P);
    const size_t smem = ((size_t)BW_WARPS * W * 32 + (size_t)BW_WARPS * (W + 2)) * sizeof(float);
#define XB_LAUNCH_SPARSE(FTV)                                                                                                  \
    do {                                                                                                                      \
        cudaFuncSetAttribute(conv_bwd_sparse_kernel<FTV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);            \
        conv_bwd_sparse_kernel<FTV><<<dim3(U, nsplit), BW_WARPS * 32, smem, stream>>>(                                        \
            codes, nmask, B, L, F, P, xb_words_codes(L), xb_words_mask(L), gpool, pooled, xhat, idx, idx16, pstride, partial, U); \
    } while (0)
    // both phases derive the slice count from the same arguments: the tensor-core scatter (default pool, one slice per
    // CTA) when the caller's `partial` has room for it, else the shared-memory scatter
    const int tc_slices = (idx && psize == XB_POOL && pstride == XB_POOL) ? xb_conv_bwd_tc_slices(B, L, U, F, nsplit, gpool, pooled, xhat) : 0;
    nsplit = tc_slices > 0 ? tc_slices : std::min(nsplit, sparse_nsplit(U));
    if ((phases & 1) && tc_slices > 0) {
        int rc = xb_conv_bwd_tc(codes, nmask, B, L, U, F, gpool, pooled, xhat, idx, partial, nsplit, stream);
        if (rc != XB_OK) return rc;
    } else if (phases & 1) {
        if (F == 19) XB_LAUNCH_SPARSE(19);
        else if (F == 20) XB_LAUNCH_SPARSE(20);
        else XB_LAUNCH_SPARSE(0);
        int rc = xb_check_launch("conv_bwd_sparse");
        if (rc != XB_OK) return rc;
    }
#undef XB_LAUNCH_SPARSE
    if (phases & 2) {
        const double N = n_pos_total > 0 ? (double)n_pos_total : (double)B * S;
        conv_bwd_final_kernel<<<U, 4 * BF_COLS, 0, stream>>>(partial, nsplit, U, F, M, Mw, bias, mean, var, gamma, eps, N, dW, dbias, dgamma, dbeta);
    }
    return xb_check_launch("conv_bwd_final");
}

extern "C" int explainn_conv_pool_bwd(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                      const float *gpool, const float *pooled, const float *xhat, const unsigned char *idx,
                                      const int *M, const float *Mw, const float *bias, const float *mean, const float *var,
                                      const float *gamma, float eps, float *partial, int nsplit, float *dW, float *dbias,
                                      float *dgamma, float *dbeta, cudaStream_t stream)
{
    return conv_pool_bwd_impl(3, 0, codes, nmask, B, L, U, F, XB_POOL, XB_POOL, gpool, pooled, xhat, idx, nullptr, M, Mw, bias, mean, var, gamma,
                              eps, partial, nsplit, dW, dbias, dgamma, dbeta, stream);
}

// any (pool_size, pool_stride); idx16 (B,U,P) u16 = offset of the window maximum (explainn_pool_general)
extern "C" int explainn_conv_pool_bwd_general(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                              int pool_size, int pool_stride, const float *gpool, const float *pooled,
                                              const float *xhat, const unsigned short *idx16, const int *M, const float *Mw,
                                              const float *bias, const float *mean, const float *var, const float *gamma, float eps,
                                              float *partial, int nsplit, float *dW, float *dbias, float *dgamma, float *dbeta,
                                              cudaStream_t stream)
{
    return conv_pool_bwd_impl(3, 0, codes, nmask, B, L, U, F, pool_size, pool_stride, gpool, pooled, xhat, nullptr, idx16, M, Mw, bias, mean, var,
                              gamma, eps, partial, nsplit, dW, dbias, dgamma, dbeta, stream);
}

// explainn_conv_pool_bwd in two separately callable phases (1 sparse scatter, 2 dense correction) for synchronised
// BatchNorm: between them the caller averages `partial` over the ranks; M / Mw / mean / var then describe the whole
// data-parallel batch and n_pos_total is its B*S.
extern "C" int explainn_conv_pool_bwd_phase(int phases, long long n_pos_total, const uint32_t *codes, const uint32_t *nmask,
                                            long long B, int L, int U, int F, const float *gpool, const float *pooled,
                                            const float *xhat, const unsigned char *idx, const int *M, const float *Mw,
                                            const float *bias, const float *mean, const float *var, const float *gamma, float eps,
                                            float *partial, int nsplit, float *dW, float *dbias, float *dgamma, float *dbeta,
                                            cudaStream_t stream)
{
    XB_REQUIRE(phases >= 1 && phases <= 3, "conv_pool_bwd_phase: bad phase mask");
    return conv_pool_bwd_impl(phases, n_pos_total, codes, nmask, B, L, U, F, XB_POOL, XB_POOL, gpool, pooled, xhat, idx, nullptr, M, Mw,
                              bias, mean, var, gamma, eps, partial, nsplit, dW, dbias, dgamma, dbeta, stream);
}

extern "C" long long explainn_loss_scratch_bytes(void) { return (long long)LOSS_MAX_BLOCKS * sizeof(double) + 16; }

extern "C" int explainn_loss_fwd_bwd(const float *logits, const float *target, long long B, int C, int kind, float grad_scale,
                                     float *loss, float *dlogits, void *scratch, cudaStream_t stream)
{
    XB_REQUIRE(logits && target && loss && scratch && B >= 1 && C >= 1, "loss_fwd_bwd: bad argument");
    XB_REQUIRE(kind == EXPLAINN_LOSS_BCE_WITH_LOGITS || kind == EXPLAINN_LOSS_MSE, "loss_fwd_bwd: unknown loss %d", kind);
    const long long n = B * C;
    int blocks = (int)std::min<long long>(LOSS_MAX_BLOCKS, (n + 1023) / 1024);
    if (blocks < 1) blocks = 1;
    double *partial = reinterpret_cast<double *>(scratch);
    unsigned int *ticket = reinterpret_cast<unsigned int *>(partial + LOSS_MAX_BLOCKS);
    loss_kernel<<<blocks, 256, 0, stream>>>(logits, target, n, kind, grad_scale, loss, dlogits, partial, ticket);
    return xb_check_launch("loss_fwd_bwd");
}

extern "C" int explainn_adam_step(float *param, const float *grad, float *exp_avg, float *exp_avg_sq, long long n, float *step,
                                  float lr, float beta1, float beta2, float eps, float grad_scale, cudaStream_t stream)
{
    XB_REQUIRE(param && grad && exp_avg && exp_avg_sq && step && n >= 0, "adam_step: bad argument");
    adam_tick_kernel<<<1, 1, 0, stream>>>(step);
    if (n > 0) {
        long long blocks = (n / 4 + 255) / 256 + 1;
        const long long cap = 4LL * xb_num_sms();
        if (blocks > cap) blocks = cap;
        adam_kernel<<<(unsigned)blocks, 256, 0, stream>>>(param, grad, exp_avg, exp_avg_sq, n, step, lr, beta1, beta2, eps, grad_scale);
    }
    return xb_check_launch("adam_step");
}

// ---- up to four int64 device counters += 1 in one launch (BatchNorm num_batches_tracked, the dropout step counter) ----
__global__ void counters_inc_kernel(long long *c0, long long *c1, long long *c2, long long *c3)
{
    long long *c[4] = {c0, c1, c2, c3};
    if (threadIdx.x < 4 && c[threadIdx.x]) *c[threadIdx.x] += 1;
}

extern "C" int explainn_counters_inc(long long *c0, long long *c1, long long *c2, long long *c3, cudaStream_t stream)
{
    counters_inc_kernel<<<1, 32, 0, stream>>>(c0, c1, c2, c3);
    return xb_check_launch("counters_inc");
}

extern "C" int explainn_reduce_partials(float *out, const float *partials, long long n, int k, cudaStream_t stream)
{
    XB_REQUIRE(out && partials && n >= 0 && k >= 1, "reduce_partials: bad argument");
    if (n == 0) return XB_OK;
    if (n % 4 == 0 && ((reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(partials)) & 15u) == 0)
        reduce_partials4_kernel<<<(unsigned)((n / 4 + 255) / 256), 256, 0, stream>>>(reinterpret_cast<float4 *>(out),
                                                                                  reinterpret_cast<const float4 *>(partials), n / 4, k);
    else
        reduce_partials_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(out, partials, n, k);
    return xb_check_launch("reduce_partials");
}
