// conv_tc.cu -- tensor-core (tcgen05 / TMEM) variant of the fused conv + BN + exp + maxpool forward.
//
// Replaces the same reference lines as conv.cu (explainn/models/networks.py:98-102,153) for the two paths
// that tolerate fp32 rounding differences (inference and the training forward); the interpretation pass keeps
// the bit-exact table kernel of conv.cu.
//
// Formulation.  The conv on one-hot DNA is the GEMM  D[u][n] = sum_k A[u][k] * X[n][k]  with
//   A[u][k = 4*tap + c] = w'[u][c][tap]  (BatchNorm folded in),  A[u][4F] = shift[u],
//   X[n][4*tap + c]     = 1 if base(n + tap) == c,               X[n][4F] = 1.
// X is exactly representable in fp16; the fp32 weights are split A = A_hi + A_lo (two fp16 pieces, ~22
// significant bits), so  D = A_hi X^T + A_lo X^T  with every product exact and fp32 accumulation in TMEM.
// Units are the M dimension (one TMEM lane per unit), positions the N dimension (columns), so each epilogue
// thread owns one unit and pools 7 consecutive columns in registers.
//
// One CTA per SM, persistent.  A (one 128-unit tile, hi+lo, 40 KB) stays in shared memory; for every
// (sequence, 34-window chunk) a builder warp expands the 2-bit codes into the one-hot operand (16 bytes per
// position, read by the tensor core through overlapping K-major core matrices), an issue warp issues 2*K/16
// tcgen05.mma into one of two TMEM accumulators (240 columns each) and the epilogue warps drain the other one.
#include "tc_common.cuh"
#include "../../include/explainn_b200.h"
#include <stdlib.h>
#include <algorithm>

namespace {

constexpr int TC_M = 128;          // units per tile (TMEM lanes)
constexpr int TC_N = 240;          // positions per tile (TMEM columns), 34 windows of 7 = 238 used
constexpr int TC_WIN = 34;         // pool windows per tile
constexpr int TC_THREADS = 512;    // 16 warps: lane quarter = warp & 3, column range = warp >> 2

// Phase clocks (experiments only, -DXB_PHASE_CLOCKS): CTA 0 accumulates clock64() differences per phase; slots 0-7 by
// the producer lane, 8-15 by thread 0 (epilogue group 0).  tools/phase_clocks_conv.py prints them per tile.
#ifdef XB_PHASE_CLOCKS
__device__ long long g_conv_clk[16];
#define XB_CLK_DECL long long clk_t = clock64(), clk_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}
#define XB_CLK(i) do { const long long clk_n = clock64(); clk_acc[i] += clk_n - clk_t; clk_t = clk_n; } while (0)
#define XB_CLK_FLUSH(base, cond, ntiles) do { clk_acc[7] = (ntiles); if (blockIdx.x == 0 && (cond)) for (int ci = 0; ci < 8; ++ci) g_conv_clk[(base) + ci] = clk_acc[ci]; } while (0)
#else
#define XB_CLK_DECL do { } while (0)
#define XB_CLK(i) do { } while (0)
#define XB_CLK_FLUSH(base, cond, ntiles) do { } while (0)
#endif

// ------------------------------------------------------------------------------------------------
// A-operand images: per 128-unit tile, [hi | lo] x canonical K-major layout
//   byte offset of element (row r, k) inside one image = (r/8)*KC*128 + (k/8)*128 + (r%8)*16 + (k%8)*2
// Row u holds  sgn_u * 2^e_u * w'[u][:]  with w' the BatchNorm-folded filter:
//   2^e_u  brings the row's largest tap into [128, 256), so that neither piece of the fp16 split can overflow
//          (|w'| > 65504) or lose its low piece to the fp16 subnormal range (|w'| < 2^-3 * row maximum at worst:
//          absolute error 2^-25 against a row maximum of 2^7);
//   sgn_u  = sign of the training-mode BatchNorm gain (sign_of == NULL: +1): max-pooling z = gamma * x_hat + beta is
//          then a plain maximum of the accumulators for either sign of gamma.
// shift_rs[u] = shift (fp32), shift_rs[U + u] = sgn_u * 2^-e_u: the epilogue evaluates fma(acc, rs, shift), which is
// exact in the scale (a power of two) and so equal to the unscaled sum.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) tc_prepare_kernel(const float *__restrict__ w, const float *__restrict__ bias,
                                                         const float *__restrict__ gamma, const float *__restrict__ beta,
                                                         const float *__restrict__ mean, const float *__restrict__ var, float eps,
                                                         const float *__restrict__ sign_of, int U, int F, int K,
                                                         __half *__restrict__ images, float *__restrict__ shift_rs)
{
    const int mt = blockIdx.x, KC = K / 8;
    __half *hi = images + (size_t)mt * 2 * TC_M * K, *lo = hi + (size_t)TC_M * K;
    // grid (n_mt, 8): block y handles 16 rows of the tile, 16 threads per row
    const int r = blockIdx.y * 16 + (threadIdx.x >> 4), l16 = threadIdx.x & 15;
    const int u = mt * TC_M + r;
    float sc = 0.f;
    if (u < U) sc = __fdiv_rn(gamma[u], __fsqrt_rn(__fadd_rn(var[u], eps)));
    auto tap = [&](int k) -> float {
        if (u >= U || k >= 4 * F) return 0.0f;
        return __fmul_rn(w[((size_t)u * 4 + (k & 3)) * F + (k >> 2)], sc);
    };
    float mx = 0.f;
    for (int k = l16; k < K; k += 16) mx = fmaxf(mx, fabsf(tap(k)));
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    int e = 0;
    if (mx > 0.f && mx < INFINITY) {
        e = 7 - ilogbf(mx);
        e = max(-100, min(100, e));
    }
    const float sg = (sign_of && u < U && sign_of[u] < 0.f) ? -1.0f : 1.0f;
    const float up = ldexpf(sg, e);
    for (int k = l16; k < K; k += 16) {
        const float v = __fmul_rn(tap(k), up);                      // exact: power of two
        const __half h = __float2half_rn(v);
        const __half l = __float2half_rn(v - __half2float(h));
        const size_t off = (size_t)(r >> 3) * KC * 64 + (size_t)(k >> 3) * 64 + (r & 7) * 8 + (k & 7);
        hi[off] = h;
        lo[off] = l;
    }
    if (l16 == 0 && u < U) {
        shift_rs[u] = __fadd_rn(__fmul_rn(__fsub_rn(bias[u], mean[u]), sc), beta[u]);
        shift_rs[U + u] = ldexpf(sg, -e);
    }
}

// 3-input maximum (one FMNMX3 instead of two FMNMX)
__device__ __forceinline__ float fmax3(float a, float b, float c)
{
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// exp(z) to ~2 ulp in 10 instructions: z * log2(e) is split into an integer n (magic-number rounding) and a fraction
// f = z*log2e - n evaluated with a two-term log2(e) (error ~3e-8 absolute), 2^f comes from MUFU.EX2 (2^-22 relative),
// 2^n is added into the exponent field.  (__expf multiplies first and loses |z| * 2^-24 relative; the 25-instruction
// det_expf of the table kernel exists for bit-equality with the CPU oracle, which this kernel does not claim.)
__device__ __forceinline__ float tc_expf(float z)
{
    const float tm = __fmaf_rn(z, 1.44269502f, 12582912.0f);
    const float n = __fsub_rn(tm, 12582912.0f);
    float f = __fmaf_rn(z, 1.44269502f, -n);
    f = __fmaf_rn(z, 1.92596303e-8f, f);
    float p;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(p) : "f"(f));
    float r = __int_as_float(__float_as_int(p) + (int)((unsigned)__float_as_int(tm) << 23));
    r = z < -87.0f ? 0.0f : r;
    r = z > 88.7228394f ? __int_as_float(0x7f800000) : r;
    return z != z ? z : r;
}

struct TcArgs {
    const uint32_t *codes, *nmask;
    long long B;
    int L, F, U, K, P, wc, wm;
    const __half *images;
    const float *shift;     // (2U): shift, then the signed power-of-two row scale (see tc_prepare_kernel)
    float *pooled;          // (B,U,P)
    float *xhat;            // train only
    unsigned char *idx;     // train only
    const float *gamma, *beta;
    int gsz;                // unit tiles (of 128) whose weights stay resident in shared memory together
    int single;             // 1: only the high fp16 piece of the weights is multiplied (stated-tolerance mode)
#ifdef XB_EXPERIMENT
    int dbg;                // XB_TC_DBG bitmask (experiment build only): 1 no stores, 2 no V build, 4 no drain, 8 no MMA
#endif
};
#ifdef XB_EXPERIMENT
#define XB_DBG(a, bit) ((a).dbg & (bit))
#else
#define XB_DBG(a, bit) 0
#endif

constexpr int TC_VROWS = TC_N + 64;     // dinucleotide entries per tile: positions n0 .. n0 + 240 + 2*KC
constexpr int TC_VBUFS = 4;             // V tiles in flight: tile n may be built while the MMAs of n-1, n-2 run

// Epilogue of one accumulator for the columns [C0, C1) (whole pool windows), one unit per thread.
// The accumulators are  sgn * 2^e * (z - shift)  (tc_prepare_kernel), so the window maximum is taken on the RAW values
// (three 3-input maxima per window), and the affine map and the exponential run once per window, not per position.
// Training also needs the first position of the maximum (MaxPool1d backward routing): found by comparing the seven
// values with the maximum afterwards.  Tie rule: first maximum of the normalised conv output x_hat (times sign(gamma));
// gamma == 0 makes z constant, where the first position wins as in the reference.
// Tensor-memory loads are double-buffered in batches of 32 columns (the loads of batch k+1 are in flight while batch k
// is pooled); warps whose lane quarter holds no valid unit skip the drain altogether (caller).
template <bool TRAIN, int C0, int C1>
__device__ __forceinline__ void tc_drain(uint32_t taddr, int row, int nwin, int sst, float *stage, unsigned char *stage_i, float shift,
                                         float rs, float gam, float bet)
{
    constexpr int plane = TC_M * (TC_WIN + 1);
    constexpr int P0 = C0 / 16, P1 = (C1 + 15) / 16;
    const bool gzero = TRAIN && gam == 0.0f;
    float m = 0.f, pa = 0.f, pb_ = 0.f, w0 = 0.f;      // running maximum, the one or two values waiting for their partner
    int ib = 0;
    auto column = [&](int col, uint32_t bits) {
        static_assert(XB_POOL == 7, "window maximum written for 7 positions: (0,1,2) (3,4) (5,6)");
        const int ph = col % XB_POOL, win = col / XB_POOL;
        const float v = __uint_as_float(bits);
        if (ph == 0) { pa = v; if (TRAIN) w0 = v; }
        else if (ph == 1) pb_ = v;
        else if (ph == 2) {
            m = fmax3(pa, pb_, v);
            if (TRAIN) ib = pa == m ? 0 : (pb_ == m ? 1 : 2);
        } else if (ph == 3 || ph == 5) pa = v;
        else {                                          // ph == 4 or 6: fold the pair (ph-1, ph) into the maximum
            const float m2 = fmax3(m, pa, v);
            if (TRAIN) ib = m2 == m ? ib : (pa == m2 ? ph - 1 : ph);       // an earlier maximum keeps its position
            m = m2;
        }
        if (ph == XB_POOL - 1 && win < nwin) {
            if (TRAIN) {
                if (gzero) { ib = 0; m = w0; }
                const float xh = __fmaf_rn(m, rs, shift);
                stage[row * sst + win] = tc_expf(__fmaf_rn(gam, xh, bet));
                stage[plane + row * sst + win] = xh;
                stage_i[row * sst + win] = (unsigned char)ib;
            } else {
                // inference: the 2-instruction exp (relative error |z| * 2^-24 from the multiplication by log2(e), within
                // the 1e-5 the tests ask of this backend); the epilogue's instruction count sets the pace of this kernel
                stage[row * sst + win] = __expf(__fmaf_rn(m, rs, shift));
            }
        }
    };
    constexpr int BATCH = TRAIN ? 1 : 2;        // the training epilogue keeps three outputs per window: fewer columns in flight
    constexpr int NB = (P1 - P0 + BATCH - 1) / BATCH;
    uint32_t rr[2][BATCH * 16] = {};
#pragma unroll
    for (int p = 0; p < BATCH; ++p)
        if (P0 + p < P1) tc_ld16_nowait(taddr + (P0 + p) * 16, rr[0] + p * 16);
#pragma unroll
    for (int bi = 0; bi < NB; ++bi) {
        const int pb = P0 + bi * BATCH;
        uint32_t *r = rr[bi & 1];
        if constexpr (BATCH == 2) tc_wait_ld32(r); else tc_wait_ld16(r);
        if (bi + 1 < NB) {
#pragma unroll
            for (int p = 0; p < BATCH; ++p)
                if (pb + BATCH + p < P1) tc_ld16_nowait(taddr + (pb + BATCH + p) * 16, rr[(bi + 1) & 1] + p * 16);
        }
#pragma unroll
        for (int i = 0; i < BATCH * 16; ++i) {
            const int col = pb * 16 + i;
            if (col >= C0 && col < C1) column(col, r[i]);
        }
    }
}

// item iterator shared by both roles: (tile group, sequence, chunk), advanced without divisions
struct TcIter {
    int grp, chunk, nchunk;
    long long b, B, it, it1;
    __device__ TcIter(long long it0, long long it1_, long long B_, int nchunk_) : nchunk(nchunk_), B(B_), it(it0), it1(it1_)
    {
        const long long per_grp = B_ * nchunk_;
        grp = (int)(it0 / per_grp);
        const long long rem = it0 - (long long)grp * per_grp;
        b = rem / nchunk_;
        chunk = (int)(rem - b * nchunk_);
    }
    __device__ bool valid() const { return it < it1; }
    __device__ void next()
    {
        ++it;
        if (++chunk == nchunk) {
            chunk = 0;
            if (++b == B) { b = 0; ++grp; }
        }
    }
};

// Warp-specialised persistent kernel.
//   warp 17      builder: expands the packed codes of one (sequence, chunk) into the V operand ONCE (ring of four
//                buffers, v_full / v_empty mbarriers)
//   warp 16      issuer: one elected lane issues the tcgen05.mma of every resident unit tile against that operand (the
//                weights of up to `gsz` 128-unit tiles stay in shared memory) and commits each accumulator to an
//                mbarrier; the last commit of an item hands the operand buffer back to the builder
//   warps 0..15  epilogue: drain the accumulator (lane quarter = warp & 3, window range = warp >> 2), pool, exp,
//                stage, TMA bulk store
// Two TMEM accumulators (tmem_full / tmem_empty mbarriers), four V buffers and two staging buffers keep the
// tensor pipe, the epilogue and the stores of consecutive tiles in flight.
template <bool TRAIN>
__global__ void __launch_bounds__(TC_THREADS + 64, 1) conv_tc_kernel(const TcArgs a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar_a[4], tmem_full[2], tmem_empty[2], v_full[TC_VBUFS], v_empty[TC_VBUFS];
    __shared__ uint32_t tmem_base_s;

    const int K = a.K, KC = K / 8, KS = K / 16;
    const uint32_t a_bytes = (uint32_t)TC_M * K * 2;        // one image (hi or lo)
    unsigned char *sA = smem;                               // gsz tiles x (hi, lo)
    // X operand: V[pos] = 16 bytes = one-hot fp16 of (base(pos), base(pos+1)), 16-byte pitch.  Row n of the im2col
    // matrix is V[n], V[n+2], V[n+4], ... so the K-major core-matrix layout is the V array itself with
    // LBO = 32 B (next K chunk = two positions on) and SBO = 128 B (next 8 rows): no expansion is ever written.
    uint4 *sV = reinterpret_cast<uint4 *>(sA + (size_t)a.gsz * 2 * a_bytes);   // TC_VBUFS buffers x TC_VROWS
    // staging, two buffers: [128][sst] floats (pooled) (+ x_hat floats and argmax bytes when TRAIN).  With a single
    // chunk per sequence the row stride is P itself: a plane is the exact image of pooled[b][u0..u0+127][:] and
    // leaves with ONE TMA bulk store.
    const int sst = a.P <= TC_WIN ? a.P : TC_WIN + 1;
    constexpr int plane = TC_M * (TC_WIN + 1);
    constexpr int stage_bytes = (TRAIN ? 9 : 4) * plane;
    unsigned char *stage_base = reinterpret_cast<unsigned char *>(sV + TC_VBUFS * TC_VROWS);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_mt = (a.U + TC_M - 1) / TC_M;
    const int n_grp = (n_mt + a.gsz - 1) / a.gsz;
    const int nchunk = (a.P + TC_WIN - 1) / TC_WIN;
    const long long total = a.B * nchunk * n_grp;
    const long long it0 = total * blockIdx.x / gridDim.x, it1 = total * (blockIdx.x + 1) / gridDim.x;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; ++i) mbar_init(&bar_a[i], 1);
        mbar_init(&tmem_full[0], 1);
        mbar_init(&tmem_full[1], 1);
        mbar_init(&tmem_empty[0], TC_THREADS / 64);      // the 8 warps of the epilogue group that owns the buffer
        mbar_init(&tmem_empty[1], TC_THREADS / 64);
        for (int i = 0; i < TC_VBUFS; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
        mbar_fence_init();
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;

    const int role = tc_warp_uniform();              // warp-uniform role branch: MMA operands stay in uniform registers
    if (role == TC_THREADS / 32) {
        // =========================== issue warp ===========================
        // waits for an operand buffer (built by the builder warp) and a free accumulator, issues the MMAs of every
        // resident unit tile and commits; the last commit of an item also hands its operand buffer back
        uint32_t a_parity = 1;                             // toggled when a group's load is issued: first group waits on 0
        int cur_grp = -1;
        long long n = 0;                                   // tiles issued by this CTA
        long long vi = 0;                                  // items (V tiles) consumed by this CTA
        XB_CLK_DECL;
        for (TcIter t(it0, it1, a.B, nchunk); t.valid(); t.next(), ++vi) {
            XB_CLK(0);                                     // loop overhead
            const int mt0 = t.grp * a.gsz, mtn = min(a.gsz, n_mt - mt0);
            if (t.grp != cur_grp) {
                // new group of unit tiles: every MMA that reads sA must have completed before TMA overwrites it
                if (n > 0) mbar_wait(&tmem_full[(n - 1) & 1], (uint32_t)(((n - 1) >> 1) & 1));
                // one barrier per image: the MMAs of the first tile start as soon as ITS weights have landed
                if (lane == 0) {
                    for (int m = 0; m < mtn; ++m) {
                        mbar_arrive_expect_tx(&bar_a[m], 2 * a_bytes);
                        tma_bulk_g2s(sA + (size_t)m * 2 * a_bytes, reinterpret_cast<const unsigned char *>(a.images) + (size_t)(mt0 + m) * 2 * a_bytes,
                                     2 * a_bytes, &bar_a[m]);
                    }
                }
                a_parity ^= 1;
                cur_grp = t.grp;
            }
            const int slot = (int)(vi & (TC_VBUFS - 1));
            const uint4 *V = sV + slot * TC_VROWS;
            // f16 x f16 -> f32, M = 128, N = the columns this chunk's pool windows need (multiple of 16, at most 240):
            // a short last chunk (P = 140: 4 windows) or a short sequence (P = 32) does not pay for 240 columns
            const int ncol = (min(TC_WIN, a.P - t.chunk * TC_WIN) * XB_POOL + 15) & ~15;
            const uint32_t idesc = (1u << 4) | ((uint32_t)(ncol >> 3) << 17) | ((uint32_t)(TC_M >> 4) << 24);
            mbar_wait(&v_full[slot], (uint32_t)((vi / TC_VBUFS) & 1));
            XB_CLK(1);                                     // wait for the operand buffer
            for (int m = 0; m < mtn; ++m, ++n) {
                const int buf = (int)(n & 1);
                const uint32_t use = (uint32_t)(n >> 1);   // how often this accumulator has been used before
                mbar_wait(&bar_a[m], a_parity);                      // this tile's weights (immediate after the first item)
                mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);          // drained by the epilogue warps
                XB_CLK(2);                                 // wait for a free accumulator
                // descriptors of K step 0, computed by the whole warp; a K step advances A by 256 bytes and V by 64
                const uint32_t d = tmem_base + (uint32_t)buf * 256u;
                const uint32_t a_hi = smem_u32(sA) + (uint32_t)m * 2 * a_bytes;
                const uint64_t dA_hi = make_smem_desc(a_hi, 128, (uint32_t)KC * 128), dA_lo = make_smem_desc(a_hi + a_bytes, 128, (uint32_t)KC * 128);
                const uint64_t dV = make_smem_desc(smem_u32(V), 32, 128);
                if (tc_elect_one()) {
                    tc_fence_after();
                    for (int ks = 0; ks < (XB_DBG(a, 8) ? 0 : KS); ++ks) {
                        tc_mma_f16(d, dA_hi + ks * 16, dV + ks * 4, idesc, ks > 0 ? 1u : 0u);
                        if (!a.single) tc_mma_f16(d, dA_lo + ks * 16, dV + ks * 4, idesc, 1u);
                    }
                    tc_commit(&tmem_full[buf]);
                    if (m == mtn - 1) tc_commit(&v_empty[slot]);     // all MMAs reading this operand buffer have retired
                }
                __syncwarp();
                XB_CLK(3);                                 // MMA issue
            }
        }
        XB_CLK_FLUSH(0, lane == 0, n);
    } else if (role == TC_THREADS / 32 + 1) {
        // =========================== builder warp ===========================
        // expands the packed codes of one (sequence, chunk) into the V operand, TC_VBUFS buffers ahead of the issue warp
        long long vi = 0;
        // an item touches positions n0 .. n0 + TC_VROWS: at most 20 code words and 11 mask words, one per lane;
        // they are prefetched two items ahead
        auto load_words = [&](const TcIter &t, uint32_t &cwv, uint32_t &mwv) {
            cwv = 0u; mwv = 0u;
            if (!t.valid()) return;
            const int n0 = t.chunk * TC_WIN * XB_POOL;
            const int ci = (n0 >> 4) + lane, mi = (n0 >> 5) + lane;
            if (ci < a.wc) cwv = a.codes[t.b * a.wc + ci];
            if (mi < a.wm) mwv = a.nmask[t.b * a.wm + mi];
        };
        uint32_t cw0, mw0, cw1, mw1;
        TcIter tp(it0, it1, a.B, nchunk);                  // prefetch cursor
        load_words(tp, cw0, mw0); tp.next();
        load_words(tp, cw1, mw1); tp.next();
        for (TcIter t(it0, it1, a.B, nchunk); t.valid(); t.next(), ++vi) {
            const uint32_t cwr = cw0, mwr = mw0;
            cw0 = cw1; mw0 = mw1;
            load_words(tp, cw1, mw1); tp.next();
            const int n0 = t.chunk * TC_WIN * XB_POOL;
            const int slot = (int)(vi & (TC_VBUFS - 1));
            mbar_wait(&v_empty[slot], (uint32_t)(((vi / TC_VBUFS) & 1) ^ 1));    // its previous user's MMAs have retired
            uint4 *V = sV + slot * TC_VROWS;
            {
                const uint32_t w_hi = __shfl_down_sync(0xffffffffu, cwr, 1);
                const uint32_t c16 = __funnelshift_r(cwr, w_hi, 2 * (n0 & 15));          // codes of my 16 positions
                const uint32_t c_nx = __shfl_down_sync(0xffffffffu, c16, 1) & 3u;         // code of position n0+16(l+1)
                const int mpos = (n0 & 31) + 16 * lane;
                const uint32_t mw_a = __shfl_sync(0xffffffffu, mwr, mpos >> 5);
                const uint32_t mw_b = __shfl_sync(0xffffffffu, mwr, (mpos >> 5) + 1);
                uint32_t m16 = __funnelshift_r(mw_a, mw_b, mpos & 31) & 0xFFFFu;          // N flags of my 16 positions
                const int nvalid_pos = a.L - (n0 + 16 * lane);                            // positions >= L count as N
                if (nvalid_pos < 16) m16 |= nvalid_pos <= 0 ? 0xFFFFu : (0xFFFFu << nvalid_pos) & 0xFFFFu;
                const uint32_t m_nx = __shfl_down_sync(0xffffffffu, m16, 1) & 1u;
                // one-hot fp16 quadruple of each of the 17 bases: (x,y) = halves (A,C) and (G,T)
                uint32_t ox[17], oy[17];
#pragma unroll
                for (int j = 0; j < 17; ++j) {
                    const uint32_t c = j < 16 ? (c16 >> (2 * j)) & 3u : c_nx;
                    const uint32_t isn = j < 16 ? (m16 >> j) & 1u : m_nx;
                    const uint32_t one = isn ? 0u : (0x3C00u << ((c & 1u) * 16));
                    ox[j] = (c & 2u) ? 0u : one;
                    oy[j] = (c & 2u) ? one : 0u;
                }
                if (lane < TC_VROWS / 16 && !(XB_DBG(a, 2) && vi >= TC_VBUFS)) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) V[16 * lane + j] = make_uint4(ox[j], oy[j], ox[j + 1], oy[j + 1]);
                }
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&v_full[slot]);
        }
    } else {
        // =========================== epilogue warps ===========================
        // Two groups of 8 warps; group g owns accumulator g and staging buffer g, i.e. every other tile, so the
        // drain / stage / store chain of one tile overlaps the next tile's.  Inside a group: lane quarter =
        // warp & 3, column half = (warp >> 2) & 1 (17 whole windows each).
        const int grp = warp >> 3;
        const int gtid = threadIdx.x & 255;                // thread index inside the group
        const int row = 32 * (warp & 3) + lane;            // TMEM lane = unit inside the tile
        float *stage = reinterpret_cast<float *>(stage_base + (size_t)grp * stage_bytes);
        unsigned char *stage_i = reinterpret_cast<unsigned char *>(stage) + 8 * plane;
        const uint32_t taddr = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)grp * 256u;
        constexpr int SPLIT = 17 * XB_POOL;
        long long n = 0;
        XB_CLK_DECL;
        for (TcIter t(it0, it1, a.B, nchunk); t.valid(); t.next()) {
            const int mt0 = t.grp * a.gsz, mtn = min(a.gsz, n_mt - mt0);
            const int w0 = t.chunk * TC_WIN;
            const int nwin = min(TC_WIN, a.P - w0);
            for (int m = 0; m < mtn; ++m, ++n) {
                if ((int)(n & 1) != grp) continue;
                const uint32_t use = (uint32_t)(n >> 1);
                const int u0 = (mt0 + m) * TC_M;
                const int u = u0 + row;
                const float shift = u < a.U ? __ldg(a.shift + u) : 0.f;
                const float rs = u < a.U ? __ldg(a.shift + a.U + u) : 1.f;
                float gam = 1.f, bet = 0.f;
                if (TRAIN && u < a.U) { gam = __ldg(a.gamma + u); bet = __ldg(a.beta + u); }
                XB_CLK(0);                                 // loop overhead, coefficient loads
                // the previous tile of this group was staged in the same buffer: its bulk stores must be done
                // reading before anybody overwrites it (a second staging buffer per group was tried: no gain)
                if (gtid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                XB_CLK(1);                                 // previous bulk store finished reading the staging buffer
                asm volatile("bar.sync %0, 256;" ::"r"(1 + grp) : "memory");
                XB_CLK(2);                                 // group barrier 1
                mbar_wait(&tmem_full[grp], use & 1);
                tc_fence_after();
                XB_CLK(3);                                 // wait for the accumulator
                if (u0 + 32 * (warp & 3) < a.U && !XB_DBG(a, 4)) {   // lane quarters without a valid unit are not read at all
                    if (((warp >> 2) & 1) == 0) tc_drain<TRAIN, 0, SPLIT>(taddr, row, nwin, sst, stage, stage_i, shift, rs, gam, bet);
                    else tc_drain<TRAIN, SPLIT, TC_WIN * XB_POOL>(taddr, row, nwin, sst, stage, stage_i, shift, rs, gam, bet);
                }
                // accumulator free again for the producer
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tmem_empty[grp]);
                XB_CLK(4);                                 // drain
                if (XB_DBG(a, 1)) continue;

                // ---- store the staged tile ----
                fence_proxy_async();
                asm volatile("bar.sync %0, 256;" ::"r"(1 + grp) : "memory");
                XB_CLK(5);                                 // group barrier 2
                const int nvalid = min(TC_M, a.U - u0);
                const size_t o = ((size_t)t.b * a.U + u0) * a.P;
                const uint32_t bytes = (uint32_t)nvalid * a.P * 4;
                const bool bulk_f = sst == a.P && (bytes & 15u) == 0 && ((o * 4) & 15u) == 0 && !XB_DBG(a, 16);
                const bool bulk_i = TRAIN && sst == a.P && ((bytes / 4) & 15u) == 0 && (o & 15u) == 0;
                if (gtid == 0 && (bulk_f || bulk_i)) {
                    if (bulk_f) {
                        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(a.pooled + o), "r"(smem_u32(stage)), "r"(bytes) : "memory");
                        if (TRAIN)
                            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(a.xhat + o), "r"(smem_u32(stage + plane)), "r"(bytes) : "memory");
                    }
                    if (bulk_i)
                        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(a.idx + o), "r"(smem_u32(stage_i)), "r"(bytes / 4) : "memory");
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                }
                if (sst == a.P) {
                    // single chunk per sequence: whatever cannot leave as a bulk store is copied flat
                    const int nel = nvalid * a.P;
                    if (!bulk_f)
                        for (int i = gtid; i < nel; i += 256) {
                            a.pooled[o + i] = stage[i];
                            if (TRAIN) a.xhat[o + i] = stage[plane + i];
                        }
                    if (TRAIN && !bulk_i)
                        for (int i = gtid; i < nel; i += 256) a.idx[o + i] = stage_i[i];
                } else {
                    for (int r = warp & 7; r < nvalid; r += 8) {
                        const size_t orow = o + (size_t)r * a.P + w0;
                        for (int c = lane; c < nwin; c += 32) {
                            a.pooled[orow + c] = stage[r * sst + c];
                            if (TRAIN) {
                                a.xhat[orow + c] = stage[plane + r * sst + c];
                                a.idx[orow + c] = stage_i[r * sst + c];
                            }
                        }
                    }
                }
                XB_CLK(6);                                 // store issue
            }
        }
        XB_CLK_FLUSH(8, threadIdx.x == 0, n);
        if (gtid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

int tc_K(int F) { return ((4 * F + 15) / 16) * 16; }

}  // namespace
#ifdef XB_PHASE_CLOCKS
extern "C" int xb_dbg_conv_clocks(long long *host) { return (int)cudaMemcpyFromSymbol(host, g_conv_clk, sizeof(g_conv_clk)); }
#endif
namespace {

}  // namespace

extern "C" long long explainn_conv_tc_image_halfs(int U, int F)
{
    const long long n_mt = (U + TC_M - 1) / TC_M;
    return n_mt * 2 * TC_M * tc_K(F);
}

extern "C" int explainn_conv_tc_prepare(const float *w, const float *bias, const float *gamma, const float *beta,
                                        const float *mean, const float *var, float eps, const float *sign_of, int U, int F,
                                        void *images, float *shift, cudaStream_t stream)
{
    XB_REQUIRE(w && bias && gamma && beta && mean && var && images && shift && U >= 1 && F >= 1, "conv_tc_prepare: bad argument");
    if (F > 4 * XB_MAX_G) {
        xb_set_error("conv_tc_prepare: filter_size %d not supported (max %d)", F, 4 * XB_MAX_G);
        return XB_ERR_UNSUPPORTED;
    }
    const int n_mt = (U + TC_M - 1) / TC_M;
    tc_prepare_kernel<<<dim3(n_mt, 8), 256, 0, stream>>>(w, bias, gamma, beta, mean, var, eps, sign_of, U, F, tc_K(F), (__half *)images, shift);
    return xb_check_launch("conv_tc_prepare");
}

static int launch_tc(bool train, bool single, const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F, const void *images,
                     const float *shift, const float *gamma, const float *beta, float *pooled, float *xhat, unsigned char *idx,
                     cudaStream_t stream)
{
    XB_REQUIRE(codes && nmask && images && shift && pooled && B >= 0 && U >= 1, "conv_pool_fwd_tc: bad argument");
    if (F < 1 || F > 4 * XB_MAX_G || L < F + XB_POOL - 1) {
        xb_set_error("conv_pool_fwd_tc: unsupported geometry L=%d filter_size=%d", L, F);
        return XB_ERR_UNSUPPORTED;
    }
    if (B == 0) return XB_OK;
    TcArgs a;
    a.codes = codes; a.nmask = nmask; a.B = B; a.L = L; a.F = F; a.U = U; a.K = tc_K(F);
    a.P = ((L - F + 1) - XB_POOL) / XB_POOL + 1;
    a.wc = xb_words_codes(L); a.wm = xb_words_mask(L);
#ifdef XB_EXPERIMENT
    { const char *e = getenv("XB_TC_DBG"); a.dbg = e ? atoi(e) : 0; }
#endif
    a.single = single ? 1 : 0;
    a.images = (const __half *)images; a.shift = shift; a.pooled = pooled; a.xhat = xhat; a.idx = idx; a.gamma = gamma; a.beta = beta;
    const int n_mt = (U + TC_M - 1) / TC_M, nchunk = (a.P + TC_WIN - 1) / TC_WIN;
    const size_t tile_bytes = (size_t)2 * TC_M * a.K * 2;
    const size_t fixed = (size_t)TC_VBUFS * TC_VROWS * 16 + (size_t)2 * (train ? 9 : 4) * TC_M * (TC_WIN + 1);
    const size_t budget = 227 * 1024 - 512;
    if (fixed + tile_bytes > budget) {
        xb_set_error("conv_pool_fwd_tc: %zu bytes of shared memory needed (L=%d, filter_size=%d)", fixed + tile_bytes, L, F);
        return XB_ERR_UNSUPPORTED;
    }
    a.gsz = (int)std::min<size_t>(std::min<size_t>((budget - fixed) / tile_bytes, (size_t)n_mt), 4);      // four weight barriers
    const size_t smem = fixed + (size_t)a.gsz * tile_bytes;
    const int n_grp = (n_mt + a.gsz - 1) / a.gsz;
    long long total = (long long)n_grp * B * nchunk;
    int grid = xb_num_sms();
    if (total < grid) grid = (int)total;
    cudaError_t e;
    if (train) {
        e = cudaFuncSetAttribute(conv_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) conv_tc_kernel<true><<<grid, TC_THREADS + 64, smem, stream>>>(a);
    } else {
        e = cudaFuncSetAttribute(conv_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) conv_tc_kernel<false><<<grid, TC_THREADS + 64, smem, stream>>>(a);
    }
    if (e != cudaSuccess) {
        xb_set_error("conv_pool_fwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    return xb_check_launch("conv_pool_fwd_tc");
}

extern "C" int explainn_conv_pool_fwd_tc(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                         const void *images, const float *shift, float *pooled, cudaStream_t stream)
{
    return launch_tc(false, false, codes, nmask, B, L, U, F, images, shift, nullptr, nullptr, pooled, nullptr, nullptr, stream);
}

extern "C" int explainn_conv_pool_fwd_tc_fp16(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                              const void *images, const float *shift, float *pooled, cudaStream_t stream)
{
    return launch_tc(false, true, codes, nmask, B, L, U, F, images, shift, nullptr, nullptr, pooled, nullptr, nullptr, stream);
}

extern "C" int explainn_conv_pool_fwd_train_tc(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                               const void *images, const float *shift, const float *gamma, const float *beta,
                                               float *pooled, float *xhat, unsigned char *idx, cudaStream_t stream)
{
    XB_REQUIRE(gamma && beta && xhat && idx, "conv_pool_fwd_train_tc: null pointer");
    return launch_tc(true, false, codes, nmask, B, L, U, F, images, shift, gamma, beta, pooled, xhat, idx, stream);
}
