// interp_tc.cu -- the two interpretation passes at tensor-core speed, with results identical to the table kernel.
//
// What it replaces in the reference (paths relative to /root/reference):
//   np.amax(activations, axis=(0, 2))  and  activations[j, i, :] > threshold[i]
//       explainn/interpretation/interpretation.py:187, 209  (on the activation map of :143-147)
//
// The bit-exact quantity is the canonical fp32 pre-activation z of conv.cu / oracle/explainn_oracle.c.  The tensor-core
// GEMM of conv_tc.cu yields z~ with |z~ - z| <= delta_u (explainn_interp_tc_bounds: 2^-15 of the largest partial sum of
// the unit, several times the worst case of the fp16 split, the fp32 accumulation in tensor memory and the canonical
// roundings).  So the GEMM DECIDES every position that is further than delta from a decision boundary, and only the
// rest -- a vanishing fraction -- is re-evaluated in the canonical order by a resolve kernel:
//   pass A  a position can only be the unit's maximum if z~ >= (running maximum of z~) - 2 delta: those candidates are
//           listed and their exact z is folded into zmax with an atomic maximum;
//   pass B  z~ > zhi + delta is a hit, z~ <= zlo - delta is not; in between the position is listed, evaluated exactly
//           and its bit is set by the resolve kernel.  Hit bits then go through hit_count_kernel of conv.cu.
// Kernel structure (builder warp / issue warp / 16 epilogue warps, two TMEM accumulators, operand as overlapping
// K-major core matrices) is that of conv_tc_kernel; tiles are 256 positions wide so that one tile is 8 whole words of
// the hit-bit rows and the two accumulators fill tensor memory exactly.
#include "tc_common.cuh"
#include "../../include/explainn_b200.h"
#include <algorithm>

namespace {

constexpr int IT_M = 128;
constexpr int IT_N = 256;               // positions per tile = 8 words of hit bits
constexpr int IT_THREADS = 512;
constexpr int IT_VROWS = IT_N + 64;
constexpr int IT_VBUFS = 4;

enum { IT_MAX = 0, IT_HITS = 1 };

struct ItArgs {
    const uint32_t *codes, *nmask;
    long long B;
    int L, F, U, K, S, W, wc, wm, gsz;
    const __half *images;
    const float *shift_rs;              // (2U) shift, signed inverse row scale (conv_tc_prepare)
    const float *delta;                 // (U)
    // pass A
    const float *zmax_in;               // (U) exact maximum so far (-inf at the start)
    uint32_t *vrun;                     // (U) running maximum of the raw accumulators, order-preserving encoding, shared by
                                        //     all CTAs (seeded from zmax_in): keeps the candidate lists short
    // pass B
    const float *zlo, *zhi;             // (U) bracket of the float16 cut-over (interp_thresholds)
    uint32_t *hitbits;                  // (B,U,W)
    // both: list of positions the resolve kernel re-evaluates
    uint2 *list;                        // {sequence, unit << 16 | position}
    unsigned int *count;                // entries appended (may exceed cap: overflow)
    unsigned int cap;
};

struct ItIter {
    int grp, chunk, nchunk;
    long long b, B, it, it1;
    __device__ ItIter(long long it0, long long it1_, long long B_, int nchunk_) : nchunk(nchunk_), B(B_), it(it0), it1(it1_)
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

__device__ __forceinline__ float fmax3i(float a, float b, float c)
{
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

__device__ __forceinline__ uint32_t ord_enc(float f)
{
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float ord_dec(uint32_t e) { return __uint_as_float((e & 0x80000000u) ? (e & 0x7FFFFFFFu) : ~e); }

// append one entry; lanes of a warp that arrive together share one atomic
__device__ __forceinline__ void it_append(const ItArgs &a, uint32_t n, uint32_t u, uint32_t s)
{
    const unsigned m = __activemask();
    const int leader = __ffs(m) - 1, lane = threadIdx.x & 31;
    unsigned base = 0;
    if (lane == leader) base = atomicAdd(a.count, (unsigned)__popc(m));
    base = __shfl_sync(m, base, leader);
    const unsigned i = base + __popc(m & ((1u << lane) - 1u));
    if (i < a.cap) a.list[i] = make_uint2(n, (u << 16) | s);
}

template <int MODE>
__global__ void __launch_bounds__(IT_THREADS + 64, 1) interp_tc_kernel(const ItArgs a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar_a[4], tmem_full[2], tmem_empty[2], v_full[IT_VBUFS], v_empty[IT_VBUFS];
    __shared__ uint32_t tmem_base_s;

    const int K = a.K, KC = K / 8, KS = K / 16;
    const uint32_t a_bytes = (uint32_t)IT_M * K * 2;
    unsigned char *sA = smem;
    uint4 *sV = reinterpret_cast<uint4 *>(sA + (size_t)a.gsz * 2 * a_bytes);
    float *run_s = reinterpret_cast<float *>(sV + IT_VBUFS * IT_VROWS);          // [gsz][512] running maxima (pass A)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n_mt = (a.U + IT_M - 1) / IT_M;
    const int n_grp = (n_mt + a.gsz - 1) / a.gsz;
    const int nchunk = (a.S + IT_N - 1) / IT_N;
    const long long total = a.B * nchunk * n_grp;
    const long long it0 = total * blockIdx.x / gridDim.x, it1 = total * (blockIdx.x + 1) / gridDim.x;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; ++i) mbar_init(&bar_a[i], 1);
        mbar_init(&tmem_full[0], 1);
        mbar_init(&tmem_full[1], 1);
        mbar_init(&tmem_empty[0], IT_THREADS / 64);
        mbar_init(&tmem_empty[1], IT_THREADS / 64);
        for (int i = 0; i < IT_VBUFS; ++i) { mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1); }
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

    const int role = tc_warp_uniform();
    if (role == IT_THREADS / 32) {
        // =========================== issue warp (as in conv_tc_kernel) ===========================
        uint32_t a_parity = 1;
        int cur_grp = -1;
        long long n = 0, vi = 0;
        for (ItIter t(it0, it1, a.B, nchunk); t.valid(); t.next(), ++vi) {
            const int mt0 = t.grp * a.gsz, mtn = min(a.gsz, n_mt - mt0);
            if (t.grp != cur_grp) {
                if (n > 0) mbar_wait(&tmem_full[(n - 1) & 1], (uint32_t)(((n - 1) >> 1) & 1));
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
            const int slot = (int)(vi & (IT_VBUFS - 1));
            const uint4 *V = sV + slot * IT_VROWS;
            const int ncol = (min(IT_N, a.S - t.chunk * IT_N) + 15) & ~15;
            const uint32_t idesc = (1u << 4) | ((uint32_t)(ncol >> 3) << 17) | ((uint32_t)(IT_M >> 4) << 24);
            mbar_wait(&v_full[slot], (uint32_t)((vi / IT_VBUFS) & 1));
            for (int m = 0; m < mtn; ++m, ++n) {
                const int buf = (int)(n & 1);
                const uint32_t use = (uint32_t)(n >> 1);
                mbar_wait(&bar_a[m], a_parity);
                mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);
                const uint32_t d = tmem_base + (uint32_t)buf * 256u;
                const uint32_t a_hi = smem_u32(sA) + (uint32_t)m * 2 * a_bytes;
                const uint64_t dA_hi = make_smem_desc(a_hi, 128, (uint32_t)KC * 128), dA_lo = make_smem_desc(a_hi + a_bytes, 128, (uint32_t)KC * 128);
                const uint64_t dV = make_smem_desc(smem_u32(V), 32, 128);
                if (tc_elect_one()) {
                    tc_fence_after();
                    for (int ks = 0; ks < KS; ++ks) {
                        tc_mma_f16(d, dA_hi + ks * 16, dV + ks * 4, idesc, ks > 0 ? 1u : 0u);
                        tc_mma_f16(d, dA_lo + ks * 16, dV + ks * 4, idesc, 1u);
                    }
                    tc_commit(&tmem_full[buf]);
                    if (m == mtn - 1) tc_commit(&v_empty[slot]);
                }
                __syncwarp();
            }
        }
    } else if (role == IT_THREADS / 32 + 1) {
        // =========================== builder warp (as in conv_tc_kernel, 256-position tiles) ===========================
        long long vi = 0;
        auto load_words = [&](const ItIter &t, uint32_t &cwv, uint32_t &mwv) {
            cwv = 0u; mwv = 0u;
            if (!t.valid()) return;
            const int n0 = t.chunk * IT_N;
            const int ci = (n0 >> 4) + lane, mi = (n0 >> 5) + lane;
            if (ci < a.wc) cwv = a.codes[t.b * a.wc + ci];
            if (mi < a.wm) mwv = a.nmask[t.b * a.wm + mi];
        };
        uint32_t cw0, mw0, cw1, mw1;
        ItIter tp(it0, it1, a.B, nchunk);
        load_words(tp, cw0, mw0); tp.next();
        load_words(tp, cw1, mw1); tp.next();
        for (ItIter t(it0, it1, a.B, nchunk); t.valid(); t.next(), ++vi) {
            const uint32_t cwr = cw0, mwr = mw0;
            cw0 = cw1; mw0 = mw1;
            load_words(tp, cw1, mw1); tp.next();
            const int n0 = t.chunk * IT_N;                  // multiple of 256: word aligned in both arrays
            const int slot = (int)(vi & (IT_VBUFS - 1));
            mbar_wait(&v_empty[slot], (uint32_t)(((vi / IT_VBUFS) & 1) ^ 1));
            uint4 *V = sV + slot * IT_VROWS;
            {
                const uint32_t c16 = cwr;                                                  // codes of my 16 positions
                const uint32_t c_nx = __shfl_down_sync(0xffffffffu, c16, 1) & 3u;
                const uint32_t mw_a = __shfl_sync(0xffffffffu, mwr, lane >> 1);
                uint32_t m16 = (mw_a >> (16 * (lane & 1))) & 0xFFFFu;
                const int nvalid_pos = a.L - (n0 + 16 * lane);
                if (nvalid_pos < 16) m16 |= nvalid_pos <= 0 ? 0xFFFFu : (0xFFFFu << nvalid_pos) & 0xFFFFu;
                const uint32_t m_nx = __shfl_down_sync(0xffffffffu, m16, 1) & 1u;
                uint32_t ox[17], oy[17];
#pragma unroll
                for (int j = 0; j < 17; ++j) {
                    const uint32_t c = j < 16 ? (c16 >> (2 * j)) & 3u : c_nx;
                    const uint32_t isn = j < 16 ? (m16 >> j) & 1u : m_nx;
                    const uint32_t one = isn ? 0u : (0x3C00u << ((c & 1u) * 16));
                    ox[j] = (c & 2u) ? 0u : one;
                    oy[j] = (c & 2u) ? one : 0u;
                }
                if (lane < IT_VROWS / 16) {
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
        // two groups of 8 warps, group g owns accumulator g (every other tile); inside a group lane quarter = warp & 3,
        // column half = (warp >> 2) & 1: one thread = one unit x 128 positions = 4 words of hit bits
        const int grp = warp >> 3;
        const int half = (warp >> 2) & 1;
        const int row = 32 * (warp & 3) + lane;
        const uint32_t taddr = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)grp * 256u + (uint32_t)half * 128u;
        long long n = 0;
        if (MODE == IT_MAX)
            for (int i = threadIdx.x; i < a.gsz * IT_THREADS; i += IT_THREADS) run_s[i] = -INFINITY;
        int cur_grp = -1;
        for (ItIter t(it0, it1, a.B, nchunk); t.valid(); t.next()) {
            const int mt0 = t.grp * a.gsz, mtn = min(a.gsz, n_mt - mt0);
            const int n0 = t.chunk * IT_N;
            const int cvalid = min(IT_N, a.S - n0) - half * 128;       // valid columns of my half (may be <= 0)
            if (MODE == IT_MAX && t.grp != cur_grp) {
                // new unit tiles: the running maxima restart from the exact maximum found so far
                for (int m = 0; m < mtn; ++m) {
                    const int u = (mt0 + m) * IT_M + row;
                    float r0 = -INFINITY;
                    if (u < a.U) {
                        const float zi = __ldg(a.zmax_in + u);
                        if (zi > -INFINITY) r0 = __fdiv_rn(__fsub_rn(zi, __ldg(a.shift_rs + u)), __ldg(a.shift_rs + a.U + u));
                    }
                    run_s[m * IT_THREADS + threadIdx.x] = r0;
                }
                cur_grp = t.grp;
            }
            for (int m = 0; m < mtn; ++m, ++n) {
                if ((int)(n & 1) != grp) continue;
                const uint32_t use = (uint32_t)(n >> 1);
                const int u = (mt0 + m) * IT_M + row;
                const bool uvalid = u < a.U;
                float sh = 0.f, rs = 1.f, dl = 0.f;
                if (uvalid) { sh = __ldg(a.shift_rs + u); rs = __ldg(a.shift_rs + a.U + u); dl = __ldg(a.delta + u); }
                float grun = -INFINITY;            // the unit's running maximum over ALL CTAs (may be stale: any lower bound works)
                if (MODE == IT_MAX && uvalid) grun = ord_dec(__ldcg(a.vrun + u));
                mbar_wait(&tmem_full[grp], use & 1);
                tc_fence_after();
                const bool active = (mt0 + m) * IT_M + 32 * (warp & 3) < a.U && cvalid > 0;      // warp-uniform
                if (active) {
                    if (MODE == IT_MAX) {
                        // ---- pass A: running maximum of the raw accumulators; near-maximal positions are listed ----
                        float run = fmaxf(run_s[m * IT_THREADS + threadIdx.x], grun);
                        const float run_in = run;
                        const float tol = __fdiv_rn(2.0f * dl, rs);            // 2 delta in accumulator units (rs > 0 in eval mode)
                        uint32_t rr[2][32];
                        tc_ld16_nowait(taddr, rr[0]);
                        tc_ld16_nowait(taddr + 16, rr[0] + 16);
#pragma unroll
                        for (int bi = 0; bi < 4; ++bi) {
                            uint32_t *r = rr[bi & 1];
                            tc_wait_ld32(r);
                            if (bi + 1 < 4) {
                                tc_ld16_nowait(taddr + 32 * (bi + 1), rr[(bi + 1) & 1]);
                                tc_ld16_nowait(taddr + 32 * (bi + 1) + 16, rr[(bi + 1) & 1] + 16);
                            }
                            const int cv = cvalid - 32 * bi;                  // valid columns of this batch
                            float v[32];
#pragma unroll
                            for (int i = 0; i < 32; ++i) v[i] = i < cv ? __uint_as_float(r[i]) : -INFINITY;
                            float mx = fmax3i(v[0], v[1], v[2]);
#pragma unroll
                            for (int i = 3; i + 1 < 32; i += 2) mx = fmax3i(mx, v[i], v[i + 1]);
                            mx = fmaxf(mx, v[31]);
                            if (uvalid && mx >= run - tol) {
#pragma unroll 1
                                for (int i = 0; i < 32; ++i) {
                                    float vi = -INFINITY;
#pragma unroll
                                    for (int q = 0; q < 32; ++q) vi = q == i ? v[q] : vi;     // (select chain: no dynamic register index)
                                    if (i < cv && vi >= run - tol) it_append(a, (uint32_t)t.b, (uint32_t)u, (uint32_t)(n0 + half * 128 + 32 * bi + i));
                                    run = fmaxf(run, vi);
                                }
                            }
                            run = fmaxf(run, mx);
                        }
                        run_s[m * IT_THREADS + threadIdx.x] = run;
                        if (uvalid && run > run_in) atomicMax(a.vrun + u, ord_enc(run));
                    } else {
                        // ---- pass B: sure / possible hit bits by comparing the raw accumulators with the scaled bracket ----
                        float hi_t = INFINITY, lo_t = INFINITY;
                        if (uvalid) {
                            const float zl = __ldg(a.zlo + u), zh = __ldg(a.zhi + u);
                            if (zh < INFINITY) hi_t = __fdiv_rn(__fsub_rn(__fadd_rn(zh, dl), sh), rs);
                            if (zl < INFINITY) lo_t = __fdiv_rn(__fsub_rn(__fsub_rn(zl, dl), sh), rs);
                        }
                        uint32_t sure[4], poss[4];
                        uint32_t rr[2][32];
                        tc_ld16_nowait(taddr, rr[0]);
                        tc_ld16_nowait(taddr + 16, rr[0] + 16);
#pragma unroll
                        for (int bi = 0; bi < 4; ++bi) {
                            uint32_t *r = rr[bi & 1];
                            tc_wait_ld32(r);
                            if (bi + 1 < 4) {
                                tc_ld16_nowait(taddr + 32 * (bi + 1), rr[(bi + 1) & 1]);
                                tc_ld16_nowait(taddr + 32 * (bi + 1) + 16, rr[(bi + 1) & 1] + 16);
                            }
                            uint32_t sb = 0u, pb = 0u;
#pragma unroll
                            for (int i = 31; i >= 0; --i) {                   // the sign of (threshold - v) is the hit bit
                                const float v = __uint_as_float(r[i]);
                                sb = __funnelshift_l(__float_as_uint(__fsub_rn(hi_t, v)), sb, 1);
                                pb = __funnelshift_l(__float_as_uint(__fsub_rn(lo_t, v)), pb, 1);
                            }
                            const int cv = cvalid - 32 * bi;
                            const uint32_t vm = cv >= 32 ? 0xFFFFFFFFu : (cv <= 0 ? 0u : ((1u << cv) - 1u));
                            sure[bi] = sb & vm;
                            poss[bi] = pb & vm;
                        }
                        if (uvalid) {
                            uint32_t *dst = a.hitbits + ((size_t)t.b * a.U + u) * a.W + (size_t)t.chunk * 8 + half * 4;
                            const int wleft = a.W - (t.chunk * 8 + half * 4);
#pragma unroll
                            for (int bi = 0; bi < 4; ++bi) {
                                if (bi < wleft) dst[bi] = sure[bi];
                                uint32_t amb = poss[bi] & ~sure[bi];
                                while (amb) {
                                    const int bit = __ffs(amb) - 1;
                                    amb &= amb - 1;
                                    it_append(a, (uint32_t)t.b, (uint32_t)u, (uint32_t)(n0 + half * 128 + 32 * bi + bit));
                                }
                            }
                        }
                    }
                } else if (MODE == IT_HITS && uvalid && cvalid <= 0) {
                    // my half lies past the last position, but its words may still belong to the row: they are zero
                    uint32_t *dst = a.hitbits + ((size_t)t.b * a.U + u) * a.W + (size_t)t.chunk * 8 + half * 4;
                    const int wleft = a.W - (t.chunk * 8 + half * 4);
                    for (int bi = 0; bi < 4 && bi < wleft; ++bi) dst[bi] = 0u;
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&tmem_empty[grp]);
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
    }
}

}  // namespace

namespace {

// canonical fp32 pre-activation of one (sequence, unit, position): the per-tap expression of conv.cu / the oracle
//   t_g = ((a_4g + a_4g+1) + a_4g+2) + a_4g+3,  z = (((t_0 + t_1) + ...) + t_{G-1}) + shift,  a_k = +0 for N columns
__device__ float canon_z(const uint32_t *__restrict__ cw, const uint32_t *__restrict__ mw, int F, const float *__restrict__ wq_tile,
                         int ul, float shift, int s)
{
    const int G = (F + 3) / 4;
    float v = 0.0f;
    for (int g = 0; g < G; ++g) {
        float tg = 0.0f;
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
            const int k = 4 * g + jj, pos = s + k;
            float w = 0.0f;
            if (k < F && !((mw[pos >> 5] >> (pos & 31)) & 1u)) {
                const int c = (cw[pos >> 4] >> (2 * (pos & 15))) & 3;
                w = wq_tile[(c * F + k) * XB_UT + ul];
            }
            tg = jj == 0 ? w : __fadd_rn(tg, w);
        }
        v = g == 0 ? tg : __fadd_rn(v, tg);
    }
    return __fadd_rn(v, shift);
}

__global__ void it_enc_kernel(const float *__restrict__ zmax, const float *__restrict__ shift_rs, int U, uint32_t *__restrict__ zenc,
                              uint32_t *__restrict__ vrun)
{
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= U) return;
    const float zi = zmax[u];
    zenc[u] = ord_enc(zi);
    // the same maximum in accumulator units (z = acc * rs + shift)
    vrun[u] = ord_enc(zi > -INFINITY ? __fdiv_rn(__fsub_rn(zi, shift_rs[u]), shift_rs[U + u]) : -INFINITY);
}

struct ResolveArgs {
    const uint32_t *codes, *nmask;
    int L, F, U, W, wc, wm;
    const float *wq_t, *shift;
    const uint2 *list;
    const unsigned int *count;
    unsigned int cap;
};

__global__ void __launch_bounds__(256) it_resolve_max_kernel(const ResolveArgs a, uint32_t *__restrict__ zenc)
{
    const unsigned n = min(*a.count, a.cap);
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint2 e = a.list[i];
        const int u = (int)(e.y >> 16), s = (int)(e.y & 0xFFFFu);
        const float z = canon_z(a.codes + (size_t)e.x * a.wc, a.nmask + (size_t)e.x * a.wm, a.F,
                                a.wq_t + (size_t)(u / XB_UT) * 4 * a.F * XB_UT, u % XB_UT, a.shift[u], s);
        atomicMax(&zenc[u], ord_enc(z));
    }
}

__global__ void it_finish_max_kernel(const uint32_t *__restrict__ zenc, int U, const unsigned int *__restrict__ count, unsigned int cap,
                                     float *__restrict__ zmax, int *__restrict__ status)
{
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u < U) zmax[u] = ord_dec(zenc[u]);
    if (u == 0 && *count > cap) atomicOr(status, 1);
}

__global__ void __launch_bounds__(256) it_resolve_hits_kernel(const ResolveArgs a, const float *__restrict__ zlo, const float *__restrict__ zhi,
                                                              const float *__restrict__ thr, uint32_t *__restrict__ hitbits,
                                                              int *__restrict__ status)
{
    if (blockIdx.x == 0 && threadIdx.x == 0 && *a.count > a.cap) atomicOr(status, 2);
    const unsigned n = min(*a.count, a.cap);
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint2 e = a.list[i];
        const int u = (int)(e.y >> 16), s = (int)(e.y & 0xFFFFu);
        const float z = canon_z(a.codes + (size_t)e.x * a.wc, a.nmask + (size_t)e.x * a.wm, a.F,
                                a.wq_t + (size_t)(u / XB_UT) * 4 * a.F * XB_UT, u % XB_UT, a.shift[u], s);
        bool h = z > zhi[u];
        if (!h && z > zlo[u]) h = fp16_round(det_expf(z)) > thr[u];
        if (h) atomicOr(&hitbits[((size_t)e.x * a.U + u) * a.W + (s >> 5)], 1u << (s & 31));
    }
}

// delta[u] = 2^-15 * (sum_k max_c |w'[u][c][k]| + |shift[u]|): bound on |z~ - z| (see the file header)
__global__ void it_bounds_kernel(const float *__restrict__ w, const float *__restrict__ scale, const float *__restrict__ shift, int U, int F,
                                 float *__restrict__ delta)
{
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= U) return;
    float A = 0.f;
    for (int k = 0; k < F; ++k) {
        float m = 0.f;
        for (int c = 0; c < 4; ++c) m = fmaxf(m, fabsf(__fmul_rn(w[((size_t)u * 4 + c) * F + k], scale[u])));
        A += m;
    }
    delta[u] = 3.0517578125e-05f * (A + fabsf(shift[u]));
}

int it_fill(ItArgs *a, const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F, const void *images,
            const float *shift_rs, const float *delta, size_t *smem_out, int *grid_out)
{
    if (F < 1 || F > 4 * XB_MAX_G || L < F || L > 2048) {
        xb_set_error("interp_tc: unsupported geometry L=%d filter_size=%d", L, F);
        return XB_ERR_UNSUPPORTED;
    }
    a->codes = codes; a->nmask = nmask; a->B = B; a->L = L; a->F = F; a->U = U; a->K = ((4 * F + 15) / 16) * 16;
    a->S = L - F + 1; a->W = (a->S + 31) / 32; a->wc = xb_words_codes(L); a->wm = xb_words_mask(L);
    a->images = (const __half *)images; a->shift_rs = shift_rs; a->delta = delta;
    const int n_mt = (U + IT_M - 1) / IT_M, nchunk = (a->S + IT_N - 1) / IT_N;
    const size_t tile_bytes = (size_t)2 * IT_M * a->K * 2 + (size_t)IT_THREADS * 4;      // weights + running maxima
    const size_t fixed = (size_t)IT_VBUFS * IT_VROWS * 16;
    const size_t budget = 227 * 1024 - 512;
    a->gsz = (int)std::min<size_t>(std::min<size_t>((budget - fixed) / tile_bytes, (size_t)n_mt), 4);
    *smem_out = fixed + (size_t)a->gsz * tile_bytes;
    const int n_grp = (n_mt + a->gsz - 1) / a->gsz;
    const long long total = (long long)n_grp * B * nchunk;
    int grid = xb_num_sms();
    if (total < grid) grid = (int)total;
    *grid_out = grid;
    return XB_OK;
}

}  // namespace

extern "C" int explainn_interp_tc_bounds(const float *w, const float *scale, const float *shift, int U, int F, float *delta,
                                         cudaStream_t stream)
{
    XB_REQUIRE(w && scale && shift && delta && U >= 1 && F >= 1, "interp_tc_bounds: bad argument");
    it_bounds_kernel<<<(U + 127) / 128, 128, 0, stream>>>(w, scale, shift, U, F, delta);
    return xb_check_launch("interp_tc_bounds");
}

extern "C" long long explainn_interp_tc_list_bytes(int U) { return 256 + (long long)((U + 63) / 64 * 64) * 8 + (8LL << 20) * 8; }

static void it_list_layout(void *list_scratch, long long list_bytes, int U, unsigned int **count, uint32_t **zenc, uint32_t **vrun,
                           uint2 **list, unsigned int *cap)
{
    unsigned char *base = static_cast<unsigned char *>(list_scratch);
    *count = reinterpret_cast<unsigned int *>(base);
    *zenc = reinterpret_cast<uint32_t *>(base + 256);
    *vrun = *zenc + (U + 63) / 64 * 64;
    const long long off = 256 + (long long)((U + 63) / 64 * 64) * 8;
    *list = reinterpret_cast<uint2 *>(base + off);
    const long long room = (list_bytes - off) / 8;
    *cap = room > 0 ? (unsigned int)std::min<long long>(room, 0x7FFFFFFFLL) : 0u;
}

extern "C" int explainn_interp_unit_max_tc(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                           const void *images, const float *shift_rs, const float *delta, const float *wq_t,
                                           const float *shift, float *zmax, void *list_scratch, long long list_bytes, int *status,
                                           cudaStream_t stream)
{
    XB_REQUIRE(codes && nmask && images && shift_rs && delta && wq_t && shift && zmax && list_scratch && status && B >= 0 && U >= 1,
               "interp_unit_max_tc: bad argument");
    XB_REQUIRE(list_bytes >= 256 + (long long)((U + 63) / 64 * 64) * 8 + 8 * 1024, "interp_unit_max_tc: list scratch too small");
    XB_REQUIRE(U <= 65535, "interp_unit_max_tc: at most 65535 units");
    if (B == 0) return XB_OK;
    ItArgs a{};
    size_t smem; int grid;
    int rc = it_fill(&a, codes, nmask, B, L, U, F, images, shift_rs, delta, &smem, &grid);
    if (rc != XB_OK) return rc;
    uint32_t *zenc;
    it_list_layout(list_scratch, list_bytes, U, &a.count, &zenc, &a.vrun, &a.list, &a.cap);
    a.zmax_in = zmax;
    cudaMemsetAsync(a.count, 0, 4, stream);
    it_enc_kernel<<<(U + 127) / 128, 128, 0, stream>>>(zmax, shift_rs, U, zenc, a.vrun);
    cudaError_t e = cudaFuncSetAttribute(interp_tc_kernel<IT_MAX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        xb_set_error("interp_unit_max_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    interp_tc_kernel<IT_MAX><<<grid, IT_THREADS + 64, smem, stream>>>(a);
    ResolveArgs r{codes, nmask, L, F, U, a.W, a.wc, a.wm, wq_t, shift, a.list, a.count, a.cap};
    it_resolve_max_kernel<<<2 * xb_num_sms(), 256, 0, stream>>>(r, zenc);
    it_finish_max_kernel<<<(U + 127) / 128, 128, 0, stream>>>(zenc, U, a.count, a.cap, zmax, status);
    return xb_check_launch("interp_unit_max_tc");
}

extern "C" int explainn_interp_pwm_count_tc(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                            const void *images, const float *shift_rs, const float *delta, const float *wq_t,
                                            const float *shift, const float *zlo, const float *zhi, const float *thr,
                                            unsigned long long *counts, unsigned long long *hits, unsigned char *seq_hit,
                                            void *scratch, void *list_scratch, long long list_bytes, int *status, cudaStream_t stream)
{
    XB_REQUIRE(codes && nmask && images && shift_rs && delta && wq_t && shift && zlo && zhi && thr && counts && hits && scratch &&
                   list_scratch && status && B >= 0 && U >= 1,
               "interp_pwm_count_tc: bad argument");
    XB_REQUIRE(list_bytes >= 256 + (long long)((U + 63) / 64 * 64) * 8 + 8 * 1024, "interp_pwm_count_tc: list scratch too small");
    XB_REQUIRE(U <= 65535, "interp_pwm_count_tc: at most 65535 units");
    if (B == 0) return XB_OK;
    ItArgs a{};
    size_t smem; int grid;
    int rc = it_fill(&a, codes, nmask, B, L, U, F, images, shift_rs, delta, &smem, &grid);
    if (rc != XB_OK) return rc;
    uint32_t *zenc;
    it_list_layout(list_scratch, list_bytes, U, &a.count, &zenc, &a.vrun, &a.list, &a.cap);
    long long off_bmask, off_bits;
    xb_pwm_count_layout(B, L, U, F, &off_bmask, &off_bits);
    unsigned char *base = static_cast<unsigned char *>(scratch);
    uint32_t *bmask = reinterpret_cast<uint32_t *>(base + off_bmask);
    a.hitbits = reinterpret_cast<uint32_t *>(base + off_bits);
    a.zlo = zlo; a.zhi = zhi;
    cudaMemsetAsync(a.count, 0, 4, stream);
    cudaError_t e = cudaFuncSetAttribute(interp_tc_kernel<IT_HITS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        xb_set_error("interp_pwm_count_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    interp_tc_kernel<IT_HITS><<<grid, IT_THREADS + 64, smem, stream>>>(a);
    ResolveArgs r{codes, nmask, L, F, U, a.W, a.wc, a.wm, wq_t, shift, a.list, a.count, a.cap};
    it_resolve_hits_kernel<<<2 * xb_num_sms(), 256, 0, stream>>>(r, zlo, zhi, thr, a.hitbits, status);
    return xb_hit_count(codes, nmask, B, L, U, F, a.hitbits, bmask, nullptr, seq_hit, counts, hits, stream);
}
