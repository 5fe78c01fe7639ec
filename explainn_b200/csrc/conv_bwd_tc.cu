// conv_bwd_tc.cu -- sparse part of the conv / BatchNorm-1 backward on the tensor cores (default MaxPool1d(7, 7)).
//
// What it replaces in the reference (paths relative to /root/reference):
//   loss.backward() through MaxPool1d -> ExpActivation -> BatchNorm1d -> grouped Conv1d   explainn/train/train.py:65
//   (modules of explainn/models/networks.py:98-102)
//
// Only the argmax position of each pool window carries gradient: with gz[b,u,p] = d(pooled) * pooled and ix[b,u,p] the
// offset of the window maximum,
//     A[u][c*F + k] = sum_{b,p} gz[b,u,p] * X[b, c, 7p + ix[b,u,p] + k]          (the `partial` of explainn_conv_pool_bwd)
// conv_bwd_sparse_kernel (train.cu) scatters that sum through shared memory, 19 read-modify-writes per entry, and is
// bound by instruction issue.  Here it is a GEMM per sequence,
//     D[u][n] += sum_{p, j<8} G[u][(p,j)] * Xw[n][(p,j)],    G[u][(p,j)] = gz[b,u,p] if j == ix[b,u,p] else 0,
//                                                             Xw[c*F+k][(p,j)] = X[b, c, 7p + j + k],
// M = 128 units, N = roundup16(4F), K = 8 per pool window (7 offsets + one zero).  One K block of G is a single non-zero
// among 8 bfloat16 = 16 bytes = four 32-bit tensor-memory columns, so the G operand never touches shared memory: each
// thread writes the rows of ITS unit straight into tensor memory (tcgen05.st) and the MMA takes A from there.  gz is
// split into three bfloat16 pieces (24 significant bits: fp32-exact products, fp32 accumulation in tensor memory).
// Xw is expanded once per sequence from the packed codes (0 / 1 are exact in bfloat16).
//
// One CTA per SM; CTA = (128-unit tile, slice): the CTAs of a tile split the batch into contiguous ranges of sequences,
// the accumulator of the tile stays in tensor memory for the whole range and is drained once, through shared memory
// and coalesced stores, into rows [tile] of slice `blockIdx / n_tiles` of `partial` (deterministic: no atomics).
//   warps 0..15   G builders (lane quarter = warp & 3, K step inside the chunk = warp >> 2), later the drain;
//                 a chunk = 4 K steps = 96 tensor-memory columns, four chunks in flight (a_full / a_empty)
//   warp  16      MMA issue: 12 tcgen05.mma (3 pieces x 4 K steps) per chunk, A from tensor memory, B = Xw descriptor
//   warp  17      loader: d(pooled), pooled, x_hat blocks of one (sequence, tile) by TMA bulk copy (each one contiguous
//                 block of the (B,U,P) tensors), argmax bytes by 4-byte cp.async; two staging slots
//   warps 18..21  Xw builders: code words prefetched four sequences ahead (cp.async ring), indicator bit masks per
//                 base, each distinct 16-byte row expanded once through a 128-entry table; two operand buffers
// Measured on B200 (tools/time_conv_bwd.py, cfg 2, batch 512): 33 us against 64 us of the shared-memory scatter; the
// kernel reads 66 MB (16 us at the DRAM rate the loader alone reaches); a tcgen05.mma with A in tensor memory costs
// ~64 clk for any N <= 128, so the 51 MMAs per (sequence, tile) set the floor at 25 us.
#include "tc_common.cuh"
#include "../../include/explainn_b200.h"
#include <algorithm>
#include <stdlib.h>

namespace {

// ablation switches of the experiment build (-DXB_EXPERIMENT, env XB_CB_DBG): 1 no MMAs, 2 no G arithmetic / stores,
// 4 no Xw rows, 8 no output copy, 16 no argmax-byte copies, 32 no bulk copies, 64 no chunk hand-shakes
#ifdef XB_EXPERIMENT
#define CB_DBG(a, bit) ((a).dbg & (bit))
#else
#define CB_DBG(a, bit) 0
#endif

constexpr int CB_M = 128;
constexpr int CB_AWARPS = 16;
constexpr int CB_BWARPS = 4;
constexpr int CB_THREADS = (CB_AWARPS + CB_BWARPS + 2) * 32;      // 704
constexpr int CB_KSC = 4;                        // K steps (two pool windows each) per G buffer
constexpr int CB_ACOLS = 3 * CB_KSC * 8;         // tensor-memory columns of one G buffer: 3 pieces x 4 K steps x 8
constexpr int CB_NABUF = 4;                      // G buffers in flight
constexpr int CB_XJ = 3;                         // distinct Xw rows per builder thread and base: T <= 128 * CB_XJ
constexpr int CB_XR = 4;                         // windows one distinct row serves: ceil(F / 7)
constexpr int CB_XRING = 4;                      // sequences of code words in flight for the Xw builders
constexpr int CB_ACC_COLS = 128;                 // accumulator: N <= 96 columns

struct CbArgs {
    const uint32_t *codes, *nmask;
    const float *gpool, *pooled, *xhat;
    const unsigned char *idx;
    float *partial;
    long long B;
    int L, F, U, P, W, N, wc, wm, cpt, P2, mwa, dbg;
    uint32_t plane, stage_bytes, b_bytes, b_alloc, s_bytes;
};

__device__ __forceinline__ uint32_t cb_even_bits(uint32_t x)
{
    x &= 0x55555555u;
    x = (x | (x >> 1)) & 0x33333333u;
    x = (x | (x >> 2)) & 0x0F0F0F0Fu;
    x = (x | (x >> 4)) & 0x00FF00FFu;
    x = (x | (x >> 8)) & 0x0000FFFFu;
    return x;
}

__device__ __forceinline__ void cb_mma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
        "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}

// the same with the accumulate predicate constant true (no register, no setp)
__device__ __forceinline__ void cb_mma_ts_acc(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc)
{
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.eq.u32 p, 1, 1;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
        "r"(tmem_a), "l"(bdesc), "r"(idesc)
        : "memory");
}

__device__ __forceinline__ void cb_st8(uint32_t taddr, const uint32_t (&w)[8])
{
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(w[0]), "r"(w[1]), "r"(w[2]),
                 "r"(w[3]), "r"(w[4]), "r"(w[5]), "r"(w[6]), "r"(w[7])
                 : "memory");
}

__device__ __forceinline__ void cb_ld4(uint32_t taddr, float (&v)[4])
{
    uint32_t r[4];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = __uint_as_float(r[i]);
}

__device__ __forceinline__ void cb_bar(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

__device__ __forceinline__ uint32_t cb_shl(uint32_t x, uint32_t s)          // PTX shl: amounts >= 32 give 0
{
    uint32_t r;
    asm("shl.b32 %0, %1, %2;" : "=r"(r) : "r"(x), "r"(s));
    return r;
}

__global__ void __launch_bounds__(CB_THREADS, 1) conv_bwd_tc_kernel(const CbArgs a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t s_full[2], s_empty[2], b_full[2], b_empty[2], a_full[CB_NABUF], a_empty[CB_NABUF], acc_full;
    __shared__ uint32_t tmem_base_s;
    __shared__ float red[4][CB_M][2];
    __shared__ uint4 lut[128];                                       // 7 indicator bits -> 8 bfloat16 (the eighth is zero)

    unsigned char *sB = smem;                                        // two Xw operands
    unsigned char *sS = sB + 2 * (size_t)a.b_alloc;                  // two staging slots (d(pooled), pooled, x_hat planes + argmax
                                                                     // bytes); the drain stages its output tile here
    uint32_t *sMask = reinterpret_cast<uint32_t *>(sS + a.s_bytes);  // [4][mwa] indicator bits of one sequence
    uint32_t *sRing = sMask + 4 * a.mwa;                             // [CB_XRING][128][3] code / mask words on their way in

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // CTA = (unit tile, slice): tile = blockIdx % n_mt, slice j = blockIdx / n_mt handles a contiguous range of sequences
    const int n_mt = (a.U + CB_M - 1) / CB_M;
    const int tile_m = blockIdx.x % n_mt, slice = blockIdx.x / n_mt;
    const int b0 = (int)(a.B * slice / a.cpt), b1 = (int)(a.B * (slice + 1) / a.cpt);
    const int u0 = tile_m * CB_M, rows = min(CB_M, a.U - u0);
    const int KS = a.P2 / 2;                                         // K steps of 16 (two windows) per sequence
    const int nch = (KS + CB_KSC - 1) / CB_KSC;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&s_full[i], 1 + 32);                           // expect_tx arrival + the loader's 32 lanes
            mbar_init(&s_empty[i], CB_AWARPS);
            mbar_init(&b_full[i], CB_BWARPS);
            mbar_init(&b_empty[i], 1);
        }
        for (int i = 0; i < CB_NABUF; ++i) {
            mbar_init(&a_full[i], CB_AWARPS);
            mbar_init(&a_empty[i], 1);
        }
        mbar_init(&acc_full, 1);
        mbar_fence_init();
    }
    if (threadIdx.x < 128) {
        const uint32_t t = threadIdx.x;
        uint4 o;
        o.x = ((t & 1u) ? 0x3F80u : 0u) | ((t & 2u) ? 0x3F800000u : 0u);
        o.y = ((t & 4u) ? 0x3F80u : 0u) | ((t & 8u) ? 0x3F800000u : 0u);
        o.z = ((t & 16u) ? 0x3F80u : 0u) | ((t & 32u) ? 0x3F800000u : 0u);
        o.w = (t & 64u) ? 0x3F80u : 0u;
        lut[t] = o;
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t tmem_a0 = tmem_base + CB_ACC_COLS;                // G buffers behind the accumulator

    const int role = tc_warp_uniform();
    if (role == CB_AWARPS) {
        // =========================== issue warp ===========================
        const uint32_t idesc = tc_idesc(1u, CB_M, a.N);              // bf16 x bf16 -> f32
        uint32_t ai = 0;
        for (int b = b0; b < b1; ++b) {
            const uint32_t bi = (uint32_t)(b - b0);
            const int bb = (int)(bi & 1);
            mbar_wait(&b_full[bb], (bi >> 1) & 1);
            const uint64_t dB = make_smem_desc(smem_u32(sB + (size_t)bb * a.b_alloc), 128, (uint32_t)a.P2 * 128);
            for (int ch = 0; ch < (CB_DBG(a, 64) ? 0 : nch); ++ch, ++ai) {
                const uint32_t buf = ai % CB_NABUF;
                mbar_wait(&a_full[buf], (ai / CB_NABUF) & 1);
                const int nks = min(CB_KSC, KS - ch * CB_KSC);
                const uint32_t ta = tmem_a0 + buf * CB_ACOLS;
                const uint64_t dBc = dB + (uint64_t)(ch * CB_KSC * 16);               // a K step = two core matrices = 256 bytes
                const uint32_t first = (ch == 0 && b == b0) ? 0u : 1u;
                if (tc_elect_one()) {
                    tc_fence_after();
                    if (CB_DBG(a, 1)) {
                    } else if (nks == CB_KSC) {
                        // the common chunk, fully unrolled: 12 MMAs whose operands differ by compile-time offsets
                        cb_mma_ts(tmem_base, ta, dBc, idesc, first);
#pragma unroll
                        for (int i = 1; i < 3 * CB_KSC; ++i)
                            cb_mma_ts_acc(tmem_base, ta + (uint32_t)((i / CB_KSC) * CB_KSC * 8 + (i % CB_KSC) * 8), dBc + (uint64_t)((i % CB_KSC) * 16),
                                          idesc);
                    } else {
                        for (int pc = 0; pc < 3; ++pc)
                            for (int j = 0; j < nks; ++j)
                                cb_mma_ts(tmem_base, ta + (uint32_t)(pc * CB_KSC * 8 + j * 8), dBc + (uint64_t)(j * 16), idesc,
                                          (pc == 0 && j == 0) ? first : 1u);
                    }
                    tc_commit(&a_empty[buf]);
                }
                __syncwarp();
            }
            if (tc_elect_one()) tc_commit(&b_empty[bb]);             // every MMA that reads this Xw buffer has retired
            __syncwarp();
        }
        if (tc_elect_one()) tc_commit(&acc_full);
        __syncwarp();
    } else if (role == CB_AWARPS + 1) {
        // =========================== loader warp ===========================
        const int n = rows * a.P;
        for (int b = b0; b < b1; ++b) {
            const uint32_t si = (uint32_t)(b - b0);
            const int slot = (int)(si & 1);
            mbar_wait(&s_empty[slot], ((si >> 1) & 1) ^ 1);
            const size_t off = ((size_t)b * a.U + u0) * a.P;
            unsigned char *st = sS + (size_t)slot * a.stage_bytes;
            unsigned char *dsti = st + 3 * (size_t)a.plane;
            const unsigned char *srci = a.idx + off;
            // argmax bytes: asynchronous 4-byte copies, all in flight at once; each lane's arrival on s_full fires when
            // ITS copies have landed (cp.async.mbarrier.arrive.noinc), so the loader never waits for global memory
            if (((reinterpret_cast<uintptr_t>(srci) | (uintptr_t)n) & 3u) == 0) {
                const uint32_t d0 = smem_u32(dsti);
                for (int i = lane; i < (CB_DBG(a, 16) ? 0 : n / 4); i += 32)
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d0 + 4u * i), "l"(srci + 4 * (size_t)i) : "memory");
            } else {
                for (int i = lane; i < n; i += 32) dsti[i] = __ldg(srci + i);
                __threadfence_block();                               // plain stores: ordered before this lane's arrival
            }
            asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(&s_full[slot])) : "memory");
            if (lane == 0) {
                const uint32_t bytes = (uint32_t)n * 4u;
                if (CB_DBG(a, 32)) {
                    mbar_arrive(&s_full[slot]);
                } else {
                    mbar_arrive_expect_tx(&s_full[slot], 3 * bytes);
                    tma_bulk_g2s(st, a.gpool + off, bytes, &s_full[slot]);
                    tma_bulk_g2s(st + a.plane, a.pooled + off, bytes, &s_full[slot]);
                    tma_bulk_g2s(st + 2 * (size_t)a.plane, a.xhat + off, bytes, &s_full[slot]);
                }
            }
        }
    } else if (role >= CB_AWARPS + 2) {
        // =========================== Xw builder warps ===========================
        // thread = (row inside an 8-row group, window offset 0..15); a warp stores 512 contiguous bytes
        const int tid = threadIdx.x - (CB_AWARPS + 2) * 32;          // 0..127
        // Row (c, k) of window p holds the indicator bits of base c at positions 7p + k .. 7p + k + 6: it depends on
        // t = 7p + k only, so each of the 4 T distinct rows is expanded once and stored to every (k, p) it serves (at most
        // CB_XR of them: k = t % 7 + 7r).  The store offsets do not depend on the sequence: computed once, two per register.
        const int T = XB_POOL * (a.P2 - 1) + a.F;
        const uint32_t dummy = a.b_bytes / 16;                        // a spare row behind the operand takes the unused stores
        uint32_t offs[CB_XJ][4][CB_XR / 2];
#pragma unroll
        for (int j = 0; j < CB_XJ; ++j) {
            const int t = tid + CB_BWARPS * 32 * j;
#pragma unroll
            for (int c = 0; c < 4; ++c)
#pragma unroll
                for (int r = 0; r < CB_XR; ++r) {
                    const int p = t / XB_POOL - r, k = t - XB_POOL * p;
                    const int n = c * a.F + k;
                    const uint32_t o = (t < T && p >= 0 && p < a.P2 && k < a.F) ? (uint32_t)(((n >> 3) * a.P2 + p) * 8 + (n & 7)) : dummy;
                    if (r & 1) offs[j][c][r / 2] |= o << 16;
                    else offs[j][c][r / 2] = o;
                }
        }
        // rows 4F .. N-1 of both operands are zero for good
        for (int i = tid; i < 2 * (a.N - a.W) * a.P2; i += CB_BWARPS * 32) {
            const int bb = i / ((a.N - a.W) * a.P2), rem = i - bb * (a.N - a.W) * a.P2;
            const int n = a.W + rem / a.P2, pp = rem % a.P2;
            reinterpret_cast<uint4 *>(sB + (size_t)bb * a.b_alloc)[((n >> 3) * a.P2 + pp) * 8 + (n & 7)] = make_uint4(0u, 0u, 0u, 0u);
        }
        // code / mask words of a sequence: one mask word (32 positions) per thread, copied asynchronously into a ring
        // CB_XRING sequences ahead (each thread reads back only what it copied itself: no barrier needed)
        uint32_t *ring = sRing + tid * 3;
        auto fetch = [&](int b) {
            if (b < b1 && tid < a.mwa) {
                const uint32_t d = smem_u32(ring + (size_t)((b - b0) % CB_XRING) * (CB_BWARPS * 32 * 3));
                if (2 * tid < a.wc)
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(a.codes + (size_t)b * a.wc + 2 * tid) : "memory");
                if (2 * tid + 1 < a.wc)
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 4), "l"(a.codes + (size_t)b * a.wc + 2 * tid + 1) : "memory");
                if (tid < a.wm)
                    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + 8), "l"(a.nmask + (size_t)b * a.wm + tid) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        for (int i = 0; i < CB_XRING - 1; ++i) fetch(b0 + i);
        for (int b = b0; b < b1; ++b) {
            const uint32_t bi = (uint32_t)(b - b0);
            const int bb = (int)(bi & 1);
            fetch(b + CB_XRING - 1);
            asm volatile("cp.async.wait_group %0;" ::"n"(CB_XRING - 1) : "memory");      // the group of sequence b has landed
            uint32_t c0 = 0u, c1 = 0u, nm = 0xFFFFFFFFu;
            if (tid < a.mwa) {
                const uint32_t *r = ring + (size_t)(bi % CB_XRING) * (CB_BWARPS * 32 * 3);
                if (2 * tid < a.wc) c0 = r[0];
                if (2 * tid + 1 < a.wc) c1 = r[1];
                if (tid < a.wm) nm = r[2];
            }
            cb_bar(1, CB_BWARPS * 32);                               // the previous sequence's rows have been built
            if (tid < a.mwa) {
                const int w = tid;
                uint32_t valid = ~nm;
                const int left = a.L - 32 * w;
                if (left < 32) valid &= left > 0 ? (1u << left) - 1u : 0u;
                const uint32_t lo = cb_even_bits(c0) | (cb_even_bits(c1) << 16);
                const uint32_t hi = cb_even_bits(c0 >> 1) | (cb_even_bits(c1 >> 1) << 16);
                sMask[0 * a.mwa + w] = ~hi & ~lo & valid;
                sMask[1 * a.mwa + w] = ~hi & lo & valid;
                sMask[2 * a.mwa + w] = hi & ~lo & valid;
                sMask[3 * a.mwa + w] = hi & lo & valid;
            }
            cb_bar(1, CB_BWARPS * 32);
            mbar_wait(&b_empty[bb], ((bi >> 1) & 1) ^ 1);
            uint4 *dst = reinterpret_cast<uint4 *>(sB + (size_t)bb * a.b_alloc);
            if (!CB_DBG(a, 4)) {
#pragma unroll
                for (int j = 0; j < CB_XJ; ++j) {
                    const int t = min(tid + CB_BWARPS * 32 * j, T - 1);              // lanes past T repeat the last row into `dummy`
                    uint4 rowv[4];
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        const uint32_t *mk = sMask + c * a.mwa + (t >> 5);
                        rowv[c] = lut[__funnelshift_r(mk[0], mk[1], t & 31) & 0x7Fu];   // offsets 0..6; the eighth K element is zero
                    }
#pragma unroll
                    for (int c = 0; c < 4; ++c)
#pragma unroll
                        for (int r = 0; r < CB_XR; ++r) {
                            const uint32_t o = (r & 1) ? offs[j][c][r / 2] >> 16 : offs[j][c][r / 2] & 0xFFFFu;
                            dst[o] = rowv[c];
                        }
                }
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&b_full[bb]);
        }
    } else {
        // =========================== G builder warps, then drain ===========================
        const int q = warp & 3, sub = warp >> 2, row = 32 * q + lane;
        const uint32_t lane_base = (uint32_t)(32 * q) << 16;
        const int P = a.P, W = a.W;
        const uint32_t my_a = tmem_a0 + lane_base + (uint32_t)(sub * 8);
        const uint32_t pl4 = a.plane / 4;
        uint32_t ai = 0;
        float dg = 0.0f, db = 0.0f;                                  // this thread's share of d(gamma), d(beta) of its unit
        for (int b = b0; b < b1; ++b) {
            const uint32_t si = (uint32_t)(b - b0);
            const int slot = (int)(si & 1);
            mbar_wait(&s_full[slot], (si >> 1) & 1);
            const unsigned char *st = sS + (size_t)slot * a.stage_bytes;
            const float *sg = reinterpret_cast<const float *>(st) + row * P;
            const unsigned char *sidx = st + 3 * (size_t)a.plane + row * P;
            for (int ch = 0; ch < (CB_DBG(a, 64) ? 0 : nch); ++ch, ++ai) {
                const uint32_t buf = ai % CB_NABUF;
                const int ks = ch * CB_KSC + sub;                    // my K step: pool windows 2 ks, 2 ks + 1
                const bool mine = ks < KS && !CB_DBG(a, 2);
                uint32_t w[3][8];
                if (mine) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int p = 2 * ks + h;
                        // p == P (the pad window of an odd P) reads the next row's first entry: finite or not, it is dropped
                        float gz = sg[p] * sg[pl4 + p];
                        float xh = sg[2 * pl4 + p];
                        const uint32_t ix = sidx[p];
                        gz = p < P ? gz : 0.0f;
                        xh = p < P ? xh : 0.0f;
                        db += gz;
                        dg = fmaf(gz, xh, dg);
                        // three bfloat16 pieces by truncation: the two remainders are exact in fp32
                        const uint32_t hb = __float_as_uint(gz) & 0xFFFF0000u;
                        const float r1 = gz - __uint_as_float(hb);
                        const uint32_t mb = __float_as_uint(r1) & 0xFFFF0000u;
                        const float r2 = r1 - __uint_as_float(mb);
                        const uint32_t pcs[3] = {hb >> 16, mb >> 16, __float_as_uint(r2) >> 16};
                        // element ix of the window = bits [16 ix, 16 ix + 16) of its four words: word i = piece << (16 ix - 32 i),
                        // zero whenever that amount falls outside [0, 32) (unsigned wrap + PTX shift clamp)
                        const uint32_t s0 = ix << 4;
#pragma unroll
                        for (int pc = 0; pc < 3; ++pc)
#pragma unroll
                            for (int i = 0; i < 4; ++i) w[pc][4 * h + i] = cb_shl(pcs[pc], s0 - 32u * i);
                    }
                }
                // hand over the PREVIOUS chunk: its tcgen05.st had the time of this chunk's arithmetic to complete
                if (ch > 0) {
                    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&a_full[(ai - 1) % CB_NABUF]);
                }
                mbar_wait(&a_empty[buf], ((ai / CB_NABUF) & 1) ^ 1); // the MMAs that read this buffer have retired
                tc_fence_after();
                if (mine) {
                    const uint32_t ta = my_a + buf * CB_ACOLS;
#pragma unroll
                    for (int pc = 0; pc < 3; ++pc) cb_st8(ta + (uint32_t)(pc * CB_KSC * 8), w[pc]);
                }
            }
            if (!CB_DBG(a, 64)) {
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&a_full[(ai - 1) % CB_NABUF]);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&s_empty[slot]);              // staging slot read
        }
        // ---- drain: accumulator -> staging tile (the layout of `partial`) -> this CTA's rows of its slice, coalesced ----
        mbar_wait(&acc_full, 0u);
        tc_fence_after();
        const int ncq = a.N / 4;                                     // accumulator columns per warp row `sub`
        float *tile = reinterpret_cast<float *>(sS);
        for (int c4 = 0; c4 < ncq; c4 += 4) {
            float v[4];
            const int col = sub * ncq + c4;
            cb_ld4(tmem_base + lane_base + (uint32_t)col, v);
#pragma unroll
            for (int i = 0; i < 4; ++i)
                if (col + i < W) tile[row * (W + 2) + col + i] = v[i];
        }
        red[sub][row][0] = dg;
        red[sub][row][1] = db;
        cb_bar(2, CB_AWARPS * 32);
        if (sub == 0) {
            tile[row * (W + 2) + W] = ((red[0][row][0] + red[1][row][0]) + red[2][row][0]) + red[3][row][0];
            tile[row * (W + 2) + W + 1] = ((red[0][row][1] + red[1][row][1]) + red[2][row][1]) + red[3][row][1];
        }
        cb_bar(2, CB_AWARPS * 32);
        float *out = a.partial + ((size_t)slice * a.U + u0) * (W + 2);
        const int nout = CB_DBG(a, 8) ? 0 : rows * (W + 2);
        for (int i = threadIdx.x; i < nout; i += CB_AWARPS * 32) out[i] = tile[i];
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tc_dealloc(tmem_base, 512u);
}

}  // namespace

// Number of `partial` slices the tensor-core kernel writes for this problem (= its grid), 0 when the shape / alignment
// is outside what it handles and conv_bwd_sparse_kernel runs instead.
struct CbPlan {
    int grid, cpt, N, P, P2, mwa;
    uint32_t plane, stage_bytes, b_bytes, b_alloc, s_bytes;
    size_t smem;
};

static bool cb_plan(long long B, int L, int U, int F, int nsplit, const void *gpool, const void *pooled, const void *xhat, CbPlan &pl)
{
    const int S = L - F + 1;
    if (S < XB_POOL || B < 1) return false;
    pl.P = (S - XB_POOL) / XB_POOL + 1;
    pl.P2 = (pl.P + 1) & ~1;
    pl.N = (4 * F + 15) & ~15;
    if (pl.N > CB_ACC_COLS) return false;
    if (((long long)U * pl.P) % 4 != 0) return false;                // 16-byte aligned blocks for the bulk copies
    if ((reinterpret_cast<uintptr_t>(gpool) | reinterpret_cast<uintptr_t>(pooled) | reinterpret_cast<uintptr_t>(xhat)) & 15u) return false;
    const int n_mt = (U + CB_M - 1) / CB_M;
    if (n_mt > xb_num_sms()) return false;
    pl.cpt = (int)std::min<long long>(xb_num_sms() / n_mt, B);      // CTAs (= slices of `partial`) per unit tile
    pl.grid = pl.cpt * n_mt;
    pl.plane = (uint32_t)CB_M * pl.P * 4;
    pl.stage_bytes = 3 * pl.plane + (((uint32_t)CB_M * pl.P + 15u) & ~15u);
    pl.b_bytes = (uint32_t)(pl.N / 8) * pl.P2 * 128;
    pl.b_alloc = pl.b_bytes + 128;                                   // + the spare row of the Xw builders
    if (XB_POOL * (pl.P2 - 1) + F > CB_BWARPS * 32 * CB_XJ || (F + XB_POOL - 1) / XB_POOL > CB_XR || pl.b_bytes / 16 >= 65535) return false;
    pl.mwa = ((XB_POOL * (pl.P2 - 1) + F - 1) >> 5) + 2;
    if (pl.mwa > CB_BWARPS * 32) return false;                       // one mask word per Xw builder thread
    pl.s_bytes = std::max(2 * pl.stage_bytes, (uint32_t)CB_M * (4 * F + 2) * 4u);     // staging slots; the drain's output tile
    pl.smem = 2 * (size_t)pl.b_alloc + pl.s_bytes + 4 * (size_t)pl.mwa * 4 + (size_t)CB_XRING * CB_BWARPS * 32 * 3 * 4;
    int dev = 0, optin = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    if (pl.smem + 8 * 1024 > (size_t)optin) return false;            // static shared memory: barriers, table, reduction scratch
    return nsplit >= pl.cpt;
}

int xb_conv_bwd_tc_slices(long long B, int L, int U, int F, int nsplit, const void *gpool, const void *pooled, const void *xhat)
{
    CbPlan pl;
    return cb_plan(B, L, U, F, nsplit, gpool, pooled, xhat, pl) ? pl.cpt : 0;
}

int xb_conv_bwd_tc(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F, const float *gpool, const float *pooled,
                   const float *xhat, const unsigned char *idx, float *partial, int nsplit, cudaStream_t stream)
{
    CbPlan pl;
    if (!cb_plan(B, L, U, F, nsplit, gpool, pooled, xhat, pl)) return XB_ERR_UNSUPPORTED;
    CbArgs a{};
    a.codes = codes; a.nmask = nmask; a.gpool = gpool; a.pooled = pooled; a.xhat = xhat; a.idx = idx; a.partial = partial;
    a.B = B; a.L = L; a.F = F; a.U = U; a.P = pl.P; a.W = 4 * F; a.N = pl.N; a.wc = xb_words_codes(L); a.wm = xb_words_mask(L);
    a.cpt = pl.cpt; a.P2 = pl.P2; a.mwa = pl.mwa; a.plane = pl.plane; a.stage_bytes = pl.stage_bytes; a.b_bytes = pl.b_bytes; a.b_alloc = pl.b_alloc; a.s_bytes = pl.s_bytes;
#ifdef XB_EXPERIMENT
    if (const char *d = getenv("XB_CB_DBG")) a.dbg = atoi(d);
#endif
    cudaError_t e = cudaFuncSetAttribute(conv_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem);
    if (e != cudaSuccess) {
        xb_set_error("conv_bwd_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    conv_bwd_tc_kernel<<<pl.grid, CB_THREADS, pl.smem, stream>>>(a);
    return xb_check_launch("conv_bwd_tc");
}
