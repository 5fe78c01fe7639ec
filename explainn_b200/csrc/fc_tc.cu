// fc_tc.cu -- per-unit FC1 (grouped 1x1 Conv1d(U*P -> 100U), explainn/models/networks.py:105-107) on the tensor
// cores (tcgen05.mma.kind::tf32, fp32 accumulate in TMEM), with fp32-level accuracy through the 3-term split
//   x = x_hi + x_lo (x_hi = x truncated to tf32),   A.B ~= A_hi.B_hi + A_lo.B_hi + A_hi.B_lo      (error ~2^-21)
//
// MODE_TRAIN:  D[j][b] = sum_p W1[u][j][p] * pooled[b][u][p]     M = 128 (hidden channel j, 100 valid), N = 128 (b)
//              epilogue thread = one hidden channel: + bias, per-channel sum / sum of squares over its columns
//              (BatchNorm-2 batch statistics), h1[u][b][j] stored with lanes along j (coalesced)
// MODE_EVAL:   D[b][j] = same product transposed                  M = 128 (b), N = 112 (j)
//              epilogue thread = one sequence: BN-2 (running) -> ReLU -> dot with W2 -> BN-3 -> ReLU -> unit_out[b][u]
//              (replaces networks.py:105-116 in eval mode; the hidden activations never leave the SM)
// Both modes use the same two shared-memory tiles (W1[u] and the pooled tile, K-major, no-swizzle core-matrix
// layout); only which one is the A operand changes.
// fc1_tc_eval_pipe_kernel / fc1_tc_train_pipe_kernel are the software-pipelined forms used when P <= 40 (one K chunk);
// linear_tc_kernel is the final Linear for inference; bwd_fc1_tc_kernel the FC1 backward.  All MMAs are issued from one
// elected lane of a warp-uniform branch with precomputed descriptors (tc_common.cuh: tc_elect_one).
#include "tc_common.cuh"
#include "../../include/explainn_b200.h"
#include <stdlib.h>

namespace {

constexpr int H = XB_HIDDEN;          // 100
constexpr int FT_ROWS = 128;          // rows of both operand tiles
constexpr int FT_KCH = 40;            // K (= pooled positions) per chunk, multiple of 8
constexpr int FT_KC = FT_KCH / 4;     // 16-byte core columns per chunk
constexpr int FT_TILE_BYTES = FT_ROWS * FT_KCH * 4;   // 20 KB
constexpr int FT_THREADS = 256;
constexpr int FT_NEVAL = 112;         // hidden channels padded to a multiple of 16

// Fill one operand tile pair (hi, lo) from `rows` x `kvalid` fp32 values (row stride in elements), zero padded.
// Warp-level mapping: one warp store = one 8x4 core matrix (128 contiguous bytes, conflict-free).
__device__ __forceinline__ void fill_tile(unsigned char *hi, unsigned char *lo, const float *__restrict__ src, size_t row_stride,
                                          int rows, int k0, int kend)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rl = lane >> 2, kl = lane & 3;
    constexpr int NIT = (FT_ROWS / 8) * FT_KC / (FT_THREADS / 32);      // 20 core matrices per warp
    float xv[NIT];
    // all loads first (one exposed global latency), then the conversions and stores
#pragma unroll
    for (int i = 0; i < NIT; ++i) {
        const int cm = warp + (FT_THREADS / 32) * i;
        const int rg = cm / FT_KC, kc = cm - rg * FT_KC;
        const int r = rg * 8 + rl, k = k0 + kc * 4 + kl;
        xv[i] = (r < rows && k < kend) ? __ldg(src + (size_t)r * row_stride + k) : 0.f;
    }
#pragma unroll
    for (int i = 0; i < NIT; ++i) {
        const int cm = warp + (FT_THREADS / 32) * i;
        const float x = xv[i];
        const float xh = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
        const uint32_t off = (uint32_t)cm * 128 + lane * 4;
        *reinterpret_cast<float *>(hi + off) = xh;
        *reinterpret_cast<float *>(lo + off) = x - xh;
    }
}

struct FcTcArgs {
    const float *pooled;     // (B,U,P)
    long long B;
    int U, P, nchunk;        // nchunk = ceil(B/128)
    const float *W1, *b1;    // (100U,P), (100U)
    // train outputs
    float *h1;               // (U,B,100)
    float *part2;            // (U,nchunk,2,100)
    // eval
    const float *g2, *be2, *rm2, *rv2, *W2, *b2, *g3, *be3, *rm3, *rv3;
    float eps;
    float *unit_out;         // (B,U)
};

template <bool EVAL>
__global__ void __launch_bounds__(FT_THREADS, 2) fc1_tc_kernel(const FcTcArgs a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    __shared__ float cf[4][128];      // train: [0] = b1 ; eval: sc2, sh2, w2
    __shared__ float red[2][128];

    unsigned char *sWh = smem, *sWl = sWh + FT_TILE_BYTES, *sAh = sWl + FT_TILE_BYTES, *sAl = sAh + FT_TILE_BYTES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nkch = (a.P + FT_KCH - 1) / FT_KCH;

    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    if (warp == 0) tc_alloc(&tmem_base_s, 128u);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t idesc = tc_idesc(2u, 128, EVAL ? FT_NEVAL : 128);
    uint32_t parity = 0;
    // operand descriptors of K step 0, computed once in warp-uniform code (see tc_elect_one)
    const int warp_u = tc_warp_uniform();
    const uint64_t dW_hi = make_smem_desc(smem_u32(sWh), 128, FT_KC * 128), dW_lo = make_smem_desc(smem_u32(sWl), 128, FT_KC * 128);
    const uint64_t dA_hi = make_smem_desc(smem_u32(sAh), 128, FT_KC * 128), dA_lo = make_smem_desc(smem_u32(sAl), 128, FT_KC * 128);

    const long long total = (long long)a.U * a.nchunk;
    int cur_u = -1;
    // contiguous tile ranges: consecutive tiles of a CTA mostly share the unit, so W1[u] is converted once
    const long long t_begin = total * blockIdx.x / gridDim.x, t_end = total * (blockIdx.x + 1) / gridDim.x;
    for (long long tile = t_begin; tile < t_end; ++tile) {
        const int u = (int)(tile / a.nchunk), chunk = (int)(tile - (long long)u * a.nchunk);
        const long long b0 = (long long)chunk * FT_ROWS;
        const int rows = (int)min((long long)FT_ROWS, a.B - b0);
        if (u != cur_u) {
            if (threadIdx.x < 128) {
                const int j = threadIdx.x, ch = u * H + j;
                const bool v = j < H;
                if (EVAL) {
                    const float sc = v ? a.g2[ch] / sqrtf(a.rv2[ch] + a.eps) : 0.f;
                    cf[0][j] = sc;
                    cf[1][j] = v ? (a.b1[ch] - a.rm2[ch]) * sc + a.be2[ch] : 0.f;
                    cf[2][j] = v ? a.W2[ch] : 0.f;
                } else {
                    cf[0][j] = v ? a.b1[ch] : 0.f;
                }
            }
        }
        for (int kc = 0; kc < nkch; ++kc) {
            const int k0 = kc * FT_KCH;
            // W1[u] chunk (reloaded per tile only when K is chunked or the unit changed)
            if (u != cur_u || nkch > 1) fill_tile(sWh, sWl, a.W1 + (size_t)u * H * a.P, (size_t)a.P, H, k0, a.P);
            fill_tile(sAh, sAl, a.pooled + ((size_t)b0 * a.U + u) * a.P, (size_t)a.U * a.P, rows, k0, a.P);
            fence_proxy_async();
            tc_fence_before();
            __syncthreads();
            if (warp_u == 0 && tc_elect_one()) {
                tc_fence_after();
#pragma unroll
                for (int ks = 0; ks < FT_KCH / 8; ++ks) {
                    // two 16-byte core columns (256 bytes = 16 in the descriptor's address field) per K = 8 step
                    const uint64_t dWh = dW_hi + ks * 16, dWl = dW_lo + ks * 16, dAh = dA_hi + ks * 16, dAl = dA_lo + ks * 16;
                    const uint32_t acc0 = (kc > 0 || ks > 0) ? 1u : 0u;
                    if (EVAL) {      // M = sequences (pooled tile is the A operand), N = hidden channels
                        tc_mma_tf32(tmem_base, dAh, dWh, idesc, acc0);
                        tc_mma_tf32(tmem_base, dAl, dWh, idesc, 1u);
                        tc_mma_tf32(tmem_base, dAh, dWl, idesc, 1u);
                    } else {         // M = hidden channels (W1 is the A operand), N = sequences
                        tc_mma_tf32(tmem_base, dWh, dAh, idesc, acc0);
                        tc_mma_tf32(tmem_base, dWl, dAh, idesc, 1u);
                        tc_mma_tf32(tmem_base, dWh, dAl, idesc, 1u);
                    }
                }
                tc_commit(&bar);
            }
            mbar_wait(&bar, parity);
            parity ^= 1;
            tc_fence_after();
        }
        cur_u = u;

        // ---------------- epilogue ----------------
        const int q = warp & 3, half = warp >> 2;
        const int row = 32 * q + lane;                                   // TMEM lane
        const uint32_t taddr = tmem_base + ((uint32_t)(32 * q) << 16);
        if (!EVAL) {
            // row = hidden channel j; this warp handles sequences [64*half, 64*half+64)
            const int j = row;
            const float bias = cf[0][j];
            float s = 0.f, sq = 0.f;
            float *dst = a.h1 + ((size_t)u * a.B + b0) * H + j;
#pragma unroll
            for (int pb = 0; pb < 4; pb += 2) {
                uint32_t r[32];
                tc_ld16_nowait(taddr + 64 * half + pb * 16, r);
                tc_ld16_nowait(taddr + 64 * half + pb * 16 + 16, r + 16);
                tc_wait_ld();
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const int b = 64 * half + pb * 16 + i;
                    const float h = __uint_as_float(r[i]) + bias;
                    if (b < rows) {
                        s += h;
                        sq = fmaf(h, h, sq);
                        if (j < H) dst[(size_t)b * H] = h;
                    }
                }
            }
            red[half][j] = s;
            __syncthreads();
            float s_all = 0.f;
            if (threadIdx.x < 128) s_all = red[0][threadIdx.x] + red[1][threadIdx.x];
            __syncthreads();
            red[half][j] = sq;
            __syncthreads();
            if (threadIdx.x < H) {
                float *pp = a.part2 + ((size_t)u * a.nchunk + chunk) * 2 * H;
                pp[threadIdx.x] = s_all;
                pp[H + threadIdx.x] = red[0][threadIdx.x] + red[1][threadIdx.x];
            }
        } else {
            // row = sequence b; warps 0-3 handle hidden channels 0..63, warps 4-7 channels 64..111
            float dot = 0.f;
#pragma unroll
            for (int pb = 0; pb < 4; ++pb) {
                if (half == 1 && pb == 3) break;
                uint32_t r[16];
                const int c0 = 64 * half + pb * 16;
                tc_ld16_nowait(taddr + c0, r);
                tc_wait_ld();
#pragma unroll
                for (int i = 0; i < 16; ++i) {
                    const int jj = c0 + i;
                    const float v = fmaxf(fmaf(__uint_as_float(r[i]), cf[0][jj], cf[1][jj]), 0.f);
                    dot = fmaf(v, cf[2][jj], dot);
                }
            }
            red[half][row] = dot;
            __syncthreads();
            if (threadIdx.x < rows) {
                const int b = threadIdx.x;
                const float h2 = (red[0][b] + red[1][b]) + a.b2[u];
                const float sc = a.g3[u] / sqrtf(a.rv3[u] + a.eps);
                a.unit_out[(b0 + b) * a.U + u] = fmaxf((h2 - a.rm3[u]) * sc + a.be3[u], 0.f);
            }
        }
        // the accumulator and the operand tiles are reused by the next tile
        tc_fence_before();
        __syncthreads();
    }
    if (warp == 0) {
        tc_fence_after();
        tc_dealloc(tmem_base, 128u);
    }
}

// ------------------------------------------------------------------------------------------------
// Backward of FC1 on the tensor cores.  Per tile (unit u, 128 sequences):
//   G[b][j]  = d loss / d h1 (BatchNorm-2 backward of the ReLU/dropout-masked upstream gradient), built on the CUDA cores
//   D1[b][p] = sum_j G[b][j] * W1[j][p]       -> d(pooled)        M = b,  N = p (48), K = j (104)
//   D2[j][p] = sum_b G[b][j] * pooled[b][p]   -> dW1 partial      M = j,  N = p (48), K = b (128)
// kind::tf32 takes K-major operands only (MN-major descriptors produce no result on this part), so the two
// products need G with different contiguous indices.  Each thread keeps its 52 G values in registers: phase 1
// stores them j-contiguous and runs D1; once those MMAs have retired, phase 2 stores them b-contiguous into the
// SAME shared-memory region and runs D2.  W1 and the pooled tile are stored transposed once.  3-term tf32 split.
// ------------------------------------------------------------------------------------------------
constexpr int BT_J = 104;                 // hidden channels padded to a multiple of 8
constexpr int BT_P = 48;                  // pooled positions padded to a multiple of 16
constexpr int BT_G_SBO = (BT_J / 4) * 128;          // 8-row group stride of a tile with K = j  (3328)
constexpr int BT_B_SBO = (FT_ROWS / 4) * 128;       // 8-row group stride of a tile with K = b  (4096)
constexpr int BT_G_BYTES = FT_ROWS * BT_J * 4;      // G  [128 b][104 j]   53,248
constexpr int BT_GT_BYTES = BT_J * FT_ROWS * 4;     // Gt [104 j][128 b]   53,248 (accumulator rows 104..127 are never read back)
constexpr int BT_H_BYTES = FT_ROWS * H * 4;         // staged h1 tile       51,200
constexpr int BT_S_BYTES = FT_ROWS * BT_P * 4;      // output staging       24,576
constexpr int BT_W_BYTES = BT_P * BT_J * 4;         // W1t [48 p][104 j]   19,968
constexpr int BT_A_BYTES = BT_P * FT_ROWS * 4;      // At  [48 p][128 b]   24,576
constexpr int BT_KT = 7;                  // column groups (of 26) per warp in the G build: kc = warp%4 + 4t
constexpr int BT_RG = 8;                  // row groups (of 16) per warp in the G build: rg = 8*(warp/4) + g

struct FcBwdArgs {
    const float *h1, *gh2, *pooled;
    long long B;
    int U, P, nchunk;
    const float *g2, *be2, *mean2, *invstd2, *W1, *W2;
    float pdrop;
    unsigned long long seed0;
    const unsigned long long *seed_dev;
    const uint32_t *dropmask;   // (U,B,4) keep bits written by the forward, or NULL (hash is recomputed)
    const float *pb2;
    float *gpool, *dW1p, *dW2, *db1, *db2, *dg2, *dbe2;
};

__device__ __forceinline__ uint32_t core_off(int r, int k, int sbo) { return (uint32_t)(r >> 3) * sbo + (uint32_t)(k >> 2) * 128 + (r & 7) * 16 + (k & 3) * 4; }

__device__ __forceinline__ void st_split(unsigned char *hi, unsigned char *lo, uint32_t off, float x)
{
    const float xh = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
    *reinterpret_cast<float *>(hi + off) = xh;
    *reinterpret_cast<float *>(lo + off) = x - xh;
}

// Three kinds of warps, so that the phases of consecutive tiles overlap:
//   warps 0..7   workers: the h1 tile from its staging buffer into registers, G arithmetic, G stored j-contiguous (D1), then, once D1 has retired, the same
//                values b-contiguous (D2) over the same region
//   warp  12     issues the tcgen05.mma of D1 and D2 into one of two accumulator pairs, commits them to mbarriers and, once
//                a D2 has retired, the TMA bulk copy of the next tile's h1 rows
//   warps 8..11  helpers: pooled tile (global -> registers at the top of the tile, transposed hi/lo operand once the
//                previous D2 has retired), then both drains (tensor memory -> staging -> coalesced global stores) while
//                the workers are already on the next tile
// Registers: the kernel starts with 128 per thread (16 warps); each group of four warps then re-sizes its share
// (setmaxnreg): workers 192 (their 56 G values and 35 coefficients stay in registers), helpers 96, the issue warp's
// group (three of its warps only wait at the final barrier) 24.
constexpr int BT_HW = 4;
constexpr int BT_THREADS = 512;
constexpr int BT_ISSUE = FT_THREADS / 32 + BT_HW;             // warp 12
constexpr int BT_WI = FT_THREADS + 32;                        // workers + issue warp: named barrier 1

__device__ __forceinline__ void bt_bar(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

#ifdef XB_PHASE_CLOCKS
__device__ long long g_bt_clk[3 * 16];
#define BT_CLK_DECL long long ck_acc[12] = {0}; long long ck_t = clock64()
#define BT_CLK(i) do { const long long n_ = clock64(); ck_acc[i] += n_ - ck_t; ck_t = n_; } while (0)
#define BT_CLK_FLUSH(role_, cond) do { if (blockIdx.x == 0 && (cond)) for (int q_ = 0; q_ < 12; ++q_) g_bt_clk[(role_) * 16 + q_] = ck_acc[q_]; } while (0)
#else
#define BT_CLK_DECL
#define BT_CLK(i) do { } while (0)
#define BT_CLK_FLUSH(role_, cond) do { } while (0)
#endif

__global__ void __launch_bounds__(BT_THREADS, 1) bwd_fc1_tc_kernel(const FcBwdArgs a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t d1_full[2], d2_full[2], acc_free[2], at_full, bar_h, h_free;
    __shared__ uint32_t tmem_base_s;
    // per hidden channel: zb = h*c1 + c0 ;  G = h*cB + cC + (zb > 0 && kept ? gh2[row] * cA : 0)
    __shared__ float c1s[128], c0s[128], cAs[128], cBs[128], cCs[128];
    __shared__ __align__(16) uint32_t msk[2][FT_ROWS][4];       // dropout keep bits of the tile's rows: this tile / the next one
    __shared__ float gh2s[2][FT_ROWS];                          // upstream gradient of the tile's rows: this tile / the next one

    unsigned char *sR = smem;                                   // G hi | G lo, then Gt hi | Gt lo
    unsigned char *sAh = sR + 2 * BT_GT_BYTES, *sAl = sAh + BT_A_BYTES;        // pooled tile transposed, hi | lo
    // the h1 tile (128 x 100 floats, contiguous in global memory) is staged by ONE TMA bulk copy into the pooled-tile
    // buffers plus a 2 KB pad: issued as soon as the previous tile's D2 (their last reader) has retired, read into
    // registers by the workers, and only then do the helpers store the pooled operand of this tile over it
    float *sH = reinterpret_cast<float *>(sAh);
    float *sS = reinterpret_cast<float *>(sAh + BT_H_BYTES);                   // staging of D1 / D2 for coalesced stores
    static_assert(BT_H_BYTES >= 2 * BT_A_BYTES, "shared-memory plan");
    unsigned char *sWh = reinterpret_cast<unsigned char *>(sS) + BT_S_BYTES, *sWl = sWh + BT_W_BYTES;
    static_assert(BT_GT_BYTES == BT_G_BYTES, "shared-memory plan");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned long long seed = a.seed0 + (a.seed_dev ? *a.seed_dev * 0xD1B54A32D192ED03ull : 0ull);
    const unsigned thresh = a.pdrop > 0.f ? (unsigned)min(4294967295.0, (double)a.pdrop * 4294967296.0) : 0u;
    const float keep_scale = a.pdrop > 0.f ? 1.0f / (1.0f - a.pdrop) : 1.0f;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&d1_full[i], 1);
            mbar_init(&d2_full[i], 1);
            mbar_init(&acc_free[i], BT_HW);
        }
        mbar_init(&at_full, BT_HW);
        mbar_init(&bar_h, 1);
        mbar_init(&h_free, 1);
        mbar_fence_init();
    }
    if (warp == 0) tc_alloc(&tmem_base_s, 512u);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;       // accumulator pair b: D1 at columns 256 b, D2 at 256 b + 128
    // The hi and lo tiles of W1t (and of the pooled tile) are adjacent in shared memory, so one MMA with N = 2*BT_P
    // multiplies by [hi ; lo] at once: per K step  X_hi.[Y_hi ; Y_lo] -> columns 0..95  and  X_lo.Y_hi -> columns 0..47.
    // The drain adds column p and column BT_P + p.  (Two MMAs instead of three: a third less operand traffic from
    // shared memory, which is what bounds these products.)
    const uint32_t idesc = tc_idesc(2u, 128, BT_P), idesc2 = tc_idesc(2u, 128, 2 * BT_P);
    static_assert(BT_W_BYTES == (BT_P / 8) * BT_G_SBO && BT_A_BYTES == (BT_P / 8) * BT_B_SBO, "hi/lo tiles must be adjacent row groups");
    const int role = tc_warp_uniform();
    const long long total = (long long)a.U * a.nchunk;
    const long long t_begin = total * blockIdx.x / gridDim.x, t_end = total * (blockIdx.x + 1) / gridDim.x;

    if (role >= BT_ISSUE) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
    }
    if (role == BT_ISSUE) {
        // =========================== issue warp ===========================
        const uint64_t dG_hi = make_smem_desc(smem_u32(sR), 128, BT_G_SBO), dG_lo = make_smem_desc(smem_u32(sR) + BT_G_BYTES, 128, BT_G_SBO);
        const uint64_t dW_hi = make_smem_desc(smem_u32(sWh), 128, BT_G_SBO);       // [W1t hi ; W1t lo]: the lo tile follows the hi tile
        const uint64_t dT_hi = make_smem_desc(smem_u32(sR), 128, BT_B_SBO), dT_lo = make_smem_desc(smem_u32(sR) + BT_GT_BYTES, 128, BT_B_SBO);
        const uint64_t dA_hi = make_smem_desc(smem_u32(sAh), 128, BT_B_SBO);       // [pooled hi ; pooled lo]
        BT_CLK_DECL;
        auto load_h1 = [&](long long t) {                        // one lane: h1 rows of tile t -> staging
            const int tu = (int)(t / a.nchunk);
            const long long tb0 = (t - (long long)tu * a.nchunk) * FT_ROWS;
            const uint32_t bytes = (uint32_t)min((long long)FT_ROWS, a.B - tb0) * H * 4;
            mbar_arrive_expect_tx(&bar_h, bytes);
            tma_bulk_g2s(sH, a.h1 + ((size_t)tu * a.B + tb0) * H, bytes, &bar_h);
        };
        if (lane == 0 && t_begin < t_end) load_h1(t_begin);
        for (long long tile = t_begin; tile < t_end; ++tile) {
            const uint32_t i = (uint32_t)(tile - t_begin), buf = i & 1u, use = i >> 1;
            const uint32_t acc = tmem_base + 256u * buf;
            bt_bar(1, BT_WI);                                    // coefficients ready (workers only need it)
            bt_bar(1, BT_WI);                                    // G stored
            BT_CLK(0);
            mbar_wait(&acc_free[buf], (use & 1u) ^ 1u);          // the helpers have drained this pair (two tiles ago)
            BT_CLK(1);
            if (tc_elect_one()) {
                tc_fence_after();
#pragma unroll
                for (int ks = 0; ks < BT_J / 8; ++ks) {          // + 16 per K step: 256 bytes in the descriptor's address field
                    tc_mma_tf32(acc, dG_hi + ks * 16, dW_hi + ks * 16, idesc2, ks > 0 ? 1u : 0u);
                    tc_mma_tf32(acc, dG_lo + ks * 16, dW_hi + ks * 16, idesc, 1u);
                }
                tc_commit(&d1_full[buf]);
            }
            __syncwarp();
            BT_CLK(2);
            bt_bar(1, BT_WI);                                    // Gt stored
            BT_CLK(3);
            mbar_wait(&at_full, i & 1u);                         // pooled operand of this tile stored by the helpers
            BT_CLK(4);
            if (tc_elect_one()) {
                tc_fence_after();
#pragma unroll
                for (int ks = 0; ks < FT_ROWS / 8; ++ks) {
                    tc_mma_tf32(acc + 128, dT_hi + ks * 16, dA_hi + ks * 16, idesc2, ks > 0 ? 1u : 0u);
                    tc_mma_tf32(acc + 128, dT_lo + ks * 16, dA_hi + ks * 16, idesc, 1u);
                }
                tc_commit(&d2_full[buf]);
            }
            __syncwarp();
            BT_CLK(5);
            if (tile + 1 < t_end) {
                // D2 was the last reader of the pooled-tile buffers: the next tile's h1 rows go there
                mbar_wait(&d2_full[buf], use & 1u);
                if (lane == 0) load_h1(tile + 1);
                __syncwarp();
            }
            BT_CLK(6);
        }
        BT_CLK_FLUSH(0, lane == 0);
    } else if (role >= FT_THREADS / 32 && role < BT_ISSUE) {
        // =========================== helper warps ===========================
        asm volatile("setmaxnreg.dec.sync.aligned.u32 96;");
        const int hw = warp - FT_THREADS / 32;                   // 0..3: core matrices hw, hw + 4, ... of the pooled operand
        const int q = warp & 3, row = 32 * q + lane;             // tensor-memory lane quarter this warp may read
        const int htid = hw * 32 + lane;
        const uint32_t lane_base = (uint32_t)(32 * q) << 16;
        // one accumulator (rows x P, hi and lo halves added) -> staging
        auto tmem_to_staging = [&](uint32_t taddr, bool row_ok) {
            float *st = sS + row * a.P;
#pragma unroll
            for (int c = 0; c < BT_P / 16; ++c) {
                if (16 * c < a.P) {
                    uint32_t r[16], r2[16];
                    tc_ld16_nowait(taddr + 16 * c, r);
                    tc_ld16_nowait(taddr + BT_P + 16 * c, r2);
                    tc_wait_ld();
                    if (row_ok) {
#pragma unroll
                        for (int p = 0; p < 16; ++p)
                            if (16 * c + p < a.P) st[16 * c + p] = __uint_as_float(r[p]) + __uint_as_float(r2[p]);
                    }
                }
            }
        };
        BT_CLK_DECL;
        for (long long tile = t_begin; tile < t_end; ++tile) {
            const uint32_t i = (uint32_t)(tile - t_begin), buf = i & 1u, use = i >> 1;
            const int u = (int)(tile / a.nchunk), chunk = (int)(tile - (long long)u * a.nchunk);
            const long long b0 = (long long)chunk * FT_ROWS;
            const int rows = (int)min((long long)FT_ROWS, a.B - b0);
            const uint32_t acc = tmem_base + 256u * buf + lane_base;
            BT_CLK(0);
            // pooled tile transposed: rows p, K = b contiguous (B operand of D2); one warp store = one core matrix.
            // All loads are issued now; they are consumed once the previous tile's D2 no longer reads the buffers.
            constexpr int NIT = (BT_P / 8) * (FT_ROWS / 4) / BT_HW;      // 48 core matrices per warp
            float xv[NIT];
            const float *psrc = a.pooled + ((size_t)b0 * a.U + u) * a.P;
            const size_t rstride = (size_t)a.U * a.P;
#pragma unroll
            for (int k = 0; k < NIT; ++k) {
                const int cm = hw + BT_HW * k;
                const int pg = cm / (FT_ROWS / 4), bc = cm - pg * (FT_ROWS / 4);
                const int p = pg * 8 + (lane >> 2), r = bc * 4 + (lane & 3);
                xv[k] = (r < rows && p < a.P) ? __ldg(psrc + r * rstride + p) : 0.f;
            }
            BT_CLK(1);
            mbar_wait(&h_free, i & 1u);                          // the workers have read the h1 rows staged in these buffers
            BT_CLK(2);
#pragma unroll
            for (int k = 0; k < NIT; ++k) st_split(sAh, sAl, (uint32_t)(hw + BT_HW * k) * 128 + lane * 4, xv[k]);
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&at_full);
            // ---- D1 (row = sequence): staged, then written with consecutive threads on consecutive addresses ----
            BT_CLK(3);
            mbar_wait(&d1_full[buf], use & 1u);
            BT_CLK(4);
            tc_fence_after();
            tmem_to_staging(acc, true);
            bt_bar(2, BT_HW * 32);
            {
                float *dst = a.gpool + ((size_t)b0 * a.U + u) * a.P;
                const int n = rows * a.P, dr = (BT_HW * 32) / a.P, dp = BT_HW * 32 - dr * a.P;
                int rr = htid / a.P, pp = htid - rr * a.P;
                for (int k = htid; k < n; k += BT_HW * 32) {
                    dst[rr * rstride + pp] = sS[k];
                    rr += dr; pp += dp;
                    if (pp >= a.P) { pp -= a.P; ++rr; }
                }
            }
            bt_bar(2, BT_HW * 32);                               // the copy-out has finished reading the staging buffer
            BT_CLK(5);
            // ---- D2 (row = hidden channel): the 100 x P tile is contiguous in global memory ----
            mbar_wait(&d2_full[buf], use & 1u);
            BT_CLK(6);
            tc_fence_after();
            tmem_to_staging(acc + 128, row < H);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_free[buf]);          // both accumulators of the pair have been read
            bt_bar(2, BT_HW * 32);
            {
                float *dst = a.dW1p + ((size_t)chunk * a.U + u) * H * a.P;
                const int n = H * a.P;
                for (int k = htid; k < n; k += BT_HW * 32) dst[k] = sS[k];
            }
            bt_bar(2, BT_HW * 32);
            BT_CLK(7);
        }
        BT_CLK_FLUSH(1, htid == 0);
    } else if (role < FT_THREADS / 32) {
        // =========================== worker warps ===========================
        asm volatile("setmaxnreg.inc.sync.aligned.u32 192;");
        // upstream gradient and dropout keep bits of the rows of tile `t` into buffer `rbuf` (threads 0..127, one row each)
        auto row_data = [&](long long t, int rbuf) {
            if (threadIdx.x >= FT_ROWS) return;
            const int r = threadIdx.x;
            const int tu = (int)(t / a.nchunk);
            const long long tb0 = (t - (long long)tu * a.nchunk) * FT_ROWS;
            const int trows = (int)min((long long)FT_ROWS, a.B - tb0);
            const size_t e = (size_t)tu * a.B + tb0 + r;
            if (r < trows) {
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(&gh2s[rbuf][r])), "l"(a.gh2 + e) : "memory");
            } else {
                gh2s[rbuf][r] = 0.f;
            }
            if (a.pdrop > 0.f && r < trows && a.dropmask) {
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(&msk[rbuf][r][0])), "l"(a.dropmask + e * 4) : "memory");
            } else {
                uint4 mw = make_uint4(~0u, ~0u, ~0u, ~0u);
                if (a.pdrop > 0.f && r < trows) {     // no saved keep bits: regenerate them from the counter hash (slow path)
                    uint32_t w[4] = {0u, 0u, 0u, 0u};
                    for (int qd = 0; qd < H / 4; ++qd) w[qd >> 3] |= drop_keep4(seed, e * (H / 4) + qd, thresh) << (4 * (qd & 7));
                    mw = make_uint4(w[0], w[1], w[2], w[3]);
                }
                *reinterpret_cast<uint4 *>(&msk[rbuf][r][0]) = mw;
            }
        };
        if (t_begin < t_end) row_data(t_begin, 0);
        asm volatile("cp.async.commit_group;" ::: "memory");
        int cur_u = -1;
        // Warp w owns the core matrices (row group rg = 8*(w/4) + g, column group kc = w%4 + 4t), g < 8, t < 7: thread
        // (rl = lane/4, jl = lane%4) sees rows r = 8*rg + rl and only seven channels j = 4*kc + jl.
        const int kq = warp & 3, rg0 = BT_RG * (warp >> 2), rl = lane >> 2, jl = lane & 3;
        BT_CLK_DECL;
        for (long long tile = t_begin; tile < t_end; ++tile) {
            BT_CLK(0);
            const uint32_t i = (uint32_t)(tile - t_begin), buf = i & 1u, use = i >> 1;
            const int rb = (int)buf;
            const int u = (int)(tile / a.nchunk), chunk = (int)(tile - (long long)u * a.nchunk);
            const long long b0 = (long long)chunk * FT_ROWS;
            const int rows = (int)min((long long)FT_ROWS, a.B - b0);
            if (tile + 1 < t_end) row_data(tile + 1, rb ^ 1);
            asm volatile("cp.async.commit_group;" ::: "memory");
            if (u != cur_u) {
                if (threadIdx.x < 128) {
                    const int j = threadIdx.x, ch = u * H + j;
                    const bool v = j < H;
                    float acc = 0.f, s1 = 0.f, s2 = 0.f;
                    if (v) {
                        for (int k = 0; k < a.nchunk; ++k) {
                            const float *pp = a.pb2 + ((size_t)u * a.nchunk + k) * 3 * H;
                            acc += pp[j]; s1 += pp[H + j]; s2 += pp[2 * H + j];
                        }
                        if (chunk == 0) {
                            if (a.dW2) a.dW2[ch] = acc;
                            if (a.dbe2) a.dbe2[ch] = s1;
                            if (a.dg2) a.dg2[ch] = s2;
                            if (a.db1) a.db1[ch] = 0.f;          // bias followed by train-mode BatchNorm: gradient is exactly zero
                            if (a.db2 && j == 0) a.db2[u] = 0.f;
                        }
                    }
                    const float mu = v ? a.mean2[ch] : 0.f, inv = v ? a.invstd2[ch] : 0.f, gam = v ? a.g2[ch] : 0.f;
                    const float bet = v ? a.be2[ch] : 0.f, w2 = v ? a.W2[ch] : 0.f;
                    const float m1 = s1 / (float)a.B, m2 = s2 / (float)a.B;
                    const float c1 = gam * inv;
                    c1s[j] = c1;
                    c0s[j] = bet - mu * c1;
                    cAs[j] = w2 * c1 * keep_scale;
                    cBs[j] = -c1 * m2 * inv;
                    cCs[j] = c1 * (m2 * inv * mu - m1);
                }
                // W1[u] transposed: rows p, K = j contiguous (B operand of D1); one warp store = one core matrix.  The
                // previous tile's D1 (its last reader) retired before that tile's second store pass.
                {
                    constexpr int NCM = (BT_P / 8) * (BT_J / 4), NIT = (NCM + FT_THREADS / 32 - 1) / (FT_THREADS / 32);
                    float xv[NIT];
                    const float *wsrc = a.W1 + (size_t)u * H * a.P;
#pragma unroll
                    for (int k = 0; k < NIT; ++k) {
                        const int cm = warp + (FT_THREADS / 32) * k;
                        const int pg = cm / (BT_J / 4), kc = cm - pg * (BT_J / 4);
                        const int p = pg * 8 + (lane >> 2), j = kc * 4 + (lane & 3);
                        xv[k] = (cm < NCM && j < H && p < a.P) ? __ldg(wsrc + j * a.P + p) : 0.f;
                    }
#pragma unroll
                    for (int k = 0; k < NIT; ++k) {
                        const int cm = warp + (FT_THREADS / 32) * k;
                        if (cm < NCM) st_split(sWh, sWl, (uint32_t)cm * 128 + lane * 4, xv[k]);
                    }
                }
                cur_u = u;
            }
            BT_CLK(1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");      // this tile's row data (copied during the previous tile)
            bt_bar(1, BT_WI);                                          // coefficient arrays and row data visible
            BT_CLK(2);
            // the h1 values of this thread from the staged tile (its copy was issued when the previous D2, the last reader
            // of the G region as well, had retired); once every worker holds its values the buffers go to the helpers
            mbar_wait(&bar_h, i & 1u);
            float gv[BT_KT * BT_RG];
#pragma unroll
            for (int t = 0; t < BT_KT; ++t) {
                const int j = 4 * (kq + 4 * t) + jl;
#pragma unroll
                for (int g = 0; g < BT_RG; ++g) {
                    const int r = (rg0 + g) * 8 + rl;
                    gv[t * BT_RG + g] = (j < H && r < rows) ? sH[r * H + j] : 0.f;
                }
            }
            bt_bar(3, FT_THREADS);
            if (threadIdx.x == 0) mbar_arrive(&h_free);
            BT_CLK(3);
            // ---- G values and the j-contiguous operand of D1 ----
            {
                float k1[BT_KT], k0[BT_KT], kA[BT_KT], kB[BT_KT], kC[BT_KT];
#pragma unroll
                for (int t = 0; t < BT_KT; ++t) {
                    const int j = 4 * (kq + 4 * t) + jl;      // <= 111: inside the 128-entry arrays, zero beyond channel 99
                    k1[t] = c1s[j]; k0[t] = c0s[j]; kA[t] = cAs[j]; kB[t] = cBs[j]; kC[t] = cCs[j];
                }
#pragma unroll
                for (int g = 0; g < BT_RG; ++g) {
                    const int rg = rg0 + g, r = rg * 8 + rl;
                    const float gh = gh2s[rb][r];
                    const uint4 mw = *reinterpret_cast<const uint4 *>(&msk[rb][r][0]);
                    const uint32_t mws[4] = {mw.x, mw.y, mw.z, mw.w};
#pragma unroll
                    for (int t = 0; t < BT_KT; ++t) {
                        const int kc = kq + 4 * t;
                        if (kc < BT_J / 4) {                       // warp-uniform (t = 6 exists only for w%4 < 2)
                            const int j = 4 * kc + jl;
                            const float h = gv[t * BT_RG + g];
                            const float zb = fmaf(h, k1[t], k0[t]);
                            const bool kept = (mws[t >> 1] >> (j & 31)) & 1u;    // word j/32 == t/2
                            float v = fmaf(h, kB[t], kC[t]);
                            if (zb > 0.f && kept) v = fmaf(gh, kA[t], v);
                            if (r >= rows) v = 0.f;
                            gv[t * BT_RG + g] = v;
                            st_split(sR, sR + BT_G_BYTES, (uint32_t)(rg * (BT_J / 4) + kc) * 128 + lane * 4, v);      // G, j-contiguous
                        }
                    }
                }
            }
            BT_CLK(4);
            fence_proxy_async();
            bt_bar(1, BT_WI);                                          // -> the issue warp starts D1
            BT_CLK(5);
            // ---- once D1 has retired: the same values b-contiguous (Gt[j][b]) over the region it no longer reads ----
            mbar_wait(&d1_full[buf], use & 1u);
            BT_CLK(6);
#pragma unroll
            for (int t = 0; t < BT_KT; ++t) {
                const int kc = kq + 4 * t;
                if (kc < BT_J / 4) {
#pragma unroll
                    for (int g = 0; g < BT_RG; ++g)
                        st_split(sR, sR + BT_GT_BYTES, core_off(4 * kc + jl, (rg0 + g) * 8 + rl, BT_B_SBO), gv[t * BT_RG + g]);
                }
            }
            // rows 104..127 of Gt are never written: they only feed accumulator rows that are not read back
            BT_CLK(7);
            fence_proxy_async();
            bt_bar(1, BT_WI);                                          // -> the issue warp starts D2
            BT_CLK(8);
        }
        BT_CLK_FLUSH(2, threadIdx.x == 0);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tc_dealloc(tmem_base, 512u);
    }
}

// ------------------------------------------------------------------------------------------------
// Software-pipelined eval-mode FC stack (P <= 40: one K chunk).  Same arithmetic, operand order and reduction
// order as fc1_tc_kernel<true>; only the schedule differs.  Two TMEM accumulators: while the tensor core works on
// tile t, the CUDA cores issue the global loads of tile t+1 and run the epilogue of tile t-1.
// ------------------------------------------------------------------------------------------------
constexpr int FT_NIT = (FT_ROWS / 8) * FT_KC / (FT_THREADS / 32);      // 20 core matrices per warp

__device__ __forceinline__ void tile_load(float (&xv)[FT_NIT], const float *__restrict__ src, size_t row_stride, int rows, int kend)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int rl = lane >> 2, kl = lane & 3;
#pragma unroll
    for (int i = 0; i < FT_NIT; ++i) {
        const int cm = warp + (FT_THREADS / 32) * i;
        const int rg = cm / FT_KC, kc = cm - rg * FT_KC;
        const int r = rg * 8 + rl, k = kc * 4 + kl;
        xv[i] = (r < rows && k < kend) ? __ldg(src + (size_t)r * row_stride + k) : 0.f;
    }
}

__device__ __forceinline__ void tile_store(const float (&xv)[FT_NIT], unsigned char *hi, unsigned char *lo)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < FT_NIT; ++i) {
        const int cm = warp + (FT_THREADS / 32) * i;
        const float x = xv[i];
        const float xh = __uint_as_float(__float_as_uint(x) & 0xFFFFE000u);
        const uint32_t off = (uint32_t)cm * 128 + lane * 4;
        *reinterpret_cast<float *>(hi + off) = xh;
        *reinterpret_cast<float *>(lo + off) = x - xh;
    }
}

__global__ void __launch_bounds__(FT_THREADS, 2) fc1_tc_eval_pipe_kernel(const FcTcArgs a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    __shared__ __align__(16) float4 cf[2][128];     // per unit parity: {BN-2 scale, BN-2 shift (+bias), W2, -}
    __shared__ float post[2][4];                    // per unit parity: b2, rm3, BN-3 scale, be3
    __shared__ float red[2][128];

    unsigned char *sWh = smem, *sWl = sWh + FT_TILE_BYTES, *sAh = sWl + FT_TILE_BYTES, *sAl = sAh + FT_TILE_BYTES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    if (warp == 0) tc_alloc(&tmem_base_s, 256u);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t idesc = tc_idesc(2u, 128, FT_NEVAL);
    uint32_t parity = 0;
    // operand descriptors of K step 0, computed once in warp-uniform code (see tc_elect_one)
    const int warp_u = tc_warp_uniform();
    const uint64_t dW_hi = make_smem_desc(smem_u32(sWh), 128, FT_KC * 128), dW_lo = make_smem_desc(smem_u32(sWl), 128, FT_KC * 128);
    const uint64_t dA_hi = make_smem_desc(smem_u32(sAh), 128, FT_KC * 128), dA_lo = make_smem_desc(smem_u32(sAl), 128, FT_KC * 128);

    const long long total = (long long)a.U * a.nchunk;
    const long long t_begin = total * blockIdx.x / gridDim.x, t_end = total * (blockIdx.x + 1) / gridDim.x;
    const size_t row_stride = (size_t)a.U * a.P;
    float xv[FT_NIT];
    int cur_u = -1, sel = 0;
    int p_u = 0, p_rows = 0, p_sel = 0;
    long long p_b0 = 0;
    if (t_begin < t_end) {
        const int u = (int)(t_begin / a.nchunk), chunk = (int)(t_begin - (long long)u * a.nchunk);
        const long long b0 = (long long)chunk * FT_ROWS;
        tile_load(xv, a.pooled + ((size_t)b0 * a.U + u) * a.P, row_stride, (int)min((long long)FT_ROWS, a.B - b0), a.P);
    }
    int it = 0;
    for (long long tile = t_begin; tile <= t_end && t_begin < t_end; ++tile, ++it) {
        const bool have = tile < t_end;
        if (it > 0) {        // the MMAs of the previous tile have retired: its accumulator is ready, the operand tiles are free
            mbar_wait(&bar, parity);
            parity ^= 1;
            tc_fence_after();
        }
        int u = 0, rows = 0;
        long long b0 = 0;
        if (have) {
            u = (int)(tile / a.nchunk);
            const int chunk = (int)(tile - (long long)u * a.nchunk);
            b0 = (long long)chunk * FT_ROWS;
            rows = (int)min((long long)FT_ROWS, a.B - b0);
            if (u != cur_u) {
                sel ^= 1;
                if (threadIdx.x < 128) {
                    const int j = threadIdx.x, ch = u * H + j;
                    const bool v = j < H;
                    const float sc = v ? a.g2[ch] / sqrtf(a.rv2[ch] + a.eps) : 0.f;
                    cf[sel][j] = make_float4(sc, v ? (a.b1[ch] - a.rm2[ch]) * sc + a.be2[ch] : 0.f, v ? a.W2[ch] : 0.f, 0.f);
                }
                if (threadIdx.x == 128) {
                    post[sel][0] = a.b2[u];
                    post[sel][1] = a.rm3[u];
                    post[sel][2] = a.g3[u] / sqrtf(a.rv3[u] + a.eps);
                    post[sel][3] = a.be3[u];
                }
                fill_tile(sWh, sWl, a.W1 + (size_t)u * H * a.P, (size_t)a.P, H, 0, a.P);
                cur_u = u;
            }
            tile_store(xv, sAh, sAl);
            fence_proxy_async();
        }
        tc_fence_before();
        __syncthreads();
        if (have && warp_u == 0 && tc_elect_one()) {
            tc_fence_after();
            const uint32_t acc = tmem_base + (uint32_t)(it & 1) * 128u;
#pragma unroll
            for (int ks = 0; ks < FT_KCH / 8; ++ks) {
                const uint64_t dWh = dW_hi + ks * 16, dWl = dW_lo + ks * 16, dAh = dA_hi + ks * 16, dAl = dA_lo + ks * 16;
                tc_mma_tf32(acc, dAh, dWh, idesc, ks > 0 ? 1u : 0u);
                tc_mma_tf32(acc, dAl, dWh, idesc, 1u);
                tc_mma_tf32(acc, dAh, dWl, idesc, 1u);
            }
            tc_commit(&bar);
        }
        if (tile + 1 < t_end) {      // global loads of the next tile fly during this tile's MMAs and the epilogue below
            const int un = (int)((tile + 1) / a.nchunk), chunk = (int)((tile + 1) - (long long)un * a.nchunk);
            const long long bn = (long long)chunk * FT_ROWS;
            tile_load(xv, a.pooled + ((size_t)bn * a.U + un) * a.P, row_stride, (int)min((long long)FT_ROWS, a.B - bn), a.P);
        }
        if (it > 0) {
            // ---------------- epilogue of the previous tile ----------------
            const int q = warp & 3, half = warp >> 2;
            const int row = 32 * q + lane;
            const uint32_t taddr = tmem_base + (uint32_t)((it - 1) & 1) * 128u + ((uint32_t)(32 * q) << 16);
            const float4 *c = cf[p_sel];
            float dot = 0.f;
            if (half == 0) {
                uint32_t r[64];
                tc_ld16_nowait(taddr, r);
                tc_ld16_nowait(taddr + 16, r + 16);
                tc_ld16_nowait(taddr + 32, r + 32);
                tc_ld16_nowait(taddr + 48, r + 48);
                tc_wait_ld();
#pragma unroll
                for (int i = 0; i < 64; ++i) {
                    const float4 k = c[i];
                    dot = fmaf(fmaxf(fmaf(__uint_as_float(r[i]), k.x, k.y), 0.f), k.z, dot);
                }
            } else {
                uint32_t r[48];
                tc_ld16_nowait(taddr + 64, r);
                tc_ld16_nowait(taddr + 80, r + 16);
                tc_ld16_nowait(taddr + 96, r + 32);
                tc_wait_ld();
#pragma unroll
                for (int i = 0; i < 48; ++i) {
                    const float4 k = c[64 + i];
                    dot = fmaf(fmaxf(fmaf(__uint_as_float(r[i]), k.x, k.y), 0.f), k.z, dot);
                }
            }
            red[half][row] = dot;
            tc_fence_before();
            __syncthreads();
            if (threadIdx.x < p_rows) {
                const int b = threadIdx.x;
                const float h2 = (red[0][b] + red[1][b]) + post[p_sel][0];
                a.unit_out[(p_b0 + b) * a.U + p_u] = fmaxf((h2 - post[p_sel][1]) * post[p_sel][2] + post[p_sel][3], 0.f);
            }
        }
        p_u = u; p_rows = rows; p_sel = sel; p_b0 = b0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tc_dealloc(tmem_base, 256u);
    }
}

// ------------------------------------------------------------------------------------------------
// Software-pipelined training-mode FC1 (P <= 40: one K chunk): the schedule of fc1_tc_eval_pipe_kernel with the
// epilogue of fc1_tc_kernel<false> (thread = hidden channel: + bias, h1 store, BatchNorm-2 partial sums).  Same
// MMAs, operand order and summation order as the unpipelined kernel, so h1 and the partial sums are bit-identical.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(FT_THREADS, 2) fc1_tc_train_pipe_kernel(const FcTcArgs a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    __shared__ float bias_s[2][128];                // per unit parity
    __shared__ float red_s[2][128], red_q[2][128];

    unsigned char *sWh = smem, *sWl = sWh + FT_TILE_BYTES, *sAh = sWl + FT_TILE_BYTES, *sAl = sAh + FT_TILE_BYTES;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    if (warp == 0) tc_alloc(&tmem_base_s, 256u);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t idesc = tc_idesc(2u, 128, 128);
    uint32_t parity = 0;
    const int warp_u = tc_warp_uniform();
    const uint64_t dW_hi = make_smem_desc(smem_u32(sWh), 128, FT_KC * 128), dW_lo = make_smem_desc(smem_u32(sWl), 128, FT_KC * 128);
    const uint64_t dA_hi = make_smem_desc(smem_u32(sAh), 128, FT_KC * 128), dA_lo = make_smem_desc(smem_u32(sAl), 128, FT_KC * 128);

    const long long total = (long long)a.U * a.nchunk;
    const long long t_begin = total * blockIdx.x / gridDim.x, t_end = total * (blockIdx.x + 1) / gridDim.x;
    const size_t row_stride = (size_t)a.U * a.P;
    float xv[FT_NIT];
    int cur_u = -1, sel = 0;
    int p_u = 0, p_rows = 0, p_sel = 0, p_chunk = 0;
    long long p_b0 = 0;
    if (t_begin < t_end) {
        const int u = (int)(t_begin / a.nchunk), chunk = (int)(t_begin - (long long)u * a.nchunk);
        const long long b0 = (long long)chunk * FT_ROWS;
        tile_load(xv, a.pooled + ((size_t)b0 * a.U + u) * a.P, row_stride, (int)min((long long)FT_ROWS, a.B - b0), a.P);
    }
    int it = 0;
    for (long long tile = t_begin; tile <= t_end && t_begin < t_end; ++tile, ++it) {
        const bool have = tile < t_end;
        if (it > 0) {        // the MMAs of the previous tile have retired: its accumulator is ready, the operand tiles are free
            mbar_wait(&bar, parity);
            parity ^= 1;
            tc_fence_after();
        }
        int u = 0, rows = 0, chunk = 0;
        long long b0 = 0;
        if (have) {
            u = (int)(tile / a.nchunk);
            chunk = (int)(tile - (long long)u * a.nchunk);
            b0 = (long long)chunk * FT_ROWS;
            rows = (int)min((long long)FT_ROWS, a.B - b0);
            if (u != cur_u) {
                sel ^= 1;
                if (threadIdx.x < 128) bias_s[sel][threadIdx.x] = threadIdx.x < H ? a.b1[u * H + threadIdx.x] : 0.f;
                fill_tile(sWh, sWl, a.W1 + (size_t)u * H * a.P, (size_t)a.P, H, 0, a.P);
                cur_u = u;
            }
            tile_store(xv, sAh, sAl);
            fence_proxy_async();
        }
        tc_fence_before();
        __syncthreads();
        if (have && warp_u == 0 && tc_elect_one()) {
            tc_fence_after();
            const uint32_t acc = tmem_base + (uint32_t)(it & 1) * 128u;
#pragma unroll
            for (int ks = 0; ks < FT_KCH / 8; ++ks) {      // M = hidden channels (W1 is the A operand), N = sequences
                const uint64_t dWh = dW_hi + ks * 16, dWl = dW_lo + ks * 16, dAh = dA_hi + ks * 16, dAl = dA_lo + ks * 16;
                tc_mma_tf32(acc, dWh, dAh, idesc, ks > 0 ? 1u : 0u);
                tc_mma_tf32(acc, dWl, dAh, idesc, 1u);
                tc_mma_tf32(acc, dWh, dAl, idesc, 1u);
            }
            tc_commit(&bar);
        }
        if (tile + 1 < t_end) {      // global loads of the next tile fly during this tile's MMAs and the epilogue below
            const int un = (int)((tile + 1) / a.nchunk), cn = (int)((tile + 1) - (long long)un * a.nchunk);
            const long long bn = (long long)cn * FT_ROWS;
            tile_load(xv, a.pooled + ((size_t)bn * a.U + un) * a.P, row_stride, (int)min((long long)FT_ROWS, a.B - bn), a.P);
        }
        if (it > 0) {
            // ---------------- epilogue of the previous tile: row = hidden channel j, this warp's 64 sequences ----------------
            const int q = warp & 3, half = warp >> 2;
            const int j = 32 * q + lane;
            const uint32_t taddr = tmem_base + (uint32_t)((it - 1) & 1) * 128u + ((uint32_t)(32 * q) << 16);
            const float bias = bias_s[p_sel][j];
            float sm = 0.f, sq = 0.f;
            float *dst = a.h1 + ((size_t)p_u * a.B + p_b0) * H + j;
#pragma unroll
            for (int pb = 0; pb < 4; pb += 2) {
                uint32_t r[32];
                tc_ld16_nowait(taddr + 64 * half + pb * 16, r);
                tc_ld16_nowait(taddr + 64 * half + pb * 16 + 16, r + 16);
                tc_wait_ld();
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    const int b = 64 * half + pb * 16 + i;
                    const float h = __uint_as_float(r[i]) + bias;
                    if (b < p_rows) {
                        sm += h;
                        sq = fmaf(h, h, sq);
                        if (j < H) dst[(size_t)b * H] = h;
                    }
                }
            }
            red_s[half][j] = sm;
            red_q[half][j] = sq;
            tc_fence_before();
            __syncthreads();
            if (threadIdx.x < H) {
                float *pp = a.part2 + ((size_t)p_u * a.nchunk + p_chunk) * 2 * H;
                pp[threadIdx.x] = red_s[0][threadIdx.x] + red_s[1][threadIdx.x];
                pp[H + threadIdx.x] = red_q[0][threadIdx.x] + red_q[1][threadIdx.x];
            }
        }
        p_u = u; p_rows = rows; p_sel = sel; p_b0 = b0; p_chunk = chunk;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) {
        tc_fence_after();
        tc_dealloc(tmem_base, 256u);
    }
}

// ------------------------------------------------------------------------------------------------
// Final linear layer (networks.py:118,160-162: out = unit_out @ Wf^T + bias) on the tensor cores, 3-term tf32 split.
//   D[b][c] = sum_u x[b][u] * Wf[c][u]      M = 128 sequences, N = classes padded to 16 (<= 128), K = U in chunks of 40
// Both operands are K-major as they lie in memory.  One tile of 128 sequences per CTA; the global loads of K chunk
// k+1 are in flight while the MMAs of chunk k run.  The output tile is staged in shared memory and written coalesced.
// ------------------------------------------------------------------------------------------------
struct LinTcArgs {
    const float *x, *W, *bias;
    float *out;
    long long B;
    int U, C, npad;
};

__global__ void __launch_bounds__(FT_THREADS, 1) linear_tc_kernel(const LinTcArgs a)
{
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    unsigned char *sWh = smem, *sWl = sWh + FT_TILE_BYTES, *sAh = sWl + FT_TILE_BYTES, *sAl = sAh + FT_TILE_BYTES;
    float *sOut = reinterpret_cast<float *>(smem);       // reused after the last MMA: [128][C+1] floats (<= 66 KB)
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nkch = (a.U + FT_KCH - 1) / FT_KCH;

    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    if (warp == 0) tc_alloc(&tmem_base_s, 128u);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_base_s;
    const uint32_t idesc = tc_idesc(2u, 128, a.npad);
    uint32_t parity = 0;
    // operand descriptors of K step 0, computed once in warp-uniform code (see tc_elect_one)
    const int warp_u = tc_warp_uniform();
    const uint64_t dW_hi = make_smem_desc(smem_u32(sWh), 128, FT_KC * 128), dW_lo = make_smem_desc(smem_u32(sWl), 128, FT_KC * 128);
    const uint64_t dA_hi = make_smem_desc(smem_u32(sAh), 128, FT_KC * 128), dA_lo = make_smem_desc(smem_u32(sAl), 128, FT_KC * 128);

    const long long b0 = (long long)blockIdx.x * FT_ROWS;
    const int rows = (int)min((long long)FT_ROWS, a.B - b0);
    const float *xsrc = a.x + (size_t)b0 * a.U;
    float xv[FT_NIT], wv[FT_NIT];
    tile_load(xv, xsrc, (size_t)a.U, rows, a.U);
    tile_load(wv, a.W, (size_t)a.U, a.C, a.U);
    for (int kc = 0; kc < nkch; ++kc) {
        if (kc > 0) {        // MMAs of the previous chunk have retired: operand tiles are free again
            mbar_wait(&bar, parity);
            parity ^= 1;
            tc_fence_after();
        }
        tile_store(xv, sAh, sAl);
        tile_store(wv, sWh, sWl);
        fence_proxy_async();
        tc_fence_before();
        __syncthreads();
        if (warp_u == 0 && tc_elect_one()) {
            tc_fence_after();
#pragma unroll
            for (int ks = 0; ks < FT_KCH / 8; ++ks) {
                const uint64_t dWh = dW_hi + ks * 16, dWl = dW_lo + ks * 16, dAh = dA_hi + ks * 16, dAl = dA_lo + ks * 16;
                tc_mma_tf32(tmem_base, dAh, dWh, idesc, (kc > 0 || ks > 0) ? 1u : 0u);
                tc_mma_tf32(tmem_base, dAl, dWh, idesc, 1u);
                tc_mma_tf32(tmem_base, dAh, dWl, idesc, 1u);
            }
            tc_commit(&bar);
        }
        if (kc + 1 < nkch) {
            const int k0 = (kc + 1) * FT_KCH;
            tile_load(xv, xsrc + k0, (size_t)a.U, rows, a.U - k0);
            tile_load(wv, a.W + k0, (size_t)a.U, a.C, a.U - k0);
        }
    }
    mbar_wait(&bar, parity);
    tc_fence_after();

    // ---------------- epilogue: TMEM -> shared memory (+bias) -> coalesced global stores ----------------
    const int q = warp & 3, half = warp >> 2;
    const int row = 32 * q + lane;
    const uint32_t taddr = tmem_base + ((uint32_t)(32 * q) << 16);
    const int ld = a.C + 1;
    for (int c0 = 64 * half; c0 < min(a.npad, 64 * half + 64); c0 += 16) {
        uint32_t r[16];
        tc_ld16_nowait(taddr + c0, r);
        tc_wait_ld();
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (c0 + i < a.C) sOut[row * ld + c0 + i] = __uint_as_float(r[i]) + __ldg(a.bias + c0 + i);
    }
    tc_fence_before();
    __syncthreads();
    float *dst = a.out + (size_t)b0 * a.C;
    const int n = rows * a.C;
    for (int i = threadIdx.x; i < n; i += FT_THREADS) {
        const int r = i / a.C, c = i - r * a.C;
        dst[i] = sOut[r * ld + c];
    }
    if (warp == 0) {
        tc_fence_after();
        tc_dealloc(tmem_base, 128u);
    }
}

int launch_fc1_tc(bool eval, const FcTcArgs &a, cudaStream_t stream)
{
    const size_t smem = 4 * (size_t)FT_TILE_BYTES;
    const long long total = (long long)a.U * a.nchunk;
    int grid = 2 * xb_num_sms();
    if (total < grid) grid = (int)total;
    cudaError_t e;
#ifdef XB_EXPERIMENT     // ablation switch of the experiment build only (build_lib.py: XB_NVCC_EXTRA=-DXB_EXPERIMENT)
    static const bool no_pipe = [] { const char *v = getenv("XB_FC_PIPE"); return v && v[0] == '0'; }();
#else
    constexpr bool no_pipe = false;
#endif
    if (eval && a.P <= FT_KCH && !no_pipe) {
        e = cudaFuncSetAttribute(fc1_tc_eval_pipe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) fc1_tc_eval_pipe_kernel<<<grid, FT_THREADS, smem, stream>>>(a);
    } else if (eval) {
        e = cudaFuncSetAttribute(fc1_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) fc1_tc_kernel<true><<<grid, FT_THREADS, smem, stream>>>(a);
    } else if (a.P <= FT_KCH && !no_pipe) {
        e = cudaFuncSetAttribute(fc1_tc_train_pipe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) fc1_tc_train_pipe_kernel<<<grid, FT_THREADS, smem, stream>>>(a);
    } else {
        e = cudaFuncSetAttribute(fc1_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) fc1_tc_kernel<false><<<grid, FT_THREADS, smem, stream>>>(a);
    }
    if (e != cudaSuccess) {
        xb_set_error("fc1_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    return xb_check_launch("fc1_tc");
}

}  // namespace

// used by fc.cu
int xb_fc1_fwd_train_tc(const float *pooled, long long B, int U, int P, const float *W1, const float *b1, float *h1, float *part2,
                        cudaStream_t stream)
{
    FcTcArgs a{};
    a.pooled = pooled; a.B = B; a.U = U; a.P = P; a.nchunk = (int)((B + FT_ROWS - 1) / FT_ROWS);
    a.W1 = W1; a.b1 = b1; a.h1 = h1; a.part2 = part2;
    return launch_fc1_tc(false, a, stream);
}

int xb_fc_stack_fwd_eval_tc(const float *pooled, long long B, int U, int P, const float *W1, const float *b1, const float *g2,
                            const float *be2, const float *rm2, const float *rv2, const float *W2, const float *b2, const float *g3,
                            const float *be3, const float *rm3, const float *rv3, float eps, float *unit_out, cudaStream_t stream)
{
    FcTcArgs a{};
    a.pooled = pooled; a.B = B; a.U = U; a.P = P; a.nchunk = (int)((B + FT_ROWS - 1) / FT_ROWS);
    a.W1 = W1; a.b1 = b1; a.g2 = g2; a.be2 = be2; a.rm2 = rm2; a.rv2 = rv2; a.W2 = W2; a.b2 = b2;
    a.g3 = g3; a.be3 = be3; a.rm3 = rm3; a.rv3 = rv3; a.eps = eps; a.unit_out = unit_out;
    return launch_fc1_tc(true, a, stream);
}


// returns XB_ERR_UNSUPPORTED (without setting an error) when the class count does not fit one N tile
int xb_linear_tc(const float *x, long long B, int U, int C, const float *W, const float *bias, float *out, cudaStream_t stream)
{
    if (C > 128) return XB_ERR_UNSUPPORTED;
    LinTcArgs a{};
    a.x = x; a.W = W; a.bias = bias; a.out = out; a.B = B; a.U = U; a.C = C; a.npad = (C + 15) / 16 * 16;
    const size_t smem = 4 * (size_t)FT_TILE_BYTES;
    cudaError_t e = cudaFuncSetAttribute(linear_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        xb_set_error("linear_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    linear_tc_kernel<<<(unsigned)((B + FT_ROWS - 1) / FT_ROWS), FT_THREADS, smem, stream>>>(a);
    return xb_check_launch("linear_tc");
}

#ifdef XB_PHASE_CLOCKS
extern "C" int xb_dbg_phase_clocks(long long *host) { return (int)cudaMemcpyFromSymbol(host, g_bt_clk, sizeof(g_bt_clk)); }
#endif

bool xb_fc_tc_enabled()
{
#ifdef XB_EXPERIMENT     // XB_FC_TC=0 selects the CUDA-core kernels (experiment build only)
    const char *e = getenv("XB_FC_TC");
    return !(e && e[0] == '0');
#else
    return true;
#endif
}

// returns XB_ERR_UNSUPPORTED (without setting an error) when the shape does not fit the tensor-core tiles
int xb_bwd_fc1_tc(const float *h1, const float *gh2, const float *pooled, long long B, int U, int P, const float *g2, const float *be2,
                  const float *mean2, const float *invstd2, const float *W1, const float *W2, float pdrop, unsigned long long seed,
                  const unsigned long long *seed_dev, const uint32_t *dropmask, const float *pb2, float *gpool, float *dW1p, float *dW2,
                  float *db1, float *db2, float *dg2, float *dbe2, cudaStream_t stream)
{
    if (P > BT_P) return XB_ERR_UNSUPPORTED;
    FcBwdArgs a{};
    a.h1 = h1; a.gh2 = gh2; a.pooled = pooled; a.B = B; a.U = U; a.P = P; a.nchunk = (int)((B + FT_ROWS - 1) / FT_ROWS);
    a.g2 = g2; a.be2 = be2; a.mean2 = mean2; a.invstd2 = invstd2; a.W1 = W1; a.W2 = W2; a.pdrop = pdrop; a.seed0 = seed; a.seed_dev = seed_dev; a.dropmask = dropmask;
    a.pb2 = pb2; a.gpool = gpool; a.dW1p = dW1p; a.dW2 = dW2; a.db1 = db1; a.db2 = db2; a.dg2 = dg2; a.dbe2 = dbe2;
    const size_t smem = 2 * (size_t)BT_GT_BYTES + BT_H_BYTES + BT_S_BYTES + 2 * (size_t)BT_W_BYTES;
    const long long total = (long long)U * a.nchunk;
    int grid = xb_num_sms();
    if (total < grid) grid = (int)total;
    cudaError_t e = cudaFuncSetAttribute(bwd_fc1_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        xb_set_error("bwd_fc1_tc: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    bwd_fc1_tc_kernel<<<grid, BT_THREADS, smem, stream>>>(a);
    return xb_check_launch("bwd_fc1_tc");
}
