// conv.cu -- the fused one-hot grouped-conv core and the kernels built on it.
//
// What it replaces in the reference (paths relative to /root/reference):
//   x.repeat(1,U,1) + linears[0..3] = Conv1d(groups=U) -> BatchNorm1d -> exp -> MaxPool1d(7,7)
//       explainn/models/networks.py:98-102,153
//   linears[:3] activation scan cast to float16 ....... explainn/interpretation/interpretation.py:143-147
//   per-unit max and threshold/count PWM extraction .... explainn/interpretation/interpretation.py:187-214
//
// Formulation.  On one-hot DNA the 4xF convolution is a gather-sum: y[s] = sum_k w[code(s+k), k].
// Sequences arrive 2-bit packed, so four adjacent taps form an 8-bit k-mer index; the weights
// (with BatchNorm folded in) are pre-summed into G = ceil(F/4) lookup tables of 256 entries per
// unit (prep.cu).  One CTA owns a tile of 32 units (lane = unit): the G*256*32 floats of tables
// (160 KB for F = 19) are brought into shared memory by TMA bulk copies, a warp walks one sequence
// at a time, every lane adds G table entries per position (conflict-free: all lanes read the same
// k-mer row, consecutive banks) and the pool / exp / threshold logic runs on registers.  Nothing
// of size (B,U,S) is ever written unless the caller asks for the activation map itself.
//
// Canonical arithmetic order (shared with oracle/explainn_oracle.c, bit-exact):
//   t_g = ((a_4g + a_4g+1) + a_4g+2) + a_4g+3 ;  z = (((t_0 + t_1) + ...) + t_{G-1}) + shift[u]
// where a_k = 0 for N columns and for taps k >= F.  Windows that contain an N column take a
// per-tap path that evaluates exactly the same expression.
#include "common.cuh"
#include "../../include/explainn_b200.h"
#include <string.h>

namespace {

constexpr int NW = 8;             // warps per CTA
constexpr int ITER_WIN = 3;       // pool windows per unrolled iteration
constexpr int ITER_POS = ITER_WIN * XB_POOL;   // 21 positions
constexpr int TBL_GROUP_BYTES = 256 * XB_UT * 4;   // 32 KB per k-mer group

struct ConvArgs {
    const uint32_t *codes;     // (B, wc)
    const uint32_t *nmask;     // (B, wm)
    long long B;
    int L, F, U, S;
    int wc, wm;
    int seq_stride;            // words of per-warp staging (codes + mask, padded)
    int mask_off;              // word offset of the mask inside the staging buffer
    const float *tables;       // (tiles, G, 256, 32)
    const float *wq_t;         // (tiles, 4, F, 32) folded weights, lane-minor
    const float *shift;        // (U)
    int nwin;                  // windows to visit per sequence
};

__device__ __forceinline__ float lds_f32(const unsigned char *smem, uint32_t off)
{
    return *reinterpret_cast<const float *>(smem + off);
}

// ----------------------------------------------------------------------------------------------
// Sinks: what happens with the 7 z-values of one pool window.  `acc` excludes shift[u].
// ----------------------------------------------------------------------------------------------

// (1) inference pooling: pooled[b][u][p] = exp(max_window z)
template <int PCH_>
struct PoolSinkT {
    struct Params { float *pooled; int P; };
    static constexpr int PCH = PCH_;
    static __host__ __device__ int smem_bytes(int P) { int ps = (P < PCH ? P : PCH) | 1; return NW * XB_UT * ps * 4; }
    Params p;
    float *stage;
    int ps, lane, nvalid;
    float shift;
    long long out_base, row_stride;   // element offset of pooled[b][u0][0]; distance between units
    int win_base;
    __device__ void init_cta(unsigned char *smem, const Params &pp, int warp, int lane_)
    {
        p = pp;
        ps = (p.P < PCH ? p.P : PCH) | 1;
        stage = reinterpret_cast<float *>(smem) + warp * XB_UT * ps;
        lane = lane_;
    }
    __device__ void begin_tile(int u0, int nvalid_, float shift_, const ConvArgs &) { nvalid = nvalid_; shift = shift_; (void)u0; }
    __device__ void begin_seq(long long b, int u0, const ConvArgs &a)
    {
        row_stride = p.P;
        out_base = (b * a.U + u0) * (long long)p.P;
        win_base = 0;
    }
    __device__ void flush(int upto)
    {
        __syncwarp();
        const int n = upto - win_base;
        for (int r = 0; r < nvalid; ++r)
            for (int c = lane; c < n; c += 32) p.pooled[out_base + r * row_stride + win_base + c] = stage[r * ps + c];
        __syncwarp();
        win_base = upto;
    }
    __device__ __forceinline__ void window(int win, int, const float *acc, const ConvArgs &)
    {
        if (win >= p.P) return;
        if (win - win_base == PCH) flush(win);
        float m = acc[0];
#pragma unroll
        for (int j = 1; j < XB_POOL; ++j) m = fmaxf(m, acc[j]);
        stage[lane * ps + (win - win_base)] = det_expf(__fadd_rn(m, shift));
    }
    __device__ void end_seq(const ConvArgs &) { flush(p.P); }
    __device__ void end_tile(int, const ConvArgs &) {}
};

// (2) training pooling: tables hold x_hat = (y - mu) / sigma; z = gamma * x_hat + beta.
//     Emits pooled = exp(z*) , x_hat* and the first-max index inside the window.
template <int PCH_>
struct PoolTrainSinkT {
    struct Params { float *pooled; float *xhat; unsigned char *idx; const float *gamma; const float *beta; int P; };
    static constexpr int PCH = PCH_;
    static __host__ __device__ int smem_bytes(int) { return NW * XB_UT * (PCH + 1) * 12; }
    Params p;
    float *st_a, *st_x;
    int *st_i;
    int lane, nvalid;
    float shift, gam, bet;
    long long out_base, row_stride;
    int win_base;
    __device__ void init_cta(unsigned char *smem, const Params &pp, int warp, int lane_)
    {
        p = pp;
        lane = lane_;
        float *base = reinterpret_cast<float *>(smem) + warp * XB_UT * (PCH + 1) * 3;
        st_a = base;
        st_x = base + XB_UT * (PCH + 1);
        st_i = reinterpret_cast<int *>(base + 2 * XB_UT * (PCH + 1));
    }
    __device__ void begin_tile(int u0, int nvalid_, float shift_, const ConvArgs &a)
    {
        nvalid = nvalid_;
        shift = shift_;
        const int u = u0 + lane;
        gam = u < a.U ? p.gamma[u] : 1.0f;
        bet = u < a.U ? p.beta[u] : 0.0f;
    }
    __device__ void begin_seq(long long b, int u0, const ConvArgs &a)
    {
        row_stride = p.P;
        out_base = (b * a.U + u0) * (long long)p.P;
        win_base = 0;
    }
    __device__ void flush(int upto)
    {
        __syncwarp();
        const int n = upto - win_base;
        for (int r = 0; r < nvalid; ++r)
            for (int c = lane; c < n; c += 32) {
                const long long o = out_base + r * row_stride + win_base + c;
                p.pooled[o] = st_a[r * (PCH + 1) + c];
                p.xhat[o] = st_x[r * (PCH + 1) + c];
                p.idx[o] = (unsigned char)st_i[r * (PCH + 1) + c];
            }
        __syncwarp();
        win_base = upto;
    }
    __device__ __forceinline__ void window(int win, int, const float *acc, const ConvArgs &)
    {
        if (win >= p.P) return;
        if (win - win_base == PCH) flush(win);
        // first maximum of z = gamma * x_hat + beta (networks.py:100-102; MaxPool1d backward routes the
        // gradient to the first maximal element)
        float xb = __fadd_rn(acc[0], shift);
        float zb = __fmaf_rn(gam, xb, bet);
        int ib = 0;
#pragma unroll
        for (int j = 1; j < XB_POOL; ++j) {
            const float x = __fadd_rn(acc[j], shift);
            const float z = __fmaf_rn(gam, x, bet);
            if (z > zb) { zb = z; xb = x; ib = j; }
        }
        const int c = win - win_base;
        st_a[lane * (PCH + 1) + c] = det_expf(zb);
        st_x[lane * (PCH + 1) + c] = xb;
        st_i[lane * (PCH + 1) + c] = ib;
    }
    __device__ void end_seq(const ConvArgs &) { flush(p.P); }
    __device__ void end_tile(int, const ConvArgs &) {}
};

// (3) activation map: act[b,u,s] = exp(z) for every s < S, float16 or float32 (RAW: z itself, float32)
template <typename OutT, int PCH_, bool RAW = false>
struct ActSink {
    struct Params { OutT *act; };
    static constexpr int PCH = PCH_;                     // windows per flush
    static constexpr int ROW = PCH * XB_POOL + 1;
    static __host__ __device__ int smem_bytes(int) { return NW * XB_UT * ROW * (int)sizeof(OutT); }
    Params p;
    OutT *stage;
    int lane, nvalid;
    float shift;
    long long out_base;
    int win_base;
    __device__ void init_cta(unsigned char *smem, const Params &pp, int warp, int lane_)
    {
        p = pp;
        lane = lane_;
        stage = reinterpret_cast<OutT *>(smem) + warp * XB_UT * ROW;
    }
    __device__ void begin_tile(int, int nvalid_, float shift_, const ConvArgs &) { nvalid = nvalid_; shift = shift_; }
    __device__ void begin_seq(long long b, int u0, const ConvArgs &a) { out_base = (b * a.U + u0) * (long long)a.S; win_base = 0; }
    __device__ void flush(int upto_win, const ConvArgs &a)
    {
        __syncwarp();
        const int s0 = win_base * XB_POOL;
        int n = upto_win * XB_POOL - s0;
        if (s0 + n > a.S) n = a.S - s0;
        for (int r = 0; r < nvalid; ++r)
            for (int c = lane; c < n; c += 32) p.act[out_base + (long long)r * a.S + s0 + c] = stage[r * ROW + c];
        __syncwarp();
        win_base = upto_win;
    }
    __device__ __forceinline__ void window(int win, int, const float *acc, const ConvArgs &a)
    {
        if (win >= a.nwin) return;
        if (win - win_base == PCH) flush(win, a);
        const int c0 = (win - win_base) * XB_POOL;
#pragma unroll
        for (int j = 0; j < XB_POOL; ++j) {
            const float z = __fadd_rn(acc[j], shift);
            stage[lane * ROW + c0 + j] = (OutT)(RAW ? z : det_expf(z));
        }
    }
    __device__ void end_seq(const ConvArgs &a) { flush(a.nwin, a); }
    __device__ void end_tile(int, const ConvArgs &) {}
};

// (4) interpretation pass A: per-unit maximum of z over all sequences and positions s < S
struct MaxSink {
    struct Params { float *zmax_partial; int n_cta; };   // (n_cta, U) initialised to -inf by the host wrapper
    static __host__ __device__ int smem_bytes(int) { return NW * XB_UT * 4; }
    Params p;
    float *red;
    int lane, warp;
    float shift, m;
    __device__ void init_cta(unsigned char *smem, const Params &pp, int warp_, int lane_)
    {
        p = pp;
        lane = lane_;
        warp = warp_;
        red = reinterpret_cast<float *>(smem);
    }
    __device__ void begin_tile(int, int, float shift_, const ConvArgs &) { shift = shift_; m = -INFINITY; }
    __device__ void begin_seq(long long, int, const ConvArgs &) {}
    __device__ __forceinline__ void window(int win, int t, const float *acc, const ConvArgs &a)
    {
        if (win >= a.nwin) return;
#pragma unroll
        for (int j = 0; j < XB_POOL; ++j)
            if (t + j < a.S) m = fmaxf(m, acc[j]);
    }
    __device__ void end_seq(const ConvArgs &) {}
    __device__ void end_tile(int u0, const ConvArgs &a)
    {
        // adding the same shift is monotone, so max(acc) + shift == max(acc + shift) exactly
        red[warp * XB_UT + lane] = __fadd_rn(m, shift);
        __syncthreads();
        if (warp == 0) {
            float v = red[lane];
            for (int w = 1; w < NW; ++w) v = fmaxf(v, red[w * XB_UT + lane]);
            const int u = u0 + lane;
            if (u < a.U) {
                float *dst = p.zmax_partial + (long long)blockIdx.x * a.U + u;
                *dst = fmaxf(*dst, v);
            }
        }
        __syncthreads();
    }
};

// (5) interpretation pass B, first half: WHICH positions have a float16 activation above the float16 threshold
//     (interpretation.py:187,209).  hit  <=>  half(exp(z)) > thr  <=>  z > zhi, or z in (zlo, zhi] and the exact test
//     holds; zlo/zhi bracket the cut-over by 64 ulps on each side (prep.cu: interp_thresholds).
//     The sink only records the hits: one bit per position, W = ceil(S/32) words per (sequence, unit), plus one flag
//     byte per (sequence, unit).  The windows themselves are counted afterwards by hit_count_kernel with bit-parallel
//     population counts, so the cost of pass B does not depend on how many positions hit (round 1 added F shared-memory
//     atomics per hit inside the scan: 50x slower when 5 % of the positions hit).
struct HitSink {
    struct Params { const float *zlo; const float *zhi; const float *thr; uint32_t *hitbits; unsigned char *flags; int W; };
    static __host__ __device__ int smem_bytes(int) { return 0; }
    Params p;
    int lane;
    float shift, zlo, zhi, thr;
    unsigned long long acc64;
    int nb, wi;
    bool seq_any, valid;
    uint32_t *dst;
    long long flag_idx;
    __device__ void init_cta(unsigned char *, const Params &pp, int, int lane_) { p = pp; lane = lane_; }
    __device__ void begin_tile(int u0, int, float shift_, const ConvArgs &a)
    {
        shift = shift_;
        const int u = u0 + lane;
        valid = u < a.U;
        zlo = valid ? p.zlo[u] : INFINITY;
        zhi = valid ? p.zhi[u] : INFINITY;
        thr = valid ? p.thr[u] : INFINITY;
    }
    __device__ void begin_seq(long long b, int u0, const ConvArgs &a)
    {
        acc64 = 0ull; nb = 0; wi = 0; seq_any = false;
        flag_idx = b * a.U + u0 + lane;
        dst = p.hitbits + flag_idx * p.W;
    }
    __device__ __forceinline__ void window(int win, int t, const float *acc, const ConvArgs &a)
    {
        if (win >= a.nwin) return;
        unsigned hitmask = 0;
#pragma unroll
        for (int j = 0; j < XB_POOL; ++j) {
            const float z = __fadd_rn(acc[j], shift);
            bool h = z > zhi;
            if (!h && z > zlo) h = fp16_round(det_expf(z)) > thr;
            if (h && t + j < a.S) hitmask |= 1u << j;
        }
        seq_any |= hitmask != 0;
        acc64 |= (unsigned long long)hitmask << nb;
        nb += XB_POOL;
        if (nb >= 32) {
            if (valid) dst[wi] = (uint32_t)acc64;
            acc64 >>= 32;
            nb -= 32;
            ++wi;
        }
    }
    __device__ void end_seq(const ConvArgs &)
    {
        if (!valid) return;
        for (; wi < p.W; ++wi) {                 // the last, partly filled word and any word past the last window
            dst[wi] = (uint32_t)acc64;
            acc64 >>= 32;
        }
        p.flags[flag_idx] = seq_any ? 1 : 0;
    }
    __device__ void end_tile(int, const ConvArgs &) {}
};

// (6) PWM scoring (explainn/models/networks.py:393-441): per sequence and unit, the maximum or the sum of the conv
//     output over the S valid positions.  `accumulate` folds the call for the reverse strand into the same output
//     (max of both strands / sum of both strands), which is torch.cat((o, o_rev), 2) followed by max / sum.
struct ScoreSink {
    struct Params { float *out; int use_max; int accumulate; };
    static __host__ __device__ int smem_bytes(int) { return 0; }
    Params p;
    int lane;
    float shift, m, sum;
    long long out_idx;
    __device__ void init_cta(unsigned char *, const Params &pp, int, int lane_) { p = pp; lane = lane_; }
    __device__ void begin_tile(int, int, float shift_, const ConvArgs &) { shift = shift_; }
    __device__ void begin_seq(long long b, int u0, const ConvArgs &a)
    {
        m = -INFINITY;
        sum = 0.f;
        out_idx = u0 + lane < a.U ? b * a.U + u0 + lane : -1;
    }
    __device__ __forceinline__ void window(int win, int t, const float *acc, const ConvArgs &a)
    {
        if (win >= a.nwin) return;
#pragma unroll
        for (int j = 0; j < XB_POOL; ++j)
            if (t + j < a.S) {
                const float z = __fadd_rn(acc[j], shift);
                m = fmaxf(m, z);
                sum = __fadd_rn(sum, z);       // positions in increasing order
            }
    }
    __device__ void end_seq(const ConvArgs &)
    {
        if (out_idx < 0) return;
        const float v = p.use_max == 2 ? det_expf(m) : (p.use_max ? m : sum);      // 2: exp(max z), the num_fc = 0 unit output
        if (p.accumulate) p.out[out_idx] = p.use_max ? fmaxf(p.out[out_idx], v) : __fadd_rn(p.out[out_idx], v);
        else p.out[out_idx] = v;
    }
    __device__ void end_tile(int, const ConvArgs &) {}
};

// (7) per (sequence, unit): the maximum activation and the FIRST position that attains it -- torch.max(act, dim=2)
//     (explainn/models/networks.py:228,281), act = exp(z) in float32.  exp is monotone but not injective in float32,
//     so "first maximum of z" is not enough: when a new z exceeds the running maximum by less than 2^-19 (or lies
//     where the float32 exp saturates / goes subnormal) the two activations are compared themselves; beyond that
//     margin the <= 2 ulp det_expf is strictly increasing and the comparison of z decides.
struct ArgmaxSink {
    struct Params { float *amax; int *apos; };
    static __host__ __device__ int smem_bytes(int) { return 0; }
    Params p;
    int lane, pos;
    float shift, m;
    long long out_idx;
    __device__ void init_cta(unsigned char *, const Params &pp, int, int lane_) { p = pp; lane = lane_; }
    __device__ void begin_tile(int, int, float shift_, const ConvArgs &) { shift = shift_; }
    __device__ void begin_seq(long long b, int u0, const ConvArgs &a)
    {
        m = -INFINITY;
        pos = 0;
        out_idx = u0 + lane < a.U ? b * a.U + u0 + lane : -1;
    }
    __device__ __forceinline__ void window(int win, int t, const float *acc, const ConvArgs &a)
    {
        if (win >= a.nwin) return;
#pragma unroll
        for (int j = 0; j < XB_POOL; ++j) {
            const float z = __fadd_rn(acc[j], shift);
            if (t + j < a.S && z > m) {
                bool later = true;
                if (__fsub_rn(z, m) < 1.9073486328125e-06f || z > 88.0f || z < -86.0f) later = det_expf(z) > det_expf(m);
                if (later) pos = t + j;
                m = z;
            }
        }
    }
    __device__ void end_seq(const ConvArgs &)
    {
        if (out_idx < 0) return;
        p.amax[out_idx] = det_expf(m);
        p.apos[out_idx] = pos;
    }
    __device__ void end_tile(int, const ConvArgs &) {}
};

template <class Sink> struct NeedsSeq { static constexpr bool value = false; };

// One position whose window contains an N column: only the 4-tap groups that contain the N are summed tap by tap (the
// table entry's own expression with +0 for the N taps), the other groups keep their table lookup -- bit-identical to the
// oracle.  Deliberately not inlined and not unrolled: the sequence of a warp is shared by its lanes, so the call is
// warp-uniform, and the 21-position unrolled fast path of the caller stays small.
template <int G>
__device__ __noinline__ float scan_slow_position(const uint32_t *cw, const uint32_t *mw, int t, uint32_t nm, int F, const unsigned char *tbl,
                                                 const float *wq, int lane)
{
    float v = 0.0f;
#pragma unroll 1
    for (int g = 0; g < G; ++g) {
        const uint32_t o = 2u * (uint32_t)(t + 4 * g);
        const uint32_t km = __funnelshift_r(cw[o >> 5], cw[(o >> 5) + 1], o & 31u) & 0xFFu;
        const uint32_t nib = (nm >> (4 * g)) & 15u;
        float tg;
        if (nib == 0u) {
            tg = lds_f32(tbl, (uint32_t)lane * 4u + (km << 7) + (uint32_t)g * TBL_GROUP_BYTES);
        } else {
            tg = 0.0f;
#pragma unroll 1
            for (int jj = 0; jj < 4; ++jj) {
                const int k = 4 * g + jj;
                float w = 0.0f;
                if (k < F && !((nib >> jj) & 1u)) w = wq[(((km >> (2 * jj)) & 3u) * F + k) * XB_UT + lane];
                tg = jj == 0 ? w : __fadd_rn(tg, w);
            }
        }
        v = g == 0 ? tg : __fadd_rn(v, tg);
    }
    return v;
}

// ----------------------------------------------------------------------------------------------
// The scan kernel.  grid = one CTA per SM; the flattened (tile, sequence) space is split evenly.
// ----------------------------------------------------------------------------------------------
template <int G, class Sink>
__global__ void __launch_bounds__(NW * 32, 1) conv_scan_kernel(const ConvArgs a, const typename Sink::Params sp)
{
    constexpr int SPAN = 4 * (G - 1);
    constexpr int NA = ITER_POS + SPAN;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t bar;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned char *tbl = smem;                                         // G * 32 KB
    float *wq = reinterpret_cast<float *>(smem + G * TBL_GROUP_BYTES); // 4*F*32 floats
    uint32_t *seqbuf = reinterpret_cast<uint32_t *>(wq + 4 * a.F * XB_UT) + warp * a.seq_stride;
    unsigned char *sink_smem = reinterpret_cast<unsigned char *>(reinterpret_cast<uint32_t *>(wq + 4 * a.F * XB_UT) + NW * a.seq_stride);

    Sink sink;
    sink.init_cta(sink_smem, sp, warp, lane);

    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    // zero the staging buffer once: words past the sequence stay zero for the over-read taps
    for (int i = lane; i < a.seq_stride; i += 32) seqbuf[i] = 0;
    __syncthreads();

    const int n_tiles = (a.U + XB_UT - 1) / XB_UT;
    const long long total = (long long)n_tiles * a.B;
    long long it0 = total * blockIdx.x / gridDim.x;
    const long long it1 = total * (blockIdx.x + 1) / gridDim.x;
    uint32_t parity = 0;
    const uint32_t lane_base = lane * 4;
    const uint32_t *cw = seqbuf, *mw = seqbuf + a.mask_off;
    const int nit = (a.nwin + ITER_WIN - 1) / ITER_WIN;

    while (it0 < it1) {
        const int tile = (int)(it0 / a.B);
        const long long b_lo = it0 % a.B;
        long long b_hi = b_lo + (it1 - it0);
        if (b_hi > a.B) b_hi = a.B;
        const int u0 = tile * XB_UT;
        const int nvalid = min(XB_UT, a.U - u0);

        // ---- bring this tile's tables in (TMA bulk copies, one per k-mer group) ----
        if (threadIdx.x == 0) {
            mbar_arrive_expect_tx(&bar, G * TBL_GROUP_BYTES);
            const unsigned char *src = reinterpret_cast<const unsigned char *>(a.tables) + (size_t)tile * G * TBL_GROUP_BYTES;
#pragma unroll
            for (int g = 0; g < G; ++g) tma_bulk_g2s(tbl + g * TBL_GROUP_BYTES, src + (size_t)g * TBL_GROUP_BYTES, TBL_GROUP_BYTES, &bar);
        }
        for (int i = threadIdx.x; i < 4 * a.F * XB_UT; i += blockDim.x) wq[i] = a.wq_t[(size_t)tile * 4 * a.F * XB_UT + i];
        const float shift = (u0 + lane < a.U) ? a.shift[u0 + lane] : 0.0f;
        sink.begin_tile(u0, nvalid, shift, a);
        mbar_wait(&bar, parity);
        parity ^= 1;
        __syncthreads();

        // ---- sequences of this segment, one per warp ----
        long long b = b_lo + warp;
        uint32_t pre[6];
        auto prefetch = [&](long long bb) {
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const int w = lane + 32 * i;
                pre[i] = 0;
                if (bb < b_hi) {
                    if (w < a.wc) pre[i] = a.codes[bb * a.wc + w];
                    else if (w < a.wc + a.wm) pre[i] = a.nmask[bb * a.wm + (w - a.wc)];
                }
            }
        };
        prefetch(b);
        for (; b < b_hi; b += NW) {
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const int w = lane + 32 * i;
                if (w < a.wc) seqbuf[w] = pre[i];
                else if (w < a.wc + a.wm) seqbuf[a.mask_off + (w - a.wc)] = pre[i];
            }
            prefetch(b + NW);
            __syncwarp();
            if constexpr (NeedsSeq<Sink>::value) sink.set_seq(cw, mw);
            sink.begin_seq(b, u0, a);

            uint32_t addr[NA];
            // prologue: k-mers at positions 0 .. SPAN-1
            {
                const uint32_t w0 = cw[0], w1 = cw[1];
#pragma unroll
                for (int j = 0; j < SPAN; ++j) {
                    uint32_t km;
                    if (2 * j + 8 <= 32) km = w0 >> (2 * j);
                    else if (2 * j < 32) km = __funnelshift_r(w0, w1, 2 * j);
                    else km = w1 >> (2 * j - 32);
                    addr[j] = lane_base + ((km & 0xFFu) << 7);
                }
            }
            for (int it = 0; it < nit; ++it) {
                const int t0 = it * ITER_POS;
                // new k-mers at positions t0+SPAN .. t0+SPAN+20
                {
                    const uint32_t o = 2u * (uint32_t)(t0 + SPAN);
                    const uint32_t wi = o >> 5, sh = o & 31u;
                    const uint32_t w0 = cw[wi], w1 = cw[wi + 1], w2 = cw[wi + 2];
                    const uint32_t A0 = __funnelshift_r(w0, w1, sh), A1 = __funnelshift_r(w1, w2, sh);
#pragma unroll
                    for (int j = 0; j < ITER_POS; ++j) {
                        uint32_t km;
                        if (2 * j + 8 <= 32) km = A0 >> (2 * j);
                        else if (2 * j < 32) km = __funnelshift_r(A0, A1, 2 * j);
                        else km = A1 >> (2 * j - 32);
                        addr[SPAN + j] = lane_base + ((km & 0xFFu) << 7);
                    }
                }
                // N flags of the positions under this iteration's footprint (bit p = position t0 + p).  All lanes of the
                // warp work on the same sequence, so every test on them is warp-uniform.
                unsigned long long MM;
                {
                    const uint32_t wi = (uint32_t)t0 >> 5, sh = (uint32_t)t0 & 31u;
                    const uint32_t m0 = mw[wi], m1 = mw[wi + 1], m2 = mw[wi + 2];
                    MM = (unsigned long long)__funnelshift_r(m0, m1, sh) | ((unsigned long long)__funnelshift_r(m1, m2, sh) << 32);
                }
                const uint32_t foot = a.F >= 32 ? 0xFFFFFFFFu : ((1u << a.F) - 1u);     // taps >= F are +0 in the tables
                float acc[ITER_POS];
#pragma unroll
                for (int j = 0; j < ITER_POS; ++j) {
                    const uint32_t nm = (uint32_t)(MM >> j) & foot;
                    float v;
                    if (nm == 0u) {
                        v = lds_f32(tbl, addr[j]);
#pragma unroll
                        for (int g = 1; g < G; ++g) v = __fadd_rn(v, lds_f32(tbl, addr[j + 4 * g] + g * TBL_GROUP_BYTES));
                    } else {
                        // an N column under this window (rare): evaluated out of line, so that the unrolled fast path
                        // stays small enough for the instruction cache
                        v = scan_slow_position<G>(cw, mw, t0 + j, nm, a.F, tbl, wq, lane);
                    }
                    acc[j] = v;
                }
#pragma unroll
                for (int w = 0; w < ITER_WIN; ++w) sink.window(it * ITER_WIN + w, t0 + w * XB_POOL, acc + w * XB_POOL, a);
#pragma unroll
                for (int j = 0; j < SPAN; ++j) addr[j] = addr[j + ITER_POS];
            }
            sink.end_seq(a);
        }
        __syncthreads();
        sink.end_tile(u0, a);
        it0 += (b_hi - b_lo);
    }
}

// ----------------------------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------------------------
struct Geometry {
    int G, wc, wm, seq_stride, mask_off, nwin_pool, nwin_all, S, P;
};

int make_geometry(int L, int F, Geometry *g)
{
    if (F < 1 || F > 4 * XB_MAX_G || L < F) return XB_ERR_UNSUPPORTED;
    g->G = (F + 3) / 4;
    g->wc = xb_words_codes(L);
    g->wm = xb_words_mask(L);
    g->S = L - F + 1;
    g->P = g->S >= XB_POOL ? (g->S - XB_POOL) / XB_POOL + 1 : 0;
    g->nwin_pool = g->P;
    g->nwin_all = (g->S + XB_POOL - 1) / XB_POOL;
    // positions touched by the unrolled loop: nit*21 + 4G + slack; +3 words for the funnel-shift look-ahead
    const int nit = (g->nwin_all + ITER_WIN - 1) / ITER_WIN;
    const int maxpos = nit * ITER_POS + 4 * g->G + 8;
    const int cw_pad = (2 * maxpos + 31) / 32 + 3;
    const int mw_pad = (maxpos + 31) / 32 + 3;
    g->mask_off = cw_pad > g->wc ? cw_pad : g->wc;
    g->seq_stride = g->mask_off + (mw_pad > g->wm ? mw_pad : g->wm);
    if (g->wc + g->wm > 6 * 32) return XB_ERR_UNSUPPORTED;   // prefetch registers: L <= 2048
    return XB_OK;
}

template <int G, class Sink>
int launch_scan(const ConvArgs &a, const typename Sink::Params &sp, int sink_smem, cudaStream_t stream)
{
    const size_t smem = (size_t)G * TBL_GROUP_BYTES + (size_t)4 * a.F * XB_UT * 4 + (size_t)NW * a.seq_stride * 4 + sink_smem;
    if (smem > 227 * 1024) {
        xb_set_error("conv scan: %zu bytes of shared memory needed (filter_size=%d, L=%d) exceed 227 KB", smem, a.F, a.L);
        return XB_ERR_UNSUPPORTED;
    }
    auto kern = conv_scan_kernel<G, Sink>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) {
        xb_set_error("cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    const int n_tiles = (a.U + XB_UT - 1) / XB_UT;
    long long total = (long long)n_tiles * a.B;
    int grid = xb_num_sms();
    if (total < grid) grid = (int)total;
    if (grid < 1) return XB_OK;
    kern<<<grid, NW * 32, smem, stream>>>(a, sp);
    return xb_check_launch("conv_scan_kernel");
}

// Sink: staging sized for G <= 5 (160 KB of tables); Small: the same sink with a short staging buffer
// for G = 6 (192 KB of tables leave ~21 KB).
template <class Sink, class Small>
int dispatch_scan(int G, const ConvArgs &a, const typename Sink::Params &sp, int arg, cudaStream_t stream)
{
    static_assert(sizeof(typename Sink::Params) == sizeof(typename Small::Params), "sink params differ");
    const int sink_smem = Sink::smem_bytes(arg);
    switch (G) {
    case 1: return launch_scan<1, Sink>(a, sp, sink_smem, stream);
    case 2: return launch_scan<2, Sink>(a, sp, sink_smem, stream);
    case 3: return launch_scan<3, Sink>(a, sp, sink_smem, stream);
    case 4: return launch_scan<4, Sink>(a, sp, sink_smem, stream);
    case 5: return launch_scan<5, Sink>(a, sp, sink_smem, stream);
    case 6: {
        typename Small::Params sq;
        memcpy(&sq, &sp, sizeof(sq));
        return launch_scan<6, Small>(a, sq, Small::smem_bytes(arg), stream);
    }
    }
    xb_set_error("conv scan: filter_size %d not supported (max %d)", a.F, 4 * XB_MAX_G);
    return XB_ERR_UNSUPPORTED;
}

int fill_args(ConvArgs *a, Geometry *g, const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
              const float *tables, const float *wq_t, const float *shift)
{
    if (!codes || !nmask || !tables || !wq_t || !shift) {
        xb_set_error("conv scan: null pointer");
        return XB_ERR_ARG;
    }
    if (B < 0 || U < 1) {
        xb_set_error("conv scan: bad sizes B=%lld U=%d", B, U);
        return XB_ERR_ARG;
    }
    int rc = make_geometry(L, F, g);
    if (rc != XB_OK) {
        xb_set_error("conv scan: unsupported geometry L=%d filter_size=%d (need filter_size <= %d, L <= 2048)", L, F, 4 * XB_MAX_G);
        return rc;
    }
    a->codes = codes; a->nmask = nmask; a->B = B; a->L = L; a->F = F; a->U = U; a->S = g->S;
    a->wc = g->wc; a->wm = g->wm; a->seq_stride = g->seq_stride; a->mask_off = g->mask_off;
    a->tables = tables; a->wq_t = wq_t; a->shift = shift; a->nwin = 0;
    return XB_OK;
}

}  // namespace

extern "C" int explainn_conv_pool_fwd(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                      const float *tables, const float *wq_t, const float *shift, float *pooled,
                                      cudaStream_t stream)
{
    ConvArgs a; Geometry g;
    int rc = fill_args(&a, &g, codes, nmask, B, L, U, F, tables, wq_t, shift);
    if (rc != XB_OK) return rc;
    XB_REQUIRE(pooled, "conv_pool_fwd: null output");
    if (B == 0 || g.P == 0) return XB_OK;
    a.nwin = g.nwin_pool;
    PoolSinkT<36>::Params sp{pooled, g.P};
    return dispatch_scan<PoolSinkT<36>, PoolSinkT<12>>(g.G, a, sp, g.P, stream);
}

extern "C" int explainn_conv_pool_fwd_train(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                            const float *tables, const float *wq_t, const float *shift,
                                            const float *gamma, const float *beta, float *pooled, float *xhat,
                                            unsigned char *idx, cudaStream_t stream)
{
    ConvArgs a; Geometry g;
    int rc = fill_args(&a, &g, codes, nmask, B, L, U, F, tables, wq_t, shift);
    if (rc != XB_OK) return rc;
    XB_REQUIRE(pooled && xhat && idx && gamma && beta, "conv_pool_fwd_train: null pointer");
    if (B == 0 || g.P == 0) return XB_OK;
    a.nwin = g.nwin_pool;
    PoolTrainSinkT<12>::Params sp{pooled, xhat, idx, gamma, beta, g.P};
    return dispatch_scan<PoolTrainSinkT<12>, PoolTrainSinkT<3>>(g.G, a, sp, g.P, stream);
}

extern "C" int explainn_unit_activations(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                         const float *tables, const float *wq_t, const float *shift, void *act,
                                         int act_dtype, cudaStream_t stream)
{
    ConvArgs a; Geometry g;
    int rc = fill_args(&a, &g, codes, nmask, B, L, U, F, tables, wq_t, shift);
    if (rc != XB_OK) return rc;
    XB_REQUIRE(act, "unit_activations: null output");
    if (B == 0) return XB_OK;
    a.nwin = g.nwin_all;
    if (act_dtype == EXPLAINN_DTYPE_F16) {
        ActSink<__half, 6>::Params sp{(__half *)act};
        return dispatch_scan<ActSink<__half, 6>, ActSink<__half, 3>>(g.G, a, sp, 0, stream);
    } else if (act_dtype == EXPLAINN_DTYPE_F32) {
        ActSink<float, 6>::Params sp{(float *)act};
        return dispatch_scan<ActSink<float, 6>, ActSink<float, 1>>(g.G, a, sp, 0, stream);
    } else if (act_dtype == EXPLAINN_DTYPE_F32_PREACT) {
        ActSink<float, 6, true>::Params sp{(float *)act};
        return dispatch_scan<ActSink<float, 6, true>, ActSink<float, 1, true>>(g.G, a, sp, 0, stream);
    }
    xb_set_error("unit_activations: unsupported dtype %d", act_dtype);
    return XB_ERR_UNSUPPORTED;
}

extern "C" long long explainn_interp_unit_max_scratch_floats(int U) { return (long long)xb_num_sms() * U; }

__global__ void fill_f32_kernel(float *p, long long n, float v)
{
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}
__global__ void reduce_zmax_kernel(const float *partial, int n_cta, int U, float *zmax, int accumulate)
{
    const int u = blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= U) return;
    float m = accumulate ? zmax[u] : -INFINITY;
    for (int c = 0; c < n_cta; ++c) m = fmaxf(m, partial[(long long)c * U + u]);
    zmax[u] = m;
}

extern "C" int explainn_interp_unit_max(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                        const float *tables, const float *wq_t, const float *shift, float *scratch,
                                        float *zmax, int accumulate, cudaStream_t stream)
{
    ConvArgs a; Geometry g;
    int rc = fill_args(&a, &g, codes, nmask, B, L, U, F, tables, wq_t, shift);
    if (rc != XB_OK) return rc;
    XB_REQUIRE(scratch && zmax, "interp_unit_max: null pointer");
    const int n_cta = xb_num_sms();
    const long long n = (long long)n_cta * U;
    fill_f32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(scratch, n, -INFINITY);
    a.nwin = g.nwin_all;
    if (B > 0) {
        MaxSink::Params sp{scratch, n_cta};
        rc = dispatch_scan<MaxSink, MaxSink>(g.G, a, sp, 0, stream);
        if (rc != XB_OK) return rc;
    }
    reduce_zmax_kernel<<<(U + 127) / 128, 128, 0, stream>>>(scratch, n_cta, U, zmax, accumulate);
    return xb_check_launch("interp_unit_max");
}

extern "C" int explainn_pwm_score(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                  const float *tables, const float *wq_t, const float *shift, int use_max, int accumulate,
                                  float *out, cudaStream_t stream)
{
    ConvArgs a; Geometry g;
    int rc = fill_args(&a, &g, codes, nmask, B, L, U, F, tables, wq_t, shift);
    if (rc != XB_OK) return rc;
    XB_REQUIRE(out, "pwm_score: null output");
    if (B == 0) return XB_OK;
    a.nwin = g.nwin_all;
    ScoreSink::Params sp{out, use_max, accumulate};
    return dispatch_scan<ScoreSink, ScoreSink>(g.G, a, sp, 0, stream);
}

extern "C" int explainn_interp_seq_argmax(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                          const float *tables, const float *wq_t, const float *shift, float *amax,
                                          int *apos, cudaStream_t stream)
{
    ConvArgs a; Geometry g;
    int rc = fill_args(&a, &g, codes, nmask, B, L, U, F, tables, wq_t, shift);
    if (rc != XB_OK) return rc;
    XB_REQUIRE(amax && apos, "interp_seq_argmax: null output");
    if (B == 0) return XB_OK;
    a.nwin = g.nwin_all;
    ArgmaxSink::Params sp{amax, apos};
    return dispatch_scan<ArgmaxSink, ArgmaxSink>(g.G, a, sp, 0, stream);
}

// Base masks of a batch: bmask[n][c][i] bit p = 1 iff base(32 i + p) == c (N columns and positions >= L: no bit),
// i <= W.  One warp per sequence.
__global__ void __launch_bounds__(256) basemask_kernel(const uint32_t *__restrict__ codes, const uint32_t *__restrict__ nmask, long long B,
                                                       int L, int W, uint32_t *__restrict__ bmask)
{
    const int lane = threadIdx.x & 31;
    const long long n = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (n >= B) return;
    const int wc = xb_words_codes(L), wm = xb_words_mask(L);
    for (int i = 0; i <= W; ++i) {
        const int pos = 32 * i + lane;
        int c = 4;
        if (pos < L && !((nmask[n * wm + (pos >> 5)] >> (pos & 31)) & 1u)) c = (codes[n * wc + (pos >> 4)] >> (2 * (pos & 15))) & 3;
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
            const unsigned m = __ballot_sync(0xffffffffu, c == cc);
            if (lane == cc) bmask[(n * 4 + cc) * (W + 1) + i] = m;
        }
    }
}

// Pass B, second half: counts[u][c][k] += #{hit positions s of unit u : base(s + k) == c} = popc(H_u & (B_c >> k)).
// A warp owns 8 units and a range of sequences; lane k < F owns tap k (its base masks shifted by k live in registers,
// prepared once per sequence and 8-word chunk), lane 31 carries an all-ones mask and so counts the hits themselves.
// The 32 counters per lane stay in registers over all the warp's sequences: no atomics until the very end.
__global__ void __launch_bounds__(128) hit_count_kernel(const uint32_t *__restrict__ hitbits, const uint32_t *__restrict__ bmask,
                                                        const unsigned char *__restrict__ flags, unsigned char *__restrict__ seq_hit_out,
                                                        long long B, int U, int W, int F,
                                                        unsigned long long *__restrict__ counts, unsigned long long *__restrict__ hits)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int u0 = blockIdx.x * 8;
    const long long n0 = B * blockIdx.y / gridDim.y, n1 = B * (blockIdx.y + 1) / gridDim.y;
    unsigned acc[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[j][c] = 0u;
    const bool tap = lane < F, ones = lane == 31;
    for (long long n = n0 + warp; n < n1; n += 4) {
        const uint32_t *hb = hitbits + (n * U + u0) * W;
        unsigned f = 0;
        if (lane < 8 && u0 + lane < U) {
            if (flags) f = flags[n * U + u0 + lane];
            else {                                   // no flag bytes (tensor-core pass B): any bit of the row
                for (int i = 0; i < W; ++i) f |= hb[lane * W + i];
                if (seq_hit_out) seq_hit_out[n * U + u0 + lane] = f ? 1 : 0;
            }
        }
        const unsigned any = __ballot_sync(0xffffffffu, f != 0);
        if (!any) continue;
        const uint32_t *bm = bmask + n * 4 * (W + 1);
        for (int w0 = 0; w0 < W; w0 += 8) {
            unsigned m[4][8];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                unsigned prev = w0 < W ? bm[c * (W + 1) + w0] : 0u;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const unsigned next = (w0 + i < W) ? bm[c * (W + 1) + w0 + i + 1] : 0u;
                    m[c][i] = (w0 + i < W) ? (tap ? __funnelshift_r(prev, next, lane) : ((ones && c == 0) ? 0xFFFFFFFFu : 0u)) : 0u;
                    prev = next;
                }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (!((any >> j) & 1u)) continue;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const unsigned h = (w0 + i < W) ? hb[j * W + w0 + i] : 0u;
#pragma unroll
                    for (int c = 0; c < 4; ++c) acc[j][c] += __popc(h & m[c][i]);
                }
            }
        }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int u = u0 + j;
        if (u >= U) continue;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            if (!acc[j][c]) continue;
            if (tap) atomicAdd(&counts[((long long)u * 4 + c) * F + lane], (unsigned long long)acc[j][c]);
            else if (ones && c == 0) atomicAdd(&hits[u], (unsigned long long)acc[j][c]);
        }
    }
}

// shared with interp_tc.cu: base masks of the batch + window counting from hit bits.  flags == NULL: the "any hit" test
// reads the bit rows themselves and, when seq_hit_out is given, writes the per-(sequence, unit) flags there.
int xb_hit_count(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F, const uint32_t *bits, uint32_t *bmask,
                 const unsigned char *flags, unsigned char *seq_hit_out, unsigned long long *counts, unsigned long long *hits,
                 cudaStream_t stream)
{
    const int W = (L - F + 1 + 31) / 32;
    basemask_kernel<<<(unsigned)((B + 7) / 8), 256, 0, stream>>>(codes, nmask, B, L, W, bmask);
    const int groups = (U + 7) / 8;
    long long split = (8LL * xb_num_sms() + groups - 1) / groups;
    if (split > (B + 3) / 4) split = (B + 3) / 4;
    if (split > 65535) split = 65535;
    if (split < 1) split = 1;
    hit_count_kernel<<<dim3((unsigned)groups, (unsigned)split), 128, 0, stream>>>(bits, bmask, flags, seq_hit_out, B, U, W, F, counts, hits);
    return xb_check_launch("interp_pwm_count");
}

long long xb_pwm_count_layout(long long B, int L, int U, int F, long long *off_bmask, long long *off_bits);

static long long pwm_count_layout(long long B, int L, int U, int F, long long *off_bmask, long long *off_bits)
{
    const int S = L - F + 1, W = (S + 31) / 32;
    long long o = (B * U + 255) / 256 * 256;            // flag bytes
    *off_bmask = o;
    o += (B * 4 * (W + 1) * 4 + 255) / 256 * 256;
    *off_bits = o;
    o += B * U * W * 4;
    return o;
}

long long xb_pwm_count_layout(long long B, int L, int U, int F, long long *off_bmask, long long *off_bits)
{
    return pwm_count_layout(B, L, U, F, off_bmask, off_bits);
}

extern "C" long long explainn_interp_pwm_count_scratch_bytes(long long B, int L, int U, int F)
{
    if (B < 0 || U < 1 || F < 1 || L < F) return 0;
    long long a, b;
    return pwm_count_layout(B, L, U, F, &a, &b);
}

extern "C" int explainn_interp_pwm_count(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                         const float *tables, const float *wq_t, const float *shift, const float *zlo,
                                         const float *zhi, const float *thr, unsigned long long *counts,
                                         unsigned long long *hits, unsigned char *seq_hit, void *scratch, cudaStream_t stream)
{
    ConvArgs a; Geometry g;
    int rc = fill_args(&a, &g, codes, nmask, B, L, U, F, tables, wq_t, shift);
    if (rc != XB_OK) return rc;
    XB_REQUIRE(zlo && zhi && thr && counts && hits && scratch, "interp_pwm_count: null pointer");
    if (B == 0) return XB_OK;
    a.nwin = g.nwin_all;
    const int W = (g.S + 31) / 32;
    long long off_bmask, off_bits;
    pwm_count_layout(B, L, U, F, &off_bmask, &off_bits);
    unsigned char *base = static_cast<unsigned char *>(scratch);
    unsigned char *flags = seq_hit ? seq_hit : base;
    uint32_t *bmask = reinterpret_cast<uint32_t *>(base + off_bmask), *bits = reinterpret_cast<uint32_t *>(base + off_bits);
    HitSink::Params sp{zlo, zhi, thr, bits, flags, W};
    rc = dispatch_scan<HitSink, HitSink>(g.G, a, sp, 0, stream);
    if (rc != XB_OK) return rc;
    return xb_hit_count(codes, nmask, B, L, U, F, bits, bmask, flags, nullptr, counts, hits, stream);
}

// ----------------------------------------------------------------------------------------------
// explainn_interp_unit_max_noalloc -- pass A of the PWM extraction for callers that cannot allocate.
//
// Some hosts of this library may not call cudaMalloc on the path where the scan runs: a serving process whose
// device memory was fully partitioned at start-up, a plugin running inside another framework's stream-ordered
// region, or a worker thread that must never hit the allocator lock.  For them the per-CTA partial maxima that
// explainn_interp_unit_max() keeps in caller-provided scratch live here, in a fixed array that is part of the
// library image.  Every GPU that loads the library holds these 640 KB until the process exits.
//
// Limits: U <= XB_NOALLOC_MAX_U (1024) units and a GPU with at most XB_NOALLOC_MAX_CTAS (160) SMs; larger
// requests are refused with EXPLAINN_ERR_LIMIT.
//
// The array's name is only a host-side handle: host code must not pass it to a kernel.  The device address is
// obtained once per device with cudaGetSymbolAddress (with that device current) and cached.  That set-up
// happens inside the first call on a device, so make that first call outside of any stream capture.
//
// All calls on one GPU share the array, and it keeps whatever the previous call left in it; each call first
// resets the part it uses to -inf on `stream`.  Two calls that overlap on the GPU (different streams, no
// ordering between them) race on the array -- serialise them.
// ----------------------------------------------------------------------------------------------
#define XB_NOALLOC_MAX_U 1024
#define XB_NOALLOC_MAX_CTAS 160
#define XB_NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) float g_unit_max_scratch[XB_NOALLOC_MAX_CTAS * XB_NOALLOC_MAX_U];

static float *g_unit_max_scratch_addr[XB_NOALLOC_MAX_DEVICES];   // device address, per device ordinal (0 = not set up)

static int noalloc_scratch_for_current_device(float **out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) {
        xb_set_error("interp_unit_max_noalloc: cudaGetDevice: %s", cudaGetErrorString(e));
        return XB_ERR_CUDA;
    }
    if (dev < 0 || dev >= XB_NOALLOC_MAX_DEVICES) {
        xb_set_error("interp_unit_max_noalloc: device ordinal %d outside the supported range [0, %d)", dev, XB_NOALLOC_MAX_DEVICES);
        return XB_ERR_LIMIT;
    }
    float *p = __atomic_load_n(&g_unit_max_scratch_addr[dev], __ATOMIC_ACQUIRE);
    if (p == nullptr) {
        // once per device, with that device current (two racing threads would store the same address)
        e = cudaGetSymbolAddress((void **)&p, g_unit_max_scratch);
        if (e != cudaSuccess || p == nullptr) {
            xb_set_error("interp_unit_max_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return XB_ERR_CUDA;
        }
        __atomic_store_n(&g_unit_max_scratch_addr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return XB_OK;
}

extern "C" int explainn_interp_unit_max_noalloc(const uint32_t *codes, const uint32_t *nmask, long long B, int L, int U, int F,
                                                const float *tables, const float *wq_t, const float *shift, float *zmax,
                                                int accumulate, cudaStream_t stream)
{
    ConvArgs a; Geometry g;
    int rc = fill_args(&a, &g, codes, nmask, B, L, U, F, tables, wq_t, shift);
    if (rc != XB_OK) return rc;
    XB_REQUIRE(zmax, "interp_unit_max_noalloc: null output");
    const int n_cta = xb_num_sms();
    if (U > XB_NOALLOC_MAX_U || n_cta > XB_NOALLOC_MAX_CTAS) {
        xb_set_error("interp_unit_max_noalloc: %d units on %d SMs exceed the built-in scratch (at most %d units, %d SMs); "
                     "use explainn_interp_unit_max with caller scratch",
                     U, n_cta, XB_NOALLOC_MAX_U, XB_NOALLOC_MAX_CTAS);
        return XB_ERR_LIMIT;
    }
    float *scratch = nullptr;
    rc = noalloc_scratch_for_current_device(&scratch);
    if (rc != XB_OK) return rc;
    const long long n = (long long)n_cta * U;
    fill_f32_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(scratch, n, -INFINITY);   // last call's maxima go away
    a.nwin = g.nwin_all;
    if (B > 0) {
        MaxSink::Params sp{scratch, n_cta};
        rc = dispatch_scan<MaxSink, MaxSink>(g.G, a, sp, 0, stream);
        if (rc != XB_OK) return rc;
    }
    reduce_zmax_kernel<<<(U + 127) / 128, 128, 0, stream>>>(scratch, n_cta, U, zmax, accumulate);
    return xb_check_launch("interp_unit_max_noalloc");
}
