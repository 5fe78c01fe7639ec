/*
 * explainn_oracle.c -- TEST INFRASTRUCTURE ONLY (never imported by the product path).
 *
 * CPU restatement of the ExplaiNN hot path in the CANONICAL arithmetic order that the
 * CUDA kernels in explainn_b200/csrc/ also follow, so that the two can be compared
 * bit for bit.  The algorithm restated is the reference's (paths relative to
 * /root/reference):
 *   - one-hot encoding contract ........ explainn/utils/tools.py:136-181 (A,C,G,T rows, N = zero column)
 *   - grouped Conv1d (valid, stride 1) .. explainn/models/networks.py:98-99 (+ x.repeat, :153)
 *   - BatchNorm1d eval affine ........... explainn/models/networks.py:100
 *   - exp activation .................... explainn/models/networks.py:21-29
 *   - MaxPool1d(k, stride) floor mode ... explainn/models/networks.py:102
 *   - fp16 activation scan .............. explainn/interpretation/interpretation.py:126-149
 *   - threshold/count PWM extraction .... explainn/interpretation/interpretation.py:176-227
 *   - per-sequence max / first argmax ... torch.max(act, dim=2) as in explainn/models/networks.py:228,281
 *
 * Parity pinning: the reference ships no known-answer tests.  This
 * restatement is pinned against outputs of the reference itself, generated in the build
 * container by tests/golden/make_golden.py and committed under tests/golden/ ; see
 * tests/test_oracle_golden.py.
 *
 * Canonical order (all fp32, round-to-nearest-even, NO fused multiply-add except where
 * fmaf is written explicitly):
 *   scale[u] = gamma[u] / sqrtf(var[u] + eps)
 *   shift[u] = (bias[u] - mean[u]) * scale[u] + beta[u]
 *   wq[u,c,k] = w[u,c,k] * scale[u]                      (taps k >= F are +0.0f)
 *   t_g      = ((a_{4g} + a_{4g+1}) + a_{4g+2}) + a_{4g+3},   a_k = N(s+k) ? +0.0f : wq[u, code(s+k), k]
 *   z[b,u,s] = (((t_0 + t_1) + t_2) + ... ) + shift[u]
 *   act      = det_expf(z)                                (deterministic exp below)
 *   pooled   = det_expf(max_window z), argmax = first index of the maximal z
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

/* ---- deterministic expf: identical operation sequence to csrc/common.cuh::det_expf ---- */
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

ORC_API float orc_expf(float x)
{
    if (x != x) return x;
    if (x > 88.7228394f) return INFINITY;
    if (x < -104.0f) return 0.0f;
    float n = rintf(x * 1.44269502f);
    float r = fmaf(n, -0.693145751953125f, x);
    r = fmaf(n, -1.42860676533018e-06f, r);
    float p = 1.9875691500e-4f;
    p = fmaf(p, r, 1.3981999507e-3f);
    p = fmaf(p, r, 8.3334519073e-3f);
    p = fmaf(p, r, 4.1665795894e-2f);
    p = fmaf(p, r, 1.6666665459e-1f);
    p = fmaf(p, r, 5.0000001201e-1f);
    float rr = r * r;
    p = fmaf(p, rr, r);
    p = p + 1.0f;
    int ni = (int)n;
    int n1 = ni >> 1;
    int n2 = ni - n1;
    float s1 = u2f((uint32_t)(n1 + 127) << 23);
    float s2 = u2f((uint32_t)(n2 + 127) << 23);
    return (p * s1) * s2;
}

ORC_API void orc_expf_vec(const float *x, float *y, long n)
{
#pragma omp parallel for schedule(static)
    for (long i = 0; i < n; ++i) y[i] = orc_expf(x[i]);
}

/* ---- packing: 2-bit codes (16 bases / uint32, base i at bits 2*(i%16)) + N mask (32 / uint32) ---- */
ORC_API int orc_words_codes(int L) { return (L + 15) / 16; }
ORC_API int orc_words_mask(int L) { return (L + 31) / 32; }

/* x: (B,4,L) fp32 with batch stride bstride (elements).  Returns number of invalid columns
 * (neither exactly one 1.0 with three 0.0, nor all 0.0).  Bits at positions >= L stay zero. */
ORC_API long orc_pack_onehot(const float *x, long B, int L, long bstride,
                             uint32_t *codes, uint32_t *nmask)
{
    int wc = orc_words_codes(L), wm = orc_words_mask(L);
    long bad = 0;
    memset(codes, 0, (size_t)B * wc * 4);
    memset(nmask, 0, (size_t)B * wm * 4);
    for (long b = 0; b < B; ++b) {
        const float *xb = x + b * bstride;
        for (int i = 0; i < L; ++i) {
            int ones = 0, zeros = 0, code = 0;
            for (int c = 0; c < 4; ++c) {
                float v = xb[(long)c * L + i];
                if (v == 1.0f) { ones++; code = c; }
                else if (v == 0.0f) zeros++;
            }
            if (ones == 1 && zeros == 3) {
                codes[b * wc + i / 16] |= (uint32_t)code << (2 * (i % 16));
            } else {
                nmask[b * wm + i / 32] |= 1u << (i % 32);
                if (!(ones == 0 && zeros == 4)) bad++;
            }
        }
    }
    return bad;
}

static inline int get_code(const uint32_t *c, int i) { return (c[i / 16] >> (2 * (i % 16))) & 3; }
static inline int get_mask(const uint32_t *m, int i, int L) { return i >= L ? 1 : (int)((m[i / 32] >> (i % 32)) & 1); }

/* ---- BN fold ---- */
ORC_API void orc_fold_bn(const float *w, const float *bias, const float *gamma, const float *beta,
                         const float *mean, const float *var, float eps, int U, int F,
                         float *wq, float *shift, float *scale_out)
{
    for (int u = 0; u < U; ++u) {
        float sc = gamma[u] / sqrtf(var[u] + eps);
        float sh = (bias[u] - mean[u]) * sc;
        sh = sh + beta[u];
        for (int i = 0; i < 4 * F; ++i) wq[u * 4 * F + i] = w[u * 4 * F + i] * sc;
        shift[u] = sh;
        if (scale_out) scale_out[u] = sc;
    }
}

/* ---- canonical conv: z (B,U,S), S = L-F+1.  wq (U,4,F) already scaled, shift (U). ---- */
ORC_API void orc_conv_z(const uint32_t *codes, const uint32_t *nmask, long B, int L,
                        const float *wq, const float *shift, int U, int F, float *z)
{
    int wc = orc_words_codes(L), wm = orc_words_mask(L);
    int S = L - F + 1, G = (F + 3) / 4;
    /* sequences are independent: the OpenMP split changes no arithmetic */
#pragma omp parallel for schedule(static)
    for (long b = 0; b < B; ++b) {
        const uint32_t *cb = codes + b * wc, *mb = nmask + b * wm;
        /* decode once per sequence: 0..3 = base, 4 = N column (same values get_code / get_mask return per tap) */
        uint8_t cd[L + 4 * G + 4];
        for (int i = 0; i < L + 4 * G + 4; ++i) cd[i] = (i >= L || get_mask(mb, i, L)) ? 4 : (uint8_t)get_code(cb, i);
        for (int u = 0; u < U; ++u) {
            const float *wu = wq + (long)u * 4 * F;
            float *zr = z + ((long)b * U + u) * S;
            for (int s = 0; s < S; ++s) {
                float acc = 0.0f;
                for (int g = 0; g < G; ++g) {
                    float t = 0.0f;
                    for (int j = 0; j < 4; ++j) {
                        int k = 4 * g + j;
                        float a = 0.0f;
                        if (k < F && cd[s + k] < 4) a = wu[cd[s + k] * F + k];
                        t = (j == 0) ? a : t + a;
                    }
                    acc = (g == 0) ? t : acc + t;
                }
                zr[s] = acc + shift[u];
            }
        }
    }
}

/* ---- max pool (k, stride, floor) on z, exp applied to the window maximum ---- */
ORC_API void orc_pool_exp(const float *z, long BU, int S, int pk, int ps,
                          float *pooled, uint8_t *idx)
{
    int P = (S - pk) / ps + 1;
    for (long r = 0; r < BU; ++r) {
        const float *zr = z + r * S;
        for (int p = 0; p < P; ++p) {
            float m = zr[p * ps];
            int im = 0;
            for (int j = 1; j < pk; ++j) {
                float v = zr[p * ps + j];
                if (v > m) { m = v; im = j; }
            }
            pooled[r * P + p] = orc_expf(m);
            if (idx) idx[r * P + p] = (uint8_t)im;
        }
    }
}

/* ---- float16 -> float32 (exact) ---- */
static inline float half_to_float(uint16_t h)
{
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16, e = (h >> 10) & 0x1Fu, m = h & 0x3FFu;
    if (e == 0x1Fu) return u2f(sign | 0x7F800000u | (m << 13));
    if (e == 0) {
        if (m == 0) return u2f(sign);
        float f = (float)m * 5.9604644775390625e-08f;        /* m * 2^-24, exact */
        return sign ? -f : f;
    }
    return u2f(sign | ((e + 112u) << 23) | (m << 13));
}

/* ---- per-unit maximum of float16 activations, numpy semantics (np.amax propagates NaN):
 *      interpretation.py:187  np.amax(activations, axis=(0, 2)).  act16 (N,U,S) raw float16 bits,
 *      amax16 (U) raw float16 bits out. ---- */
ORC_API void orc_unit_max_f16(const uint16_t *act16, long N, int U, int S, uint16_t *amax16)
{
#pragma omp parallel for schedule(static)
    for (int u = 0; u < U; ++u) {
        float m = -INFINITY;
        uint16_t mb = 0xFC00u;
        int nan = 0;
        for (long n = 0; n < N && !nan; ++n) {
            const uint16_t *row = act16 + ((long)n * U + u) * S;
            for (int s = 0; s < S; ++s) {
                float v = half_to_float(row[s]);
                if (v != v) { nan = 1; mb = row[s]; break; }
                if (v > m) { m = v; mb = row[s]; }
            }
        }
        amax16[u] = mb;
    }
}

/* ---- threshold / count stage of get_pwms_explainn (interpretation.py:203-214) on the caller's float16
 *      activations.  thr16 (U) raw float16 bits (0.5 * max, computed by the caller in float16 as numpy does);
 *      onehot (N,4,L) float32; a hit is act > thr in float16 (the conversion to float32 is exact, so the
 *      comparison is the same).  counts (U,4,F) and hits (U) int64, accumulated into.  seq_hit (N,U) u8 or NULL. ---- */
ORC_API void orc_pwm_counts_f16(const uint16_t *act16, long N, int U, int S, const uint16_t *thr16,
                                const float *onehot, int L, int F, long long *counts, long long *hits,
                                uint8_t *seq_hit)
{
#pragma omp parallel for schedule(dynamic, 1)
    for (int u = 0; u < U; ++u) {
        const float thr = half_to_float(thr16[u]);
        long long *cu = counts + (long)u * 4 * F;
        for (long n = 0; n < N; ++n) {
            const uint16_t *row = act16 + ((long)n * U + u) * S;
            const float *x = onehot + (long)n * 4 * L;
            int any = 0;
            for (int s = 0; s < S; ++s) {
                if (!(half_to_float(row[s]) > thr)) continue;
                any = 1;
                hits[u] += 1;
                for (int c = 0; c < 4; ++c)
                    for (int k = 0; k < F; ++k)
                        if (x[(long)c * L + s + k] != 0.0f) cu[c * F + k] += 1;
            }
            if (seq_hit) seq_hit[n * U + u] = (uint8_t)any;
        }
    }
}

/* ---- per (sequence, unit): maximum activation and the FIRST position that attains it, the rule of
 *      torch.max(act, dim=2) (networks.py:228,281).  act = orc_expf(z) in float32.  z (BU,S). ---- */
ORC_API void orc_seq_argmax(const float *z, long BU, int S, float *amax, int32_t *apos)
{
#pragma omp parallel for schedule(static)
    for (long r = 0; r < BU; ++r) {
        const float *zr = z + r * S;
        float m = orc_expf(zr[0]);
        int im = 0;
        for (int s = 1; s < S; ++s) {
            float v = orc_expf(zr[s]);
            if (v > m) { m = v; im = s; }
        }
        amax[r] = m;
        apos[r] = im;
    }
}
