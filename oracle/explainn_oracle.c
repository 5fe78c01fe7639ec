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
    for (long b = 0; b < B; ++b) {
        const uint32_t *cb = codes + b * wc, *mb = nmask + b * wm;
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
                        if (k < F && !get_mask(mb, s + k, L)) a = wu[get_code(cb, s + k) * F + k];
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
