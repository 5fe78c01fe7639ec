"""ctypes front-end of the canonical C restatement plus the numpy half of the interpretation
pass.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Reference lines restated here (relative to /root/reference):
  explainn/interpretation/interpretation.py:147   activations cast to float16
  explainn/interpretation/interpretation.py:187   threshold = 0.5 * max over (N, S), in float16
  explainn/interpretation/interpretation.py:203-224  windows above threshold summed, column-normalised
  explainn/interpretation/interpretation.py:230-258  MEME text serialisation
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

_lib = None


def lib():
    global _lib
    if _lib is None:
        path = _build.build()
        L = C.CDLL(path)
        L.orc_expf.restype = C.c_float
        L.orc_expf.argtypes = [C.c_float]
        L.orc_pack_onehot.restype = C.c_long
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def words_codes(L):
    return (L + 15) // 16


def words_mask(L):
    return (L + 31) // 32


def expf(x):
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.empty_like(x)
    lib().orc_expf_vec(_p(x), _p(y), C.c_long(x.size))
    return y


def pack_onehot(x):
    """x: (B,4,L) float32 -> (codes uint32 (B,Wc), nmask uint32 (B,Wm), n_invalid_columns)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    B, four, L = x.shape
    assert four == 4
    codes = np.zeros((B, words_codes(L)), np.uint32)
    nmask = np.zeros((B, words_mask(L)), np.uint32)
    bad = lib().orc_pack_onehot(_p(x), C.c_long(B), C.c_int(L), C.c_long(4 * L), _p(codes), _p(nmask))
    return codes, nmask, int(bad)


def fold_bn(w, bias, gamma, beta, mean, var, eps=1e-5):
    w = np.ascontiguousarray(w, np.float32)
    U, four, F = w.shape
    arrs = [np.ascontiguousarray(a, np.float32) for a in (bias, gamma, beta, mean, var)]
    wq = np.empty_like(w)
    shift = np.empty(U, np.float32)
    scale = np.empty(U, np.float32)
    lib().orc_fold_bn(_p(w), *[_p(a) for a in arrs], C.c_float(eps), C.c_int(U), C.c_int(F),
                      _p(wq), _p(shift), _p(scale))
    return wq, shift, scale


def conv_z(codes, nmask, L, wq, shift):
    """Canonical folded conv: returns z (B,U,S) float32."""
    B = codes.shape[0]
    U, _, F = wq.shape
    S = L - F + 1
    z = np.empty((B, U, S), np.float32)
    lib().orc_conv_z(_p(codes), _p(nmask), C.c_long(B), C.c_int(L), _p(np.ascontiguousarray(wq)),
                     _p(np.ascontiguousarray(shift)), C.c_int(U), C.c_int(F), _p(z))
    return z


def pool_exp(z, pool_size=7, pool_stride=7):
    B, U, S = z.shape
    P = (S - pool_size) // pool_stride + 1
    pooled = np.empty((B, U, P), np.float32)
    idx = np.empty((B, U, P), np.uint8)
    lib().orc_pool_exp(_p(np.ascontiguousarray(z)), C.c_long(B * U), C.c_int(S), C.c_int(pool_size),
                       C.c_int(pool_stride), _p(pooled), _p(idx))
    return pooled, idx


def activations_fp16(z):
    """float16(det_expf(z)) -- the array interpretation.py:147 builds."""
    return expf(z).astype(np.float16)


def decode_onehot(codes, nmask, L):
    """Packed form -> (B,4,L) float32 one-hot (N columns all zero)."""
    B = codes.shape[0]
    i = np.arange(L)
    c = (codes[:, i // 16] >> (2 * (i % 16)).astype(np.uint32)) & 3
    m = (nmask[:, i // 32] >> (i % 32).astype(np.uint32)) & 1
    x = np.zeros((B, 4, L), np.float32)
    bb, ii = np.nonzero(m == 0)
    x[bb, c[bb, ii], ii] = 1.0
    return x


def unit_max_fp16(act16):
    """interpretation.py:187 -- per-unit max over (N,S) of the float16 activations."""
    return np.amax(act16, axis=(0, 2))


def pwm_counts(act16, onehot, filter_size, amax16=None):
    """Integer restatement of interpretation.py:187-214.

    act16: (N,U,S) float16; onehot: (N,4,L) 0/1.  Returns (counts int64 (U,4,F), hits int64 (U,),
    thr float16 (U,)).  counts[u,c,k] = number of (n,s) with act16[n,u,s] > thr[u] and
    onehot[n,c,s+k] == 1.  `amax16` lets a sharded caller pass the global maxima.
    """
    if amax16 is None:
        amax16 = unit_max_fp16(act16)
    thr = (np.float16(0.5) * amax16.astype(np.float16)).astype(np.float16)   # float16 arithmetic, as numpy does at :187
    N, U, S = act16.shape
    F = filter_size
    oh = np.asarray(onehot)
    counts = np.zeros((U, 4, F), np.int64)
    hits = np.zeros(U, np.int64)
    n_i, u_i, s_i = np.nonzero(act16 > thr[None, :, None])
    hits += np.bincount(u_i, minlength=U)
    ohi = (oh != 0).astype(np.int64)
    for k in range(F):
        col = ohi[n_i, :, s_i + k]            # (hits, 4)
        np.add.at(counts[:, :, k], u_i, col)
    return counts, hits, thr


def unit_max_fp16_c(act16):
    """interpretation.py:187 in C (OpenMP over units) -- same result as unit_max_fp16, for the large test shapes."""
    a = np.ascontiguousarray(act16, np.float16)
    N, U, S = a.shape
    out = np.empty(U, np.float16)
    lib().orc_unit_max_f16(_p(a), C.c_long(N), C.c_int(U), C.c_int(S), _p(out))
    return out


def pwm_counts_c(act16, onehot, filter_size, amax16=None, counts=None, hits=None, want_seq_hit=False):
    """interpretation.py:187-214 in C (OpenMP over units): same integers as pwm_counts (tests/test_oracle_golden.py
    holds the two against each other and against the reference's PWMs), fast enough for >= 16k sequences at
    U=300.  `counts`/`hits` may be passed in to accumulate over chunks; `amax16` carries the global maxima then."""
    a = np.ascontiguousarray(act16, np.float16)
    x = np.ascontiguousarray(onehot, np.float32)
    N, U, S = a.shape
    L = x.shape[2]
    F = int(filter_size)
    assert x.shape[0] == N and x.shape[1] == 4 and S + F - 1 <= L
    if amax16 is None:
        amax16 = unit_max_fp16_c(a)
    thr = (np.float16(0.5) * np.asarray(amax16, np.float16)).astype(np.float16)
    if counts is None:
        counts = np.zeros((U, 4, F), np.int64)
        hits = np.zeros(U, np.int64)
    seq_hit = np.zeros((N, U), np.uint8) if want_seq_hit else None
    lib().orc_pwm_counts_f16(_p(a), C.c_long(N), C.c_int(U), C.c_int(S), _p(thr), _p(x), C.c_int(L), C.c_int(F),
                             _p(counts), _p(hits), _p(seq_hit) if want_seq_hit else None)
    if want_seq_hit:
        return counts, hits, thr, seq_hit
    return counts, hits, thr


def seq_argmax(z):
    """Per (sequence, unit): max activation (float32 det_expf(z)) and its FIRST position -- torch.max(act, dim=2),
    networks.py:228,281.  z (B,U,S) float32 -> (amax (B,U) float32, apos (B,U) int32)."""
    z = np.ascontiguousarray(z, np.float32)
    B, U, S = z.shape
    amax = np.empty((B, U), np.float32)
    apos = np.empty((B, U), np.int32)
    lib().orc_seq_argmax(_p(z), C.c_long(B * U), C.c_int(S), _p(amax), _p(apos))
    return amax, apos


def pwm_from_counts_hits(counts, hits):
    """interpretation.py:196,217-224 -- float32 column normalisation widened to float64; units
    for which no window was collected (hits == 0) keep 0.25 everywhere."""
    U, four, F = counts.shape
    pwm = np.full((U, 4, F), 0.25, np.float64)
    for u in range(U):
        if hits[u] == 0:
            continue
        c = counts[u].astype(np.float32)
        tot = c.sum(axis=0)
        tot[tot == 0] = 1.0
        pwm[u] = np.nan_to_num(c / tot)
    return pwm


def pwm_to_meme_text(pwm, names=None):
    """interpretation.py:230-258 as a string (same bytes the reference writes)."""
    n = len(pwm)
    if names is None:
        names = {i: "filter%d" % i for i in range(n)}
    out = ["MEME version 4\n"]
    for i in range(n):
        if np.sum(pwm[i]) > 0:
            out.append("\n")
            out.append("MOTIF %s \n" % names[i])
            out.append("letter-probability matrix: alength= 4 w= %d \n"
                       % np.count_nonzero(np.sum(pwm[i], axis=0)))
        for j in range(pwm[i].shape[1]):
            if np.sum(pwm[i][:, j]) > 0:
                out.append("\t".join(str(pwm[i][c, j]) for c in range(4)) + "\n")
    return "".join(out)


_DNA_CODE = {"A": 0, "C": 1, "G": 2, "T": 3, "0": 0, "1": 1, "2": 2, "3": 3}


def dna_one_hot(seq, seq_len=None):
    """explainn/utils/tools.py:136-181 restated without the per-character try/except: (4,L) float16.

    The reference upper-cases, replaces A,C,G,T by the digits '0'..'3' and indexes with int(char), so the digits
    '0'..'3' in the input also select a row; every other character (and every position outside the placed
    sequence) is an all-zero column.  Longer than seq_len: centre crop from (len-seq_len)//2 (tools.py:153-155);
    shorter: placed at column (seq_len-len)//2 (tools.py:158)."""
    if seq_len is None:
        seq_len = len(seq)
        start = 0
    elif seq_len <= len(seq):
        trim = (len(seq) - seq_len) // 2
        seq = seq[trim:trim + seq_len]
        start = 0
    else:
        start = (seq_len - len(seq)) // 2
    seq = seq.upper()
    x = np.zeros((4, seq_len), np.float16)
    for i, ch in enumerate(seq):
        c = _DNA_CODE.get(ch)
        if c is not None:
            x[c, start + i] = 1.0
    return x


def revcomp_onehot(x):
    """`_flip(_flip(x, 1), 2)` of networks.py:430 / `s[::-1, ::-1]` of fasta2predictions.py:160 on (B,4,L)."""
    return np.ascontiguousarray(x[:, ::-1, ::-1])


def pwm_scores(onehot, pwms, scoring="sum"):
    """networks.PWM.forward (networks.py:428-437) in the canonical arithmetic: grouped conv (bias 0) on both strands,
    then max or a float32 sum over positions in increasing order (forward strand first, then the reverse strand).
    onehot (B,4,L) float, pwms (G,4,F) float32 -> (B,G) float32."""
    G, four, F = pwms.shape
    ident = (np.zeros(G, np.float32), np.ones(G, np.float32), np.zeros(G, np.float32), np.zeros(G, np.float32),
             np.ones(G, np.float32))
    wq, shift, _ = fold_bn(np.ascontiguousarray(pwms, np.float32), ident[0], ident[1], ident[2], ident[3], ident[4], eps=0.0)
    outs = []
    for x in (onehot, revcomp_onehot(onehot)):
        codes, nmask, _ = pack_onehot(np.ascontiguousarray(x, np.float32))
        outs.append(conv_z(codes, nmask, x.shape[2], wq, shift))            # (B,G,S)
    if scoring == "max":
        return np.maximum(outs[0].max(axis=2), outs[1].max(axis=2)).astype(np.float32)
    acc = []
    for z in outs:
        s = np.zeros(z.shape[:2], np.float32)
        for t in range(z.shape[2]):
            s = (s + z[:, :, t]).astype(np.float32)
        acc.append(s)
    return (acc[0] + acc[1]).astype(np.float32)
