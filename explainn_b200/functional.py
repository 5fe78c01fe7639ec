"""Thin tensor-level wrappers over the C ABI (one function per kernel group).

All tensors are CUDA tensors owned by PyTorch; the C layer never allocates.
"""
from collections import namedtuple

import torch

from . import _cabi
from ._cabi import call, ptr, stream

HIDDEN = 100          # networks.py:106
POOL = 7

Packed = namedtuple("Packed", "codes nmask B L invalid")
ConvPrep = namedtuple("ConvPrep", "tables wq_t shift scale U F")


def pooled_len(L, F, pool_size=7, pool_stride=7):
    """networks.py:105 -- int(((L-(F-1)) - (pool_size-1) - 1)/pool_stride + 1)."""
    return int(((L - (F - 1)) - (pool_size - 1) - 1) / pool_stride + 1)


def pack_onehot(x, num_cnns=1, invalid=None):
    """(B,4,L) or U-times tiled (B,4U,L) float one-hot on the GPU -> Packed (2-bit codes + N mask).

    `invalid` is a 1-element int32 CUDA counter that accumulates the number of columns that are
    neither one-hot nor all-zero; it is read lazily (see check_invalid) to avoid a sync per batch.
    """
    if x.dim() != 3:
        raise ValueError("expected (B,4,L) input, got shape %s" % (tuple(x.shape),))
    B, ch, L = x.shape
    if ch != 4 and ch != 4 * num_cnns:
        raise ValueError("expected 4 (or 4*num_cnns) input channels, got %d" % ch)
    if x.dtype == torch.float32:
        dt = _cabi.DTYPE_F32
    elif x.dtype == torch.float16:
        dt = _cabi.DTYPE_F16
    else:
        x = x.float()
        dt = _cabi.DTYPE_F32
    if not x.is_contiguous():
        x = x.contiguous()
    dev = x.device
    codes = torch.empty((B, (L + 15) // 16), dtype=torch.int32, device=dev)
    nmask = torch.empty((B, (L + 31) // 32), dtype=torch.int32, device=dev)
    if invalid is None:
        invalid = torch.zeros(1, dtype=torch.int32, device=dev)
    if B:
        call("explainn_pack_onehot", ptr(x), dt, B, L, ch * L, ptr(codes), ptr(nmask), ptr(invalid), stream())
    return Packed(codes, nmask, B, L, invalid)


def encode_dna(seqs, seq_len, device="cuda"):
    """List of DNA strings -> Packed, encoded on the GPU (semantics of `dna_one_hot`, tools.py:136-181).

    The host only concatenates the bytes (1 byte per base crosses PCIe instead of 16 for an fp32 one-hot)."""
    import numpy as np
    B = len(seqs)
    L = int(seq_len)
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("explainn_b200 runs only on CUDA devices (there is no CPU fallback)")
    raw = [s if isinstance(s, bytes) else str(s).encode("latin-1", "replace") for s in seqs]
    offs = np.zeros(B + 1, dtype=np.int64)
    if B:
        np.cumsum([len(r) for r in raw], out=offs[1:])
    text = np.frombuffer(b"".join(raw) + b"\0", dtype=np.uint8)       # never empty
    text_d = torch.from_numpy(text.copy()).to(dev, non_blocking=True)
    offs_d = torch.from_numpy(offs).to(dev, non_blocking=True)
    codes = torch.empty((B, (L + 15) // 16), dtype=torch.int32, device=dev)
    nmask = torch.empty((B, (L + 31) // 32), dtype=torch.int32, device=dev)
    if B:
        call("explainn_encode_dna", ptr(text_d), ptr(offs_d), B, L, ptr(codes), ptr(nmask), stream())
    return Packed(codes, nmask, B, L, torch.zeros(1, dtype=torch.int32, device=dev))


def encode_dna_array(text, seq_len=None, device=None):
    """(B, Lt) uint8 tensor of ASCII bases (host, ideally pinned, or device) -> Packed, encoded on the GPU with the
    rules of `dna_one_hot` (tools.py:136-181): 1 byte per base crosses PCIe instead of 16 for an fp32 one-hot."""
    if text.dim() != 2 or text.dtype != torch.uint8:
        raise ValueError("expected a (B, L) uint8 tensor of ASCII bases, got %s %s" % (tuple(text.shape), text.dtype))
    B, Lt = text.shape
    L = int(seq_len) if seq_len is not None else Lt
    if device is None:
        device = text.device if text.is_cuda else torch.device("cuda", torch.cuda.current_device())
    dev = torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError("explainn_b200 runs only on CUDA devices (there is no CPU fallback)")
    text_d = text.to(dev, non_blocking=True).contiguous()
    key = (str(dev), B, Lt)
    offs_d = _OFFSETS.get(key)
    if offs_d is None:
        offs_d = torch.arange(B + 1, dtype=torch.int64, device=dev) * Lt
        # under CUDA-graph capture the tensor lives in the graph's private pool (filled by captured kernels): not cached
        if not torch.cuda.is_current_stream_capturing():
            if len(_OFFSETS) > 64:
                _OFFSETS.clear()
            _OFFSETS[key] = offs_d
    codes = torch.empty((B, (L + 15) // 16), dtype=torch.int32, device=dev)
    nmask = torch.empty((B, (L + 31) // 32), dtype=torch.int32, device=dev)
    if B:
        call("explainn_encode_dna", ptr(text_d), ptr(offs_d), B, L, ptr(codes), ptr(nmask), stream())
    return Packed(codes, nmask, B, L, torch.zeros(1, dtype=torch.int32, device=dev))


_OFFSETS = {}


def unpack_onehot(pk, dtype=torch.float32):
    """Packed -> (B,4,L) one-hot tensor (N columns all zero)."""
    x = torch.empty((pk.B, 4, pk.L), dtype=dtype, device=pk.codes.device)
    dt = _cabi.DTYPE_F32 if dtype == torch.float32 else _cabi.DTYPE_F16
    if dtype not in (torch.float32, torch.float16):
        raise ValueError("unpack_onehot: float32 or float16")
    if pk.B:
        call("explainn_unpack_onehot", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, ptr(x), dt, stream())
    return x


def pwm_score(pk, prep, use_max, out=None):
    """Max or sum over positions of the grouped conv output; with `out` given the result is folded into it."""
    acc = out is not None
    if out is None:
        out = torch.empty((pk.B, prep.U), dtype=torch.float32, device=pk.codes.device)
    if pk.B:
        call("explainn_pwm_score", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, prep.U, prep.F, ptr(prep.tables),
             ptr(prep.wq_t), ptr(prep.shift), int(use_max), int(acc), ptr(out), stream())      # 0 sum, 1 max, 2 exp(max)
    return out


def fc1_stack_fwd_eval(pooled, W, b1, bn, eps):
    """num_fc = 1 unit in eval mode: pooled (B,U,P) -> relu(bn(W . pooled + b1)) as (B,U)."""
    B, U, P = pooled.shape
    g, be, rm, rv = bn
    out = torch.empty((B, U), dtype=torch.float32, device=pooled.device)
    if B:
        call("explainn_fc1_stack_fwd_eval", ptr(pooled), B, U, P, ptr(W.contiguous()), ptr(b1), ptr(g), ptr(be), ptr(rm), ptr(rv),
             float(eps), ptr(out), stream())
    return out


def check_invalid(invalid):
    n = int(invalid.item())
    if n:
        invalid.zero_()
        raise ValueError("explainn_b200: %d input column(s) are neither one-hot nor all-zero; "
                         "the fused kernels require one-hot DNA (there is no dense fallback)" % n)


def conv_prepare(w, bias, gamma, beta, mean, var, eps=1e-5):
    """Fold BatchNorm into the conv filter and build the k-mer tables."""
    U, four, F = w.shape
    dev = w.device
    tables = torch.empty(_cabi.lib.explainn_conv_tables_floats(U, F), dtype=torch.float32, device=dev)
    wq_t = torch.empty(_cabi.lib.explainn_conv_wq_floats(U, F), dtype=torch.float32, device=dev)
    shift = torch.empty(U, dtype=torch.float32, device=dev)
    scale = torch.empty(U, dtype=torch.float32, device=dev)
    call("explainn_conv_prepare", ptr(w.contiguous()), ptr(bias), ptr(gamma), ptr(beta), ptr(mean), ptr(var),
         float(eps), U, F, ptr(tables), ptr(wq_t), ptr(shift), ptr(scale), stream())
    return ConvPrep(tables, wq_t, shift, scale, U, F)


def conv_pool_fwd(pk, prep):
    """-> pooled (B,U,P)."""
    P = pooled_len(pk.L, prep.F)
    pooled = torch.empty((pk.B, prep.U, P), dtype=torch.float32, device=pk.codes.device)
    if pk.B == 0:
        return pooled
    call("explainn_conv_pool_fwd", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, prep.U, prep.F, ptr(prep.tables),
         ptr(prep.wq_t), ptr(prep.shift), ptr(pooled), stream())
    return pooled


def unit_activations(pk, prep, dtype=torch.float16, preact=False):
    """exp(z) as (B,U,S) in `dtype`; preact=True: z itself (float32), the input of pool_general."""
    S = pk.L - prep.F + 1
    if preact:
        dtype = torch.float32
    act = torch.empty((pk.B, prep.U, S), dtype=dtype, device=pk.codes.device)
    dt = _cabi.DTYPE_F32_PREACT if preact else (_cabi.DTYPE_F16 if dtype == torch.float16 else _cabi.DTYPE_F32)
    if pk.B == 0:
        return act
    call("explainn_unit_activations", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, prep.U, prep.F, ptr(prep.tables),
         ptr(prep.wq_t), ptr(prep.shift), ptr(act), dt, stream())
    return act


def interp_unit_max(pk, prep, zmax=None):
    """Per-unit maximum pre-activation; pass `zmax` to fold another chunk into it."""
    dev = pk.codes.device
    scratch = torch.empty(_cabi.lib.explainn_interp_unit_max_scratch_floats(prep.U), dtype=torch.float32, device=dev)
    acc = zmax is not None
    if zmax is None:
        zmax = torch.full((prep.U,), float("-inf"), dtype=torch.float32, device=dev)
        acc = True
    if pk.B == 0:
        return zmax
    call("explainn_interp_unit_max", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, prep.U, prep.F, ptr(prep.tables),
         ptr(prep.wq_t), ptr(prep.shift), ptr(scratch), ptr(zmax), int(acc), stream())
    return zmax


def interp_thresholds(zmax):
    U = zmax.numel()
    thr, zlo, zhi, amax = (torch.empty(U, dtype=torch.float32, device=zmax.device) for _ in range(4))
    call("explainn_interp_thresholds", ptr(zmax), U, ptr(thr), ptr(zlo), ptr(zhi), ptr(amax), stream())
    return thr, zlo, zhi, amax


def interp_seq_argmax(pk, prep):
    """Per (sequence, unit) maximum activation (float32) and the first position attaining it (int32), both (B,U)."""
    dev = pk.codes.device
    amax = torch.empty((pk.B, prep.U), dtype=torch.float32, device=dev)
    apos = torch.empty((pk.B, prep.U), dtype=torch.int32, device=dev)
    if pk.B:
        call("explainn_interp_seq_argmax", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, prep.U, prep.F, ptr(prep.tables),
             ptr(prep.wq_t), ptr(prep.shift), ptr(amax), ptr(apos), stream())
    return amax, apos


def _act_dtype(act):
    if act.dtype == torch.float16:
        return _cabi.DTYPE_F16
    if act.dtype == torch.float32:
        return _cabi.DTYPE_F32
    raise ValueError("activations must be float16 or float32, got %s" % act.dtype)


def act_unit_max(act, state=None):
    """Fold a chunk of (N,U,S) activations into the running per-unit maximum state (see explainn_act_unit_max)."""
    N, U, S = act.shape
    if state is None:
        state = torch.zeros(2 * U, dtype=torch.int32, device=act.device)
    if N:
        call("explainn_act_unit_max", ptr(act), _act_dtype(act), N, U, S, ptr(state), stream())
    return state


def act_thresholds(state, U, dtype):
    amax = torch.empty(U, dtype=dtype, device=state.device)
    thr = torch.empty(U, dtype=dtype, device=state.device)
    call("explainn_act_thresholds", ptr(state), U, _cabi.DTYPE_F16 if dtype == torch.float16 else _cabi.DTYPE_F32,
         ptr(amax), ptr(thr), stream())
    return amax, thr


def act_pwm_count(act, pk, F, thr, counts=None, hits=None, seq_hit=None):
    N, U, S = act.shape
    if S != pk.L - F + 1 or pk.B != N:
        raise ValueError("activations of shape %s do not match %d sequences of length %d and filter_size %d"
                         % (tuple(act.shape), pk.B, pk.L, F))
    if counts is None:
        counts = torch.zeros((U, 4, F), dtype=torch.int64, device=act.device)
        hits = torch.zeros(U, dtype=torch.int64, device=act.device)
    if N:
        call("explainn_act_pwm_count", ptr(act), _act_dtype(act), ptr(pk.codes), ptr(pk.nmask), N, pk.L, U, F, ptr(thr),
             ptr(counts), ptr(hits), ptr(seq_hit), stream())
    return counts, hits


def revcomp_packed(pk):
    """Reverse complement of packed sequences (the other strand)."""
    codes = torch.empty_like(pk.codes)
    nmask = torch.empty_like(pk.nmask)
    if pk.B:
        call("explainn_revcomp_packed", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, ptr(codes), ptr(nmask), stream())
    return Packed(codes, nmask, pk.B, pk.L, pk.invalid)


def interp_pwm_count(pk, prep, thr, zlo, zhi, counts=None, hits=None, seq_hit=None):
    dev = pk.codes.device
    if counts is None:
        counts = torch.zeros((prep.U, 4, prep.F), dtype=torch.int64, device=dev)
        hits = torch.zeros(prep.U, dtype=torch.int64, device=dev)
    if pk.B == 0:
        return counts, hits
    nbytes = _cabi.lib.explainn_interp_pwm_count_scratch_bytes(pk.B, pk.L, prep.U, prep.F)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    call("explainn_interp_pwm_count", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, prep.U, prep.F, ptr(prep.tables),
         ptr(prep.wq_t), ptr(prep.shift), ptr(zlo), ptr(zhi), ptr(thr), ptr(counts), ptr(hits), ptr(seq_hit), ptr(scratch),
         stream())
    return counts, hits


def fc_stack_fwd_eval(pooled, W1, b1, bn2, W2, b2, bn3, eps=1e-5):
    """pooled (B,U,P) -> unit outputs (B,U).  bn2/bn3 = (weight, bias, running_mean, running_var)."""
    B, U, P = pooled.shape
    out = torch.empty((B, U), dtype=torch.float32, device=pooled.device)
    if B == 0:
        return out
    call("explainn_fc_stack_fwd_eval", ptr(pooled), B, U, P, ptr(W1), ptr(b1), ptr(bn2[0]), ptr(bn2[1]), ptr(bn2[2]),
         ptr(bn2[3]), ptr(W2), ptr(b2), ptr(bn3[0]), ptr(bn3[1]), ptr(bn3[2]), ptr(bn3[3]), float(eps), ptr(out), stream())
    return out


def linear_fwd(x, W, bias):
    B, U = x.shape
    Cc = W.shape[0]
    out = torch.empty((B, Cc), dtype=torch.float32, device=x.device)
    if B == 0:
        return out
    call("explainn_linear_fwd", ptr(x), B, U, Cc, ptr(W), ptr(bias), ptr(out), stream())
    return out


# ------------------------------------------------------------------------------------------ training
def input_cooc_stats(pk, F):
    M = torch.empty((4 * F, 4 * F), dtype=torch.int32, device=pk.codes.device)
    call("explainn_input_cooc_stats", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, F, ptr(M), stream())
    return M


def bn1_batch_stats(M, w, bias, n_pos, momentum, run_mean, run_var, nbt):
    U, _, F = w.shape
    dev = w.device
    mean = torch.empty(U, dtype=torch.float32, device=dev)
    var = torch.empty(U, dtype=torch.float32, device=dev)
    Mw = torch.empty((U, 4 * F), dtype=torch.float32, device=dev)
    call("explainn_bn1_batch_stats", ptr(M), ptr(w), ptr(bias), U, F, n_pos, float(momentum), ptr(run_mean), ptr(run_var),
         ptr(nbt), ptr(mean), ptr(var), ptr(Mw), stream())
    return mean, var, Mw


def conv_pool_fwd_train(pk, prep, gamma, beta):
    P = pooled_len(pk.L, prep.F)
    dev = pk.codes.device
    pooled = torch.empty((pk.B, prep.U, P), dtype=torch.float32, device=dev)
    xhat = torch.empty((pk.B, prep.U, P), dtype=torch.float32, device=dev)
    idx = torch.empty((pk.B, prep.U, P), dtype=torch.uint8, device=dev)
    call("explainn_conv_pool_fwd_train", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, prep.U, prep.F, ptr(prep.tables),
         ptr(prep.wq_t), ptr(prep.shift), ptr(gamma), ptr(beta), ptr(pooled), ptr(xhat), ptr(idx), stream())
    return pooled, xhat, idx


def fc_nchunk(B):
    return _cabi.lib.explainn_fc_nchunk(B)


# ------------------------------------------------------------------------------------------ tensor-core conv
def conv_tc_prepare(w, bias, gamma, beta, mean, var, eps=1e-5, sign_of=None):
    """Weight images of the tensor-core conv.  `sign_of` (U): training passes the BatchNorm gain whose sign is folded
    into the rows (the prepare call itself then gets gamma = 1, beta = 0 and the batch statistics)."""
    U, four, F = w.shape
    images = torch.empty(_cabi.lib.explainn_conv_tc_image_halfs(U, F), dtype=torch.float16, device=w.device)
    shift = torch.empty(2 * U, dtype=torch.float32, device=w.device)
    call("explainn_conv_tc_prepare", ptr(w.contiguous()), ptr(bias), ptr(gamma), ptr(beta), ptr(mean), ptr(var), float(eps),
         ptr(sign_of), U, F, ptr(images), ptr(shift), stream())
    return images, shift, U, F


def conv_pool_fwd_tc(pk, tcprep, fp16_weights=False):
    """fp16_weights: the stated-tolerance mode (one fp16 piece per weight, half the tensor-core work)."""
    images, shift, U, F = tcprep
    P = pooled_len(pk.L, F)
    pooled = torch.empty((pk.B, U, P), dtype=torch.float32, device=pk.codes.device)
    if pk.B:
        call("explainn_conv_pool_fwd_tc_fp16" if fp16_weights else "explainn_conv_pool_fwd_tc", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, U, F, ptr(images), ptr(shift), ptr(pooled),
             stream())
    return pooled


def conv_pool_fwd_train_tc(pk, tcprep, gamma, beta):
    images, shift, U, F = tcprep
    P = pooled_len(pk.L, F)
    dev = pk.codes.device
    pooled = torch.empty((pk.B, U, P), dtype=torch.float32, device=dev)
    xhat = torch.empty((pk.B, U, P), dtype=torch.float32, device=dev)
    idx = torch.empty((pk.B, U, P), dtype=torch.uint8, device=dev)
    call("explainn_conv_pool_fwd_train_tc", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, U, F, ptr(images), ptr(shift), ptr(gamma), ptr(beta),
         ptr(pooled), ptr(xhat), ptr(idx), stream())
    return pooled, xhat, idx


# ------------------------------------------------------------------------------------------ model variants (generic.cu)
def pool_general(z, pool_size, pool_stride, gamma=None, beta=None):
    """General max-pool of the pre-activation map z (B,U,S).  Eval (gamma None): pooled.  Train: pooled, xhat, idx16."""
    B, U, S = z.shape
    P = (S - pool_size) // pool_stride + 1
    dev = z.device
    pooled = torch.empty((B, U, P), dtype=torch.float32, device=dev)
    train = gamma is not None
    xhat = torch.empty((B, U, P), dtype=torch.float32, device=dev) if train else None
    idx = torch.empty((B, U, P), dtype=torch.int16, device=dev) if train else None
    if B:
        call("explainn_pool_general", ptr(z), B, U, S, int(pool_size), int(pool_stride), ptr(gamma), ptr(beta), ptr(pooled), ptr(xhat),
             ptr(idx), stream())
    return (pooled, xhat, idx) if train else pooled


def hidden_fwd_eval(X, W, b, bn, eps):
    """Hidden block in eval mode: X (B,U,Pin) -> (B,U,100).  bn = (weight, bias, running_mean, running_var)."""
    B, U, Pin = X.shape
    dev = X.device
    out = torch.empty((B, U, HIDDEN), dtype=torch.float32, device=dev)
    if B:
        Y = torch.empty((U, B, HIDDEN), dtype=torch.float32, device=dev)
        stat = torch.empty(2 * U * HIDDEN, dtype=torch.float32, device=dev)
        call("explainn_hidden_fwd_eval", ptr(X), B, U, Pin, ptr(W.contiguous()), ptr(b), ptr(bn[0]), ptr(bn[1]), ptr(bn[2]), ptr(bn[3]),
             float(eps), ptr(Y), ptr(stat), ptr(out), stream())
    return out


# ------------------------------------------------------------------------------------------ tensor-core interpretation
class InterpTc:
    """State of the tensor-core interpretation passes for one model: weight images, error bounds, the list workspace
    and the device status word (see include/explainn_b200.h: explainn_interp_unit_max_tc / _pwm_count_tc)."""

    def __init__(self, w, prep, tcprep):
        self.prep, self.tcprep = prep, tcprep
        dev = w.device
        U, F = prep.U, prep.F
        self.delta = torch.empty(U, dtype=torch.float32, device=dev)
        call("explainn_interp_tc_bounds", ptr(w.contiguous()), ptr(prep.scale), ptr(prep.shift), U, F, ptr(self.delta), stream())
        self.list_bytes = int(_cabi.lib.explainn_interp_tc_list_bytes(U))
        self.list = torch.empty(self.list_bytes, dtype=torch.uint8, device=dev)
        self.status = torch.zeros(1, dtype=torch.int32, device=dev)

    def unit_max(self, pk, zmax):
        images, shift_rs, U, F = self.tcprep
        if pk.B:
            call("explainn_interp_unit_max_tc", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, U, F, ptr(images), ptr(shift_rs),
                 ptr(self.delta), ptr(self.prep.wq_t), ptr(self.prep.shift), ptr(zmax), ptr(self.list), self.list_bytes,
                 ptr(self.status), stream())
        return zmax

    def pwm_count(self, pk, thr, zlo, zhi, counts, hits, seq_hit=None):
        images, shift_rs, U, F = self.tcprep
        if pk.B:
            nbytes = _cabi.lib.explainn_interp_pwm_count_scratch_bytes(pk.B, pk.L, U, F)
            scratch = torch.empty(nbytes, dtype=torch.uint8, device=pk.codes.device)
            call("explainn_interp_pwm_count_tc", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, U, F, ptr(images), ptr(shift_rs),
                 ptr(self.delta), ptr(self.prep.wq_t), ptr(self.prep.shift), ptr(zlo), ptr(zhi), ptr(thr), ptr(counts), ptr(hits),
                 ptr(seq_hit), ptr(scratch), ptr(self.list), self.list_bytes, ptr(self.status), stream())
        return counts, hits

    def check(self):
        """Raise if a pass had more positions to re-evaluate than the list workspace holds (one small D2H read)."""
        st = int(self.status.item())
        if st:
            raise RuntimeError("explainn_b200: the tensor-core interpretation pass %s overflowed its resolve list; "
                               "rerun with PwmAccumulator(model, backend='lut')" % ("A" if st & 1 else "B"))
