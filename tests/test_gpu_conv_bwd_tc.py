"""explainn_conv_pool_bwd: the tensor-core scatter (csrc/conv_bwd_tc.cu; CTA = (128-unit tile, slice of the batch), one
`partial` slice per CTA of a tile) against the shared-memory scatter kernel (csrc/train.cu, few slices) and against a float64 restatement of the sum they compute:
    A[u, c*F + k] = sum_{b,p} gpool[b,u,p] * pooled[b,u,p] * X[b, c, 7p + idx[b,u,p] + k]
(backward of MaxPool1d -> exp -> BatchNorm1d -> grouped Conv1d, explainn/models/networks.py:98-102)."""
import numpy as np
import pytest
import torch

from util import onehot_from_codes, random_codes

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _inputs(B, U, L, F, seed, n_frac=0.01):
    from explainn_b200 import functional as Fn
    g = torch.Generator().manual_seed(seed)
    S = L - F + 1
    P = (S - 7) // 7 + 1
    x = onehot_from_codes(random_codes(B, L, seed, n_frac=n_frac))
    pk = Fn.pack_onehot(torch.from_numpy(x).to(DEV))
    # gradients over many orders of magnitude and both signs, as a trained model produces them
    gpool = (torch.randn(B, U, P, generator=g) * torch.exp(4 * torch.randn(B, U, P, generator=g))).to(DEV)
    pooled = torch.exp(torch.randn(B, U, P, generator=g)).to(DEV)
    xhat = torch.randn(B, U, P, generator=g).to(DEV)
    idx = torch.randint(0, 7, (B, U, P), generator=g, dtype=torch.uint8).to(DEV)
    return x, pk, gpool, pooled, xhat, idx, P


def _run(pk, B, L, U, F, gpool, pooled, xhat, idx, nsplit):
    from explainn_b200 import _cabi
    from explainn_b200 import functional as Fn
    ptr, call, stream = _cabi.ptr, _cabi.call, _cabi.stream
    W = 4 * F
    M = Fn.input_cooc_stats(pk, F)
    f32 = dict(dtype=torch.float32, device=DEV)
    Mw = torch.zeros(U, W, **f32)
    bias, mean, var, gamma = torch.zeros(U, **f32), torch.zeros(U, **f32), torch.ones(U, **f32), torch.ones(U, **f32)
    partial = torch.full((nsplit, U, W + 2), float("nan"), **f32)
    dW, db, dg, dbe = torch.empty(U, W, **f32), torch.empty(U, **f32), torch.empty(U, **f32), torch.empty(U, **f32)
    call("explainn_conv_pool_bwd", ptr(pk.codes), ptr(pk.nmask), B, L, U, F, ptr(gpool), ptr(pooled), ptr(xhat), ptr(idx), ptr(M),
         ptr(Mw), ptr(bias), ptr(mean), ptr(var), ptr(gamma), 0.0, ptr(partial), nsplit, ptr(dW), ptr(db), ptr(dg), ptr(dbe), stream())
    torch.cuda.synchronize()
    return partial, dW, dg, dbe


def _reference_sum(x, gpool, pooled, xhat, idx, F):
    B, U, P = gpool.shape
    gz = (gpool.double() * pooled.double()).cpu().numpy()
    # the kernels round gz to fp32 first
    gz = (gpool * pooled).double().cpu().numpy()
    ix = idx.cpu().numpy().astype(np.int64)
    A = np.zeros((U, 4, F))
    pos = 7 * np.arange(P)[None, None, :] + ix                       # (B,U,P)
    xd = x.astype(np.float64)
    for k in range(F):
        for b in range(B):
            A[:, :, k] += np.einsum("up,cup->uc", gz[b], xd[b][:, pos[b] + k])
    dgamma = (gz * xhat.double().cpu().numpy()).sum(axis=(0, 2))
    dbeta = gz.sum(axis=(0, 2))
    return A.reshape(U, 4 * F), dgamma, dbeta


@pytest.mark.parametrize("B,U,L,F", [(37, 300, 251, 19), (5, 44, 251, 19), (200, 132, 200, 24), (9, 520, 120, 8), (64, 4, 251, 19)])
def test_tc_scatter_vs_sum(B, U, L, F):
    from explainn_b200 import _cabi
    x, pk, gpool, pooled, xhat, idx, P = _inputs(B, U, L, F, seed=B + U)
    assert (U * P) % 4 == 0, "shape must take the tensor-core path"
    n_tc = _cabi.lib.explainn_conv_bwd_nsplit(U)
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    assert n_tc >= sms
    part_tc, dW_tc, dg_tc, db_tc = _run(pk, B, L, U, F, gpool, pooled, xhat, idx, n_tc)
    part_sp, dW_sp, dg_sp, db_sp = _run(pk, B, L, U, F, gpool, pooled, xhat, idx, 2)
    used = min(sms // ((U + 127) // 128), B)           # CTAs per 128-unit tile, one slice each
    assert not torch.isnan(part_tc[:used]).any(), "every CTA writes its whole slice"
    assert torch.isnan(part_tc[used:]).all(), "slices beyond the grid are left alone"
    A_ref, dg_ref, db_ref = _reference_sum(x, gpool, pooled, xhat, idx, F)
    A_tc = part_tc[:used, :, :4 * F].double().sum(0).cpu().numpy()
    A_sp = part_sp[:2, :, :4 * F].double().sum(0).cpu().numpy()
    # per unit: error against the float64 sum, relative to the sum of magnitudes (what fp32 accumulation is bounded by)
    scale = np.abs(A_ref).max(axis=1, keepdims=True) + 1e-30
    err_tc = np.abs(A_tc - A_ref) / scale
    err_sp = np.abs(A_sp - A_ref) / scale
    gz = (gpool * pooled).double().abs().sum(dim=(0, 2)).cpu().numpy()[:, None] + 1e-30
    assert (np.abs(A_tc - A_ref) / gz).max() < 2e-6, (np.abs(A_tc - A_ref) / gz).max()
    assert (np.abs(A_sp - A_ref) / gz).max() < 2e-6
    assert err_tc.max() < max(10 * err_sp.max(), 1e-5), (err_tc.max(), err_sp.max())
    for got, ref in ((dg_tc, dg_ref), (db_tc, db_ref), (dg_sp, dg_ref), (db_sp, db_ref)):
        g_ = got.double().cpu().numpy()
        assert np.abs(g_ - ref).max() <= 2e-6 * (gz.max() * 4), np.abs(g_ - ref).max()
    # the assembled gradient agrees between the two kernels
    s = dW_sp.abs().max(dim=1, keepdim=True).values + 1e-30
    assert ((dW_tc - dW_sp).abs() / s).max().item() < 1e-4


def test_tc_scatter_deterministic():
    from explainn_b200 import _cabi
    B, U, L, F = 96, 300, 251, 19
    x, pk, gpool, pooled, xhat, idx, P = _inputs(B, U, L, F, seed=3)
    n_tc = _cabi.lib.explainn_conv_bwd_nsplit(U)
    a = _run(pk, B, L, U, F, gpool, pooled, xhat, idx, n_tc)
    b = _run(pk, B, L, U, F, gpool, pooled, xhat, idx, n_tc)
    for t0, t1 in zip(a[1:], b[1:]):
        assert torch.equal(t0, t1)


def test_unaligned_shape_takes_the_scatter_kernel():
    """U * P not a multiple of 4: no 16-byte aligned blocks for the bulk copies -> shared-memory scatter, same result."""
    from explainn_b200 import _cabi
    B, U, L, F = 20, 7, 251, 19          # P = 33, U*P = 231
    x, pk, gpool, pooled, xhat, idx, P = _inputs(B, U, L, F, seed=5)
    n = _cabi.lib.explainn_conv_bwd_nsplit(U)
    part, dW, dg, db = _run(pk, B, L, U, F, gpool, pooled, xhat, idx, n)
    A_ref, dg_ref, db_ref = _reference_sum(x, gpool, pooled, xhat, idx, F)
    A = torch.nan_to_num(part[:, :, :4 * F]).double().sum(0).cpu().numpy()
    gz = (gpool * pooled).double().abs().sum(dim=(0, 2)).cpu().numpy()[:, None]
    assert (np.abs(A - A_ref) / gz).max() < 2e-6
