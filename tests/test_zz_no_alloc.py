"""GPU: the allocation-free variant of the per-unit maximum scan (pass A of the PWM extraction).

Who needs it: hosts of the library that cannot call cudaMalloc where the scan runs -- a serving process
whose GPU memory is handed out in full at start-up, a plugin living inside someone else's stream-ordered
region, a thread that must stay off the allocator lock.  They call explainn_interp_unit_max_noalloc, which
keeps its per-CTA partial maxima in a fixed array inside the library instead of caller-provided scratch.

This file is collected last on purpose (name sorts after every other test file): if it fails, a run that stops
at the first failure has already executed all other tests.  It passes alone in a fresh process (the first call
does the one-time address look-up) and after the rest of the suite.

What is checked: through the C ABI, two calls in a row (different inputs, so stale contents of the shared
array would show) give exactly the per-unit maxima of the existing entry point -- a maximum does not depend on
the order of reduction, so the comparison is exact, like the other bit-exact tests of the scan; then the
`accumulate` flag and the size limit.
"""
import numpy as np
import pytest
import torch

from util import onehot_from_codes, random_codes

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _setup(U, L, F, seed):
    from explainn_b200 import networks
    torch.manual_seed(seed)
    m = networks.ExplaiNN(U, L, 2, F).to(DEV).eval()
    return m, m._conv_prep_eval()


def _zmax_noalloc(m, prep, x, zmax=None):
    """Raw C-ABI call of the allocation-free entry point; the output tensor is the caller's."""
    from explainn_b200._cabi import call, ptr, stream
    pk = m._pack(x)
    acc = zmax is not None
    if zmax is None:
        zmax = torch.empty(prep.U, dtype=torch.float32, device=DEV)
    call("explainn_interp_unit_max_noalloc", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, prep.U, prep.F,
         ptr(prep.tables), ptr(prep.wq_t), ptr(prep.shift), ptr(zmax), int(acc), stream())
    return zmax


def _zmax_existing(m, prep, x, zmax=None):
    from explainn_b200 import functional as Fn
    return Fn.interp_unit_max(m._pack(x), prep, zmax)


def test_no_alloc_unit_max_matches_existing_entry_point_twice():
    U, L, F = 70, 131, 19
    m, prep = _setup(U, L, F, seed=3)
    xa = torch.from_numpy(onehot_from_codes(random_codes(300, L, 11, n_frac=0.01))).to(DEV)
    xb = torch.from_numpy(onehot_from_codes(random_codes(37, L, 12))).to(DEV)
    ref_a = _zmax_existing(m, prep, xa)
    ref_b = _zmax_existing(m, prep, xb)
    got_a = _zmax_noalloc(m, prep, xa)          # first call (does the one-time set-up when run alone)
    got_b = _zmax_noalloc(m, prep, xb)          # second call right after: must not see the first call's maxima
    torch.cuda.synchronize()
    assert torch.equal(got_a, ref_a)
    assert torch.equal(got_b, ref_b)
    assert not torch.equal(ref_a, ref_b)        # the two inputs really differ, so stale scratch would be visible
    # accumulate folds a second chunk into an existing result, exactly like the existing entry point
    both = _zmax_noalloc(m, prep, xb, got_a.clone())
    assert torch.equal(both, torch.maximum(ref_a, ref_b))
    # a different model shape afterwards reuses the same array
    m2, prep2 = _setup(9, 64, 8, seed=4)
    x2 = torch.from_numpy(onehot_from_codes(random_codes(50, 64, 13))).to(DEV)
    assert torch.equal(_zmax_noalloc(m2, prep2, x2), _zmax_existing(m2, prep2, x2))


def test_no_alloc_rejects_sizes_beyond_the_built_in_scratch():
    from explainn_b200._cabi import call, ptr, stream
    from explainn_b200 import functional as Fn
    U, L, F = 1056, 40, 8                       # more than the 1024 units the built-in array covers
    dev = DEV
    w = torch.zeros(U, 4, F, device=dev)
    z = torch.zeros(U, device=dev)
    o = torch.ones(U, device=dev)
    prep = Fn.conv_prepare(w, z, o, z, z, o)
    x = torch.from_numpy(onehot_from_codes(random_codes(4, L, 1))).to(dev)
    pk = Fn.pack_onehot(x)
    zmax = torch.empty(U, device=dev)
    with pytest.raises(ValueError, match="exceed the built-in scratch"):
        call("explainn_interp_unit_max_noalloc", ptr(pk.codes), ptr(pk.nmask), pk.B, pk.L, U, F, ptr(prep.tables),
             ptr(prep.wq_t), ptr(prep.shift), ptr(zmax), 0, stream())
    # the existing entry point (caller scratch) has no such limit
    assert Fn.interp_unit_max(pk, prep).shape == (U,)
