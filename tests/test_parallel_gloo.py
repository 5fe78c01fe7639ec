"""CPU, world_size 2, gloo: host-side logic of the data-parallel paths (sharding, the MAX / SUM reductions of
the interpretation pass, gradient averaging, ordered gather).  The per-shard compute is done by the oracle here
(no GPU in this container); the reductions and their semantics are the product code in explainn_b200/parallel.py."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from explainn_b200 import parallel
    from oracle import canonical as cn
    from util import load_golden, onehot_from_codes, state_dict_from

    # ---- shard_range: ordered, contiguous, balanced
    for n in (0, 1, 5, 96, 97):
        parts = [parallel.shard_range(n, r, world) for r in range(world)]
        assert parts[0][0] == 0 and parts[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
        assert max(h - l for l, h in parts) - min(h - l for l, h in parts) <= 1
    assert parallel.world() == (rank, world)

    # ---- sharded PWM extraction == unsharded (oracle compute, product reductions)
    g = load_golden("eval_u40_f20")
    sd = state_dict_from(g)
    n = lambda k: sd[k].numpy()
    L, F = int(g["L"]), int(g["F"])
    x = onehot_from_codes(g["codes"])
    wq, shift, _ = cn.fold_bn(n("linears.0.weight"), n("linears.0.bias"), n("linears.1.weight"), n("linears.1.bias"),
                              n("linears.1.running_mean"), n("linears.1.running_var"))
    lo, hi = parallel.shard_range(x.shape[0])
    codes, nmask, _ = cn.pack_onehot(x[lo:hi])
    z = cn.conv_z(codes, nmask, L, wq, shift)
    zmax = torch.from_numpy(z.max(axis=(0, 2)).copy())
    parallel.allreduce_unit_max(zmax)                                   # pass A exchange
    amax16 = cn.expf(zmax.numpy()).astype(np.float16)
    a16 = cn.activations_fp16(z)
    counts, hits, _ = cn.pwm_counts(a16, x[lo:hi], F, amax16=amax16)
    counts, hits = torch.from_numpy(counts), torch.from_numpy(hits)
    parallel.allreduce_counts(counts, hits)                             # pass B exchange
    pwm = cn.pwm_from_counts_hits(counts.numpy(), hits.numpy())
    assert np.array_equal(amax16, g["amax16"])
    assert np.array_equal(pwm, g["pwm"]), "sharded PWMs differ from the reference's"

    # ---- gradient averaging factor and ordered gather
    flat = torch.full((10,), float(rank + 1))
    scale = parallel.allreduce_mean_gradients(flat)
    assert scale == 1.0 / world and torch.allclose(flat * scale, torch.full((10,), (1 + world) / 2.0))
    rows = torch.arange(lo, hi, dtype=torch.float32).unsqueeze(1)
    allrows = parallel.gather_rows(rows, x.shape[0])
    assert torch.equal(allrows[:, 0], torch.arange(x.shape[0], dtype=torch.float32))
    open(os.path.join(out_dir, "ok%d" % rank), "w").write("ok")
    dist.destroy_process_group()


def test_world_size_2_gloo(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / ("ok%d" % r)).exists() for r in range(world))


def test_single_process_helpers_are_identity():
    sys.path.insert(0, ROOT)
    from explainn_b200 import parallel
    t = torch.arange(4.0)
    assert parallel.allreduce_mean_gradients(t) == 1.0
    assert parallel.shard_range(10) == (0, 10)
    assert parallel.gather_rows(t.unsqueeze(1), 4).shape == (4, 1)
