"""Two-GPU data-parallel step over NCCL (skipped on a one-GPU box): after one FusedTrainer step with the gradient
all-reduce overlapped with the conv backward, both ranks hold identical parameters, and the reduced gradient equals
the mean of the two ranks' gradients computed by one process."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, traceback
def _hook(t, v, tb):
    open(sys.argv[2] + ".err%s" % os.environ.get("RANK", "0"), "w").write("".join(traceback.format_exception(t, v, tb)))
    sys.__excepthook__(t, v, tb)
sys.excepthook = _hook
import torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
from util import onehot_from_codes, random_codes
from explainn_b200 import networks, train as xtrain, autograd_fn
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = "cuda:%d" % rank
dist.init_process_group("nccl", device_id=torch.device(dev))
U, L, C, F, B = 20, 80, 3, 19, 32
use_graph = sys.argv[3] == "graph"
torch.manual_seed(0)
m = networks.ExplaiNN(U, L, C, F).to(dev).train()
m.linears[9].eval()                                   # dropout off: the gradients are then a deterministic function of the batch
ref = networks.ExplaiNN(U, L, C, F).to(dev).train()
ref.load_state_dict(m.state_dict()); ref.linears[9].eval()
xs = [torch.from_numpy(onehot_from_codes(random_codes(B, L, 100 + r, n_frac=0.02))).to(dev) for r in range(world)]
ys = [(torch.rand(B, C, generator=torch.Generator().manual_seed(200 + r)) < 0.5).float().to(dev) for r in range(world)]
opt = torch.optim.Adam(m.parameters(), lr=1e-2)
tr = xtrain.FusedTrainer(m, torch.nn.BCEWithLogitsLoss(), opt, dropout_seed=1, use_graph=use_graph)
for s in range(2):                                   # step 2 captures the graphs (when enabled), step 3 replays them
    tr.step(xs[rank], ys[rank])
# single-process restatement of step 3: both ranks' gradients through the autograd path of the same kernels, then the mean
ref.load_state_dict(m.state_dict())
crit = torch.nn.BCEWithLogitsLoss()
mean_g = None
for r in range(world):
    ref.zero_grad()
    crit(ref(xs[r]), ys[r]).backward()
    g = [p.grad.clone() for p in autograd_fn._params(ref)]
    mean_g = g if mean_g is None else [a + b for a, b in zip(mean_g, g)]
mean_g = [g / world for g in mean_g]
p2p = os.environ.get("XB_P2P_ADAM", "1") != "0"
if p2p:
    tr.sync_optimizer_state()                        # the moments of a slice live on its owner: gather them for the check
p0, m0, v0 = tr.flat_p.clone(), tr.flat_m.clone(), tr.flat_v.clone()
tr.step(xs[rank], ys[rank])
torch.cuda.synchronize()
if p2p:
    # the fused kernel sums the peers' buffers on the fly; every rank keeps its local gradients
    assert tr.p2p is not None
    allg = [torch.empty_like(tr.flat_g) for _ in range(world)]
    dist.all_gather(allg, tr.flat_g)
    flat = allg[0].clone()
    for q in range(1, world):
        flat += allg[q]
    flat /= world
    got = [flat[o:o + p.numel()].view_as(p) for p, o in zip(tr.params, tr.offsets)]
    # and the update itself: Adam (lr 1e-2, default betas / eps) on the mean gradient
    t = float(tr.step_t.item())
    m1 = m0 + (flat - m0) * 0.1
    v1 = v0 * 0.999 + flat * flat * 0.001
    p1 = p0 - (1e-2 / (1 - 0.9 ** t)) * (m1 / (v1.sqrt() / (1 - 0.999 ** t) ** 0.5 + 1e-8))
    assert float((tr.flat_p - p1).abs().max()) <= 1e-6 * max(float(p1.abs().max()), 1.0), "fused Adam update"
    # ... and this rank's slices of the moments (two ranges when the FC-stack part overlaps the conv backward)
    rngs = ([tr.p2p.slice_range(tr.n_conv), tr.p2p.slice_range(tr.n - tr.n_conv, tr.n_conv)] if tr.p2p_overlap
            else [tr.p2p.slice_range(tr.n)])
    for lo, hi in rngs:
        assert float((tr.flat_m[lo:hi] - m1[lo:hi]).abs().max()) <= 1e-6 * max(float(m1.abs().max()), 1e-30), "exp_avg slice"
    tr.check_health()
else:
    got = [v / world for v in tr.views_g]            # the flat buffer holds the SUM over ranks; Adam applies 1/world
err = 0.0
for i, (a, b) in enumerate(zip(got, mean_g)):
    # index 3 = BatchNorm-1 beta: its gradient is mathematically zero, both sides hold rounding noise (test_gpu_train.py)
    scale = float(mean_g[2].abs().max()) * 1e4 if i == 3 else float(b.abs().max())
    if scale > 0:
        err = max(err, float((a - b).abs().max()) / scale)
    else:
        err = max(err, float(a.abs().max()))
flat = tr.flat_p.clone()
others = [torch.empty_like(flat) for _ in range(world)]
dist.all_gather(others, flat)
same = all(torch.equal(o, others[0]) for o in others)
if rank == 0:
    open(sys.argv[2], "w").write("%d %g %d\n" % (int(same), err, int(any(c.graph2 is not None for c in tr._graphs.values()))))
dist.barrier()
dist.destroy_process_group()
'''


@pytest.mark.parametrize("overlap", ["0", "1", "p2p", "p2p_serial"])
@pytest.mark.parametrize("mode", ["eager", "graph"])
def test_two_gpu_step_allreduce(tmp_path, mode, overlap):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    out = tmp_path / "out.txt"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", str(script), ROOT, str(out), mode]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=dict(os.environ, XB_DP_OVERLAP="1" if overlap == "1" else "0", XB_P2P_ADAM="1" if overlap.startswith("p2p") else "0",
                                XB_P2P_OVERLAP="0" if overlap == "p2p_serial" else "1"))
    # "0" / "1": NCCL all-reduce (plain / overlapped); "p2p": peer-memory step, FC part beside the conv backward; "p2p_serial": one call
    errs = "".join(p.read_text() for p in sorted(tmp_path.glob("out.txt.err*")))
    assert r.returncode == 0, errs + r.stdout[-1000:] + r.stderr[-1000:]
    same, err, split = out.read_text().split()
    assert int(same) == 1, "ranks diverged"
    assert float(err) < 1e-4, (same, err, split)                    # same kernels on both paths; only summation order differs
    assert int(split) == (1 if mode == "graph" and overlap in ("1", "p2p") else 0)


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_model_on_non_current_device():
    """A model on cuda:1 while cuda:0 is the current device: launches, streams and per-device caches follow the data."""
    import numpy as np
    from explainn_b200 import interpretation, networks
    from util import onehot_from_codes, random_codes
    torch.cuda.set_device(0)
    torch.manual_seed(0)
    m = networks.ExplaiNN(40, 120, 5, 19)
    x = torch.from_numpy(onehot_from_codes(random_codes(64, 120, 3, n_frac=0.01)))
    m0 = m.to("cuda:0").eval()
    with torch.no_grad():
        ref = m0(x.to("cuda:0")).cpu()
        pwm0 = interpretation.get_pwms_explainn_fused(x, m0, "cuda:0")
    m1 = m0.to("cuda:1")
    with torch.no_grad():
        out = m1(x.to("cuda:1")).cpu()
        pwm1 = interpretation.get_pwms_explainn_fused(x, m1, "cuda:1")
    assert torch.equal(out, ref) and np.array_equal(pwm0, pwm1)
    m1.train()
    y = torch.rand(64, 5, device="cuda:1")
    torch.nn.functional.mse_loss(m1(x.to("cuda:1")), y).backward()
    assert all(torch.isfinite(p.grad).all() for p in m1.parameters())
    assert torch.cuda.current_device() == 0


SYNC_WORKER = r'''
import os, sys, traceback
def _hook(t, v, tb):
    open(sys.argv[2] + ".err%s" % os.environ.get("RANK", "0"), "w").write("".join(traceback.format_exception(t, v, tb)))
    sys.__excepthook__(t, v, tb)
sys.excepthook = _hook
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "tests"))
from util import onehot_from_codes, random_codes
from explainn_b200 import networks, train as xtrain
from oracle import torch_port as tp
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = "cuda:%d" % rank
dist.init_process_group("nccl", device_id=torch.device(dev))
U, L, C, F, Bl = 20, 80, 3, 19, 48
torch.manual_seed(0)
port = tp.PortExplaiNN(U, L, C, F)
gen = torch.Generator().manual_seed(8)
for mod in port.modules():
    if isinstance(mod, torch.nn.BatchNorm1d):
        mod.weight.data.uniform_(0.5, 1.5, generator=gen)
        mod.bias.data.normal_(0, 0.1, generator=gen)
m = networks.ExplaiNN(U, L, C, F)
m.load_state_dict(port.state_dict())
m = m.to(dev).train()
m.sync_bn = True
m.linears[9].eval()
port.train(); port.linears[9].eval()
x = torch.from_numpy(onehot_from_codes(random_codes(Bl * world, L, 10, n_frac=0.01)))
y = (torch.rand(Bl * world, C, generator=gen) < 0.5).float()
crit = torch.nn.BCEWithLogitsLoss()
# the reference step on the CONCATENATED batch (single device)
out_p = port(x)
lp = crit(out_p, y)
lp.backward()
xs, ys = x[rank * Bl:(rank + 1) * Bl].to(dev), y[rank * Bl:(rank + 1) * Bl].to(dev)
out = m(xs)
lm = crit(out, ys)
lm.backward()
def close(a, b, tol, what):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    err = np.abs(a - b).max() / max(np.abs(b).max(), 1e-30)
    assert err <= tol, "%s: %.3e of scale (tol %.1e)" % (what, err, tol)
close(out.detach().cpu().numpy(), out_p.detach().numpy()[rank * Bl:(rank + 1) * Bl], 1e-4, "logits")
lsum = lm.detach().clone(); dist.all_reduce(lsum); lsum /= world
assert abs(float(lsum) - float(lp)) <= 1e-5 * abs(float(lp)), "loss"
zero = ("linears.0.bias", "linears.6.bias", "linears.10.bias")
dgam = float(port.linears[1].weight.grad.abs().max())
for (k, p), (_, q) in zip(port.named_parameters(), m.named_parameters()):
    g = q.grad.detach().clone(); dist.all_reduce(g); g /= world           # what the data-parallel step applies
    g, ref = g.cpu().numpy(), p.grad.numpy()
    if k in zero:
        assert np.abs(g).max() == 0.0, k
    elif k == "linears.1.bias":
        assert np.abs(g).max() < 1e-3 * dgam, k
    else:
        d = np.abs(g - ref); scale = np.abs(ref).max()
        assert np.quantile(d, 0.98) <= 1e-4 * scale and d.max() <= 2e-2 * scale, (k, float(np.quantile(d, 0.98) / scale), float(d.max() / scale))
for (k, a), (_, b) in zip(port.state_dict().items(), m.state_dict().items()):
    if "running" in k:
        close(b.cpu().numpy(), a.numpy(), 1e-5, k)
# without sync_bn the local statistics give different logits (the test would be vacuous otherwise)
m.sync_bn = False
with torch.no_grad():
    out_local = m(xs)
assert float((out_local - out.detach()).abs().max()) > 1e-3 * float(out.detach().abs().max())
# the fused trainer takes the same path (eager): replicas stay identical, loss equals the reference's
m2 = networks.ExplaiNN(U, L, C, F); m2.load_state_dict(port.state_dict()); m2 = m2.to(dev).train(); m2.sync_bn = True; m2.linears[9].eval()
opt = torch.optim.Adam(m2.parameters(), lr=1e-3)
tr = xtrain.FusedTrainer(m2, crit, opt)
l1 = tr.step(xs, ys).clone(); dist.all_reduce(l1); l1 /= world
l2 = tr.step(xs, ys)
assert tr._graph is None and abs(float(l1) - float(lp)) <= 1e-5 * abs(float(lp)), "fused loss"
torch.cuda.synchronize()
flat = tr.flat_p.clone(); others = [torch.empty_like(flat) for _ in range(world)]; dist.all_gather(others, flat)
assert all(torch.equal(o, others[0]) for o in others), "replicas diverged"
if rank == 0:
    open(sys.argv[2], "w").write("ok\n")
dist.barrier()
dist.destroy_process_group()
'''


def test_sync_bn_equals_reference_on_concatenated_batch(tmp_path):
    """model.sync_bn = True under two ranks: logits, loss, rank-averaged gradients and BatchNorm running statistics equal
    the reference step on the CONCATENATED batch on one device (the port), which local statistics do not."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "worker.py"
    script.write_text(SYNC_WORKER)
    out = tmp_path / "out.txt"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29537", str(script), ROOT, str(out)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    errs = "".join(p.read_text() for p in sorted(tmp_path.glob("out.txt.err*")))
    assert r.returncode == 0, errs + r.stdout[-1000:] + r.stderr[-1000:]
    assert out.read_text().strip() == "ok"
