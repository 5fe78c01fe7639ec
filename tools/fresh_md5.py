"""Debug aid for the rare parity failure: in a fresh process, after the same preceding tests and in
the order of tests/test_gpu_train.py::test_train_step_vs_torch_port, prints an md5 of every workspace tensor and
gradient of the GPU step, and the two errors the test checks first.  `sort | uniq -c` over many runs shows a deviant."""
import sys, os, hashlib
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, torch
import test_gpu_train as T
from util import onehot_from_codes, random_codes
from explainn_b200 import networks, autograd_fn
from oracle import torch_port as tp
dev = "cuda:0"
for name in T.TRAIN_GOLDENS:
    for backend in ("tc", "lut"):
        T.test_train_forward_backward_vs_reference(name, backend)
for fused in (False, True):
    for name in T.TRAIN_GOLDENS:
        T.test_two_adam_steps_vs_reference(name, fused)
U, L, C, F, B = 300, 251, 81, 19, 160
FIELDS = ["M", "mean1", "var1", "Mw", "pooled", "xhat", "idx", "h1", "mean2", "invstd2", "h2", "mean3", "invstd3", "unit", "logits"]
real_fwd = autograd_fn.train_forward
keep = {}
def spy_fwd(*a, **k):
    ws = real_fwd(*a, **k)
    for f in FIELDS:
        keep[f] = getattr(ws, f).clone()
    return ws
autograd_fn.train_forward = spy_fwd
# ---- the body of the test, same order ----
torch.manual_seed(7)
port = tp.PortExplaiNN(U, L, C, F)
gen = torch.Generator().manual_seed(8)
for mod in port.modules():
    if isinstance(mod, torch.nn.BatchNorm1d):
        mod.weight.data.uniform_(0.5, 1.5, generator=gen)
        mod.bias.data.normal_(0, 0.1, generator=gen)
m = networks.ExplaiNN(U, L, C, F)
m.load_state_dict(port.state_dict())
m = m.to(dev).train()
m.conv_backend = "tc"
m.linears[9].eval()
port.train()
port.linears[9].eval()
x = torch.from_numpy(onehot_from_codes(random_codes(B, L, 10, n_frac=0.01)))
y = (torch.rand(B, C, generator=gen) < 0.5).float()
crit = torch.nn.BCEWithLogitsLoss()
out_p = port(x)
lp = crit(out_p, y)
lp.backward()
out = m(x.to(dev))
lm = crit(out, y.to(dev))
lm.backward()
h = lambda t: hashlib.md5(t.detach().cpu().numpy().tobytes()).hexdigest()[:6]
e_log = float((out.detach().cpu() - out_p.detach()).abs().max() / out_p.detach().abs().max())
gw, pw = m.linears[0].weight.grad.cpu(), port.linears[0].weight.grad
e_w = float((gw - pw).abs().max() / pw.abs().max())
w_gpu = h(torch.cat([p.detach().flatten() for p in m.parameters()]))
line = "elog %.2e ew %.2e W %s cpu %s %s | " % (e_log, e_w, w_gpu, h(out_p), h(pw)) + " ".join("%s %s" % (f, h(keep[f])) for f in FIELDS)
line += " | " + " ".join("%s %s" % (k.replace("linears.", "l"), h(p.grad)) for k, p in m.named_parameters() if "final" not in k)
print(line)
