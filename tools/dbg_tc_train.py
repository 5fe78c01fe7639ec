import sys, numpy as np, torch
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
from util import onehot_from_codes, random_codes
from explainn_b200 import networks, autograd_fn
from oracle import torch_port as tp
U, L, C, F, B = 300, 251, 81, 19, 160
torch.manual_seed(7)
port = tp.PortExplaiNN(U, L, C, F)
gen = torch.Generator().manual_seed(8)
for mod in port.modules():
    if isinstance(mod, torch.nn.BatchNorm1d):
        mod.weight.data.uniform_(0.5, 1.5, generator=gen); mod.bias.data.normal_(0, 0.1, generator=gen)
m = networks.ExplaiNN(U, L, C, F); m.load_state_dict(port.state_dict()); m = m.cuda().train(); m.linears[9].eval()
x = torch.from_numpy(onehot_from_codes(random_codes(B, L, 10, n_frac=0.01))).cuda()
res = {}
for be in ("lut", "tc"):
    m.conv_backend = be
    with torch.no_grad():
        ws = autograd_fn.train_forward(m, x, seed=1)
    res[be] = (ws.pooled.clone(), ws.xhat.clone(), ws.idx.clone(), ws.logits.clone())
a, b = res["lut"], res["tc"]
d = (a[0] - b[0]).abs() / a[0].abs().clamp_min(1e-30)
print("pooled max rel", d.max().item(), "at", np.unravel_index(int(d.argmax()), d.shape), "n>1e-5:", int((d > 1e-5).sum()))
i = np.unravel_index(int(d.argmax()), d.shape); print(a[0][i].item(), b[0][i].item(), "idx", a[2][i].item(), b[2][i].item(), "xhat", a[1][i].item(), b[1][i].item())
print("xhat max abs", (a[1] - b[1]).abs().max().item(), "idx mism", int((a[2] != b[2]).sum()))
print("logits max abs", (a[3] - b[3]).abs().max().item())
bad = torch.nonzero(d > 1e-5)
print(bad[:10].tolist())
# ---- gradients: lut vs tc vs port
y = (torch.rand(B, C, generator=gen) < 0.5).float()
crit = torch.nn.BCEWithLogitsLoss()
port.train(); port.linears[9].eval()
lp = crit(port(x.cpu()), y); lp.backward()
grads = {}
for be in ("lut", "tc"):
    m.conv_backend = be; m.zero_grad()
    crit(m(x), y.cuda()).backward()
    grads[be] = {k: p.grad.detach().cpu().clone() for k, p in m.named_parameters()}
for k, p in port.named_parameters():
    g = p.grad; sc = g.abs().max().item() + 1e-30
    e1 = (grads["lut"][k] - g).abs().max().item() / sc; e2 = (grads["tc"][k] - g).abs().max().item() / sc
    e3 = (grads["tc"][k] - grads["lut"][k]).abs().max().item() / sc
    print("%-18s lut-port %.2e  tc-port %.2e  tc-lut %.2e" % (k, e1, e2, e3))
k = "linears.6.weight"
d = (grads["tc"][k] - grads["lut"][k]).abs().reshape(U, 100, -1)
u = int(d.amax(dim=(1, 2)).argmax()); print("worst unit", u, d[u].max().item(), "scale unit", grads["lut"][k].reshape(U,100,-1)[u].abs().max().item())
pl, pt = res["lut"][0][:, u, :], res["tc"][0][:, u, :]
print("pooled unit: max", pl.max().item(), "mean", pl.mean().item(), "std over batch", pl.std(0).mean().item(), "max rel diff", ((pl-pt).abs()/pl).max().item())
