"""Where the three warp roles of bwd_fc1_tc_kernel spend their cycles (CTA 0, summed over its tiles).
Library built with XB_NVCC_EXTRA=-DXB_PHASE_CLOCKS."""
import ctypes, sys, torch
sys.path.insert(0, ".")
from explainn_b200 import networks, train as xtrain, _cabi
U, L, C, F, B = 300, 251, 81, 19, 512
dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(1)
codes = torch.randint(0, 4, (2048, L), generator=g, device=dev)
data = torch.zeros((2048, 4, L), device=dev); data.scatter_(1, codes.unsqueeze(1), 1.0)
labels = (torch.rand((2048, C), generator=g, device=dev) < 0.5).float()
torch.manual_seed(0)
m = networks.ExplaiNN(U, L, C, F).to(dev).train()
opt = torch.optim.Adam(m.parameters(), lr=1e-3)
tr = xtrain.FusedTrainer(m, torch.nn.BCEWithLogitsLoss(), opt, dropout_seed=1, use_graph=False)
for i in range(4):
    tr.step(data[i * B:(i + 1) * B], labels[i * B:(i + 1) * B])
torch.cuda.synchronize()
buf = (ctypes.c_longlong * (3 * 16))()
fn = _cabi.lib.xb_dbg_phase_clocks
fn.argtypes = [ctypes.c_void_p]
assert fn(buf) == 0
names = [["barriers", "wait acc_free", "D1 issue", "bar Gt", "wait at_full", "D2 issue"],
         ["top", "pooled loads", "wait D2 prev", "At stores", "wait D1", "D1 drain", "wait D2", "D2 drain"],
         ["loop", "h1 loads+coef", "bar", "wait D2 prev", "G build", "bar", "wait D1", "Gt stores", "bar"]]
for r, role in enumerate(["issue", "helper", "worker"]):
    v = [buf[r * 16 + i] for i in range(len(names[r]))]
    print("%-7s total %7d :" % (role, sum(v)), " ".join("%s=%d" % (n, x) for n, x in zip(names[r], v)))
