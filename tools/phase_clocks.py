"""Per-phase cycle counts of bwd_fc1_tc_kernel (library built with XB_NVCC_EXTRA=-DXB_PHASE_CLOCKS)."""
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
buf = (ctypes.c_longlong * (32 * 16))()
fn = _cabi.lib.xb_dbg_phase_clocks
fn.argtypes = [ctypes.c_void_p]
assert fn(buf) == 0
names = ["top", "coef+W1t", "h1 wait", "G build", "sync+D1 issue", "pooled fill", "D1 wait", "Gt stores", "sync+D2 issue",
         "D1 drain", "D2 wait", "D1/D2 sync", "D2 drain", "end sync"]
for t in range(9):
    c = [buf[t * 16 + i] for i in range(13)]
    if c[0] == 0:
        break
    d = [c[i + 1] - c[i] for i in range(12)]
    print("tile %d total %6d :" % (t, c[12] - c[0]), " ".join("%s=%d" % (n, v) for n, v in zip(
        ["coef", "h1wait", "Gbuild", "syncD1", "poolfill", "D1wait", "Gt", "syncD2", "D1drain", "D2wait+sync", "D2drain", "endsync"], d)))
