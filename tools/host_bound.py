"""Is the training loop host-bound?  Host time per FusedTrainer.step call (no synchronisation inside the loop) beside
the device time per step, cfg 2, batch 512, one GPU."""
import sys, time, torch
sys.path.insert(0, ".")
from explainn_b200 import networks, train as xtrain
U, L, C, F, B = 300, 251, 81, 19, 512
dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(1)
codes = torch.randint(0, 4, (8 * B, L), generator=g, device=dev)
data = torch.zeros((8 * B, 4, L), device=dev); data.scatter_(1, codes.unsqueeze(1), 1.0)
labels = (torch.rand((8 * B, C), generator=g, device=dev) < 0.5).float()
torch.manual_seed(0)
m = networks.ExplaiNN(U, L, C, F).to(dev).train()
opt = torch.optim.Adam(m.parameters(), lr=1e-3)
tr = xtrain.FusedTrainer(m, torch.nn.BCEWithLogitsLoss(), opt, dropout_seed=1)
for i in range(10):
    tr.step(data[:B], labels[:B])
torch.cuda.synchronize()
for K in (50, 200):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    t0 = time.perf_counter()
    for i in range(K):
        j = (i % 8) * B
        tr.step(data[j:j + B], labels[j:j + B])
    t_host = time.perf_counter() - t0
    e.record()
    torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    print("K=%d: host issue time %.1f us/step, device time %.1f us/step, wall %.1f us/step"
          % (K, t_host / K * 1e6, s.elapsed_time(e) / K * 1e3, t_all / K * 1e6))
# graph replay alone (no input copies, no Adam): the device time of forward + loss + backward
c = next(iter(tr._graphs.values()))
torch.cuda.synchronize()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for i in range(100):
    c.graph.replay()
e.record(); torch.cuda.synchronize()
print("graph replay alone: %.1f us" % (s.elapsed_time(e) / 100 * 1e3))
