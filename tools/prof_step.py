"""Small driver for ncu: two eager training steps at cfg 2 and two eval forwards / interpretation passes."""
import sys, torch
sys.path.insert(0, ".")
from explainn_b200 import networks, train as xtrain, interpretation
U, L, C, F, B = 300, 251, 81, 19, 512
dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(1)
codes = torch.randint(0, 4, (8192, L), generator=g, device=dev)
data = torch.zeros((8192, 4, L), device=dev); data.scatter_(1, codes.unsqueeze(1), 1.0)
labels = (torch.rand((8192, C), generator=g, device=dev) < 0.5).float()
torch.manual_seed(0)
m = networks.ExplaiNN(U, L, C, F).to(dev).train()
opt = torch.optim.Adam(m.parameters(), lr=1e-3)
tr = xtrain.FusedTrainer(m, torch.nn.BCEWithLogitsLoss(), opt, dropout_seed=1, use_graph=False)
for i in range(3):
    tr.step(data[i * B:(i + 1) * B], labels[i * B:(i + 1) * B])
m.eval()
with torch.no_grad():
    for i in range(2):
        m(data)
    acc = interpretation.PwmAccumulator(m)
    acc.scan_max(data); acc.finish_max(); acc.scan_count(data)
torch.cuda.synchronize()
print("done")
