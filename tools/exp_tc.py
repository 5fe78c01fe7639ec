import os, sys, torch
sys.path.insert(0, ".")
from explainn_b200 import networks, functional as Fn
U, L, C, F, B = 300, 251, 81, 19, 8192
dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(1)
codes = torch.randint(0, 4, (B, L), generator=g, device=dev)
x = torch.zeros((B, 4, L), device=dev); x.scatter_(1, codes.unsqueeze(1), 1.0)
torch.manual_seed(0)
m = networks.ExplaiNN(U, L, C, F).to(dev).eval()
pk = m._pack(x)
conv, bn = m.linears[0], m.linears[1]
tcp = Fn.conv_tc_prepare(conv.weight.detach(), conv.bias.detach(), bn.weight.detach(), bn.bias.detach(), bn.running_mean, bn.running_var, bn.eps)
def timeit(f, n=10):
    f(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): f()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n
for dbg in [0, 16, 1, 2, 4, 8, 3, 5, 6, 7, 12, 15]:
    os.environ["XB_TC_DBG"] = str(dbg)
    print("dbg=%2d  %.1f us" % (dbg, timeit(lambda: Fn.conv_pool_fwd_tc(pk, tcp)) * 1e3))
