"""Stage timings of the eval-mode forward (CUDA events, 20 repetitions each): pack, conv+pool, FC stack, final linear."""
import os, sys, torch
sys.path.insert(0, ".")
from explainn_b200 import networks, functional as Fn
U, L, C, F = 300, 251, 81, 19
dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(1)


def data(n, L):
    codes = torch.randint(0, 4, (n, L), generator=g, device=dev)
    x = torch.zeros((n, 4, L), device=dev)
    x.scatter_(1, codes.unsqueeze(1), 1.0)
    return x


def timeit(f, n=20):
    f(); f(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n):
        f()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n * 1e3


for B in (8192, 32768, 512):
    torch.manual_seed(0)
    m = networks.ExplaiNN(U, L, C, F).to(dev).eval()
    x = data(B, L)
    pk = m._pack(x)
    pooled = m._pooled_packed(pk)
    unit = m._unit_outputs_eval_packed(pk)
    Wf, bf = m.final.weight.detach(), m.final.bias.detach()
    ref = torch.nn.functional.linear(unit.double(), Wf.double(), bf.double())
    out = Fn.linear_fwd(unit, Wf, bf)
    LS = m.linears
    bn2, bn3 = LS[7], LS[11]
    fc = lambda: Fn.fc_stack_fwd_eval(pooled, LS[6].weight.detach(), LS[6].bias.detach(),
                                      (bn2.weight.detach(), bn2.bias.detach(), bn2.running_mean, bn2.running_var),
                                      LS[10].weight.detach(), LS[10].bias.detach(),
                                      (bn3.weight.detach(), bn3.bias.detach(), bn3.running_mean, bn3.running_var), bn2.eps)
    print("B=%d linear max abs err vs fp64 %.2e" % (B, (out.double() - ref).abs().max().item()))
    print("  pack %.1f us  conv+pool %.1f us  fc stack %.1f us  linear %.1f us  whole forward %.1f us" % (
        timeit(lambda: m._pack(x)), timeit(lambda: m._pooled_packed(pk)), timeit(fc),
        timeit(lambda: Fn.linear_fwd(unit, Wf, bf)), timeit(lambda: m(x))))
