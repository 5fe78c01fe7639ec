import sys, torch, numpy as np
sys.path.insert(0, ".")
from explainn_b200 import networks, functional as Fn
dev = "cuda:0"
U, L, F, B = 300, 251, 19, 8192
g = torch.Generator(device=dev).manual_seed(1)
codes = torch.randint(0, 4, (B, L), generator=g, device=dev)
x = torch.zeros((B, 4, L), device=dev); x.scatter_(1, codes.unsqueeze(1), 1.0)
pk = Fn.pack_onehot(x)
m = networks.PWM(np.random.default_rng(0).normal(size=(U, 4, F)).astype(np.float32), L, "max").to(dev)
prep = m._prep()
def timeit(f, n=5):
    f(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): f()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n * 1e3
print("prep %.1f us" % timeit(m._prep))
print("revcomp %.1f us" % timeit(lambda: Fn.revcomp_packed(pk)))
print("unit_max scan %.1f us" % timeit(lambda: Fn.interp_unit_max(pk, prep)))
print("pwm_score max %.1f us" % timeit(lambda: Fn.pwm_score(pk, prep, True)))
print("pwm_score sum %.1f us" % timeit(lambda: Fn.pwm_score(pk, prep, False)))
print("forward_packed %.1f us" % timeit(lambda: m.forward_packed(pk)))
