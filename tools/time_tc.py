import sys, torch
sys.path.insert(0, ".")
from explainn_b200 import networks, functional as Fn
U, L, C, F = 300, 251, 81, 19
dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(1)
def data(n, L):
    codes = torch.randint(0, 4, (n, L), generator=g, device=dev)
    x = torch.zeros((n, 4, L), device=dev); x.scatter_(1, codes.unsqueeze(1), 1.0); return x
def timeit(f, n=10):
    f(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): f()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n
for (U, L, B) in [(300, 251, 8192), (300, 251, 512), (500, 1000, 1024), (100, 200, 8192)]:
    torch.manual_seed(0)
    m = networks.ExplaiNN(U, L, C, F).to(dev).eval()
    x = data(B, L)
    pk = m._pack(x)
    conv, bn = m.linears[0], m.linears[1]
    prep = m._conv_prep_eval()
    tcp = Fn.conv_tc_prepare(conv.weight.detach(), conv.bias.detach(), bn.weight.detach(), bn.bias.detach(), bn.running_mean, bn.running_var, bn.eps)
    ref = Fn.conv_pool_fwd(pk, prep); out = Fn.conv_pool_fwd_tc(pk, tcp)
    rel = ((out - ref).abs() / ref.abs().clamp_min(1e-30)).max().item()
    t_lut = timeit(lambda: Fn.conv_pool_fwd(pk, prep)); t_tc = timeit(lambda: Fn.conv_pool_fwd_tc(pk, tcp))
    P = ref.shape[2]
    byt = B * ((L + 3) // 4 + (L + 7) // 8 + 4 * U * P)
    print(f"U={U} L={L} B={B}: rel={rel:.2e} LUT {t_lut*1e3:.1f} us ({byt/t_lut/1e6:.0f} GB/s)  TC {t_tc*1e3:.1f} us ({byt/t_tc/1e6:.0f} GB/s, {B/t_tc/1e3:.2f} M seq/s)")
    # train variant
    ga, be = bn.weight.detach(), bn.bias.detach()
    a1 = Fn.conv_pool_fwd_train(pk, prep, ga, be); a2 = Fn.conv_pool_fwd_train_tc(pk, tcp, ga, be)
    torch.cuda.synchronize()
    print("   train: pooled rel", ((a1[0]-a2[0]).abs()/a1[0].abs().clamp_min(1e-30)).max().item(), "xhat abs", (a1[1]-a2[1]).abs().max().item(), "idx mismatch", (a1[2]!=a2[2]).sum().item(), "of", a1[2].numel(),
          "LUT %.1f us TC %.1f us" % (timeit(lambda: Fn.conv_pool_fwd_train(pk, prep, ga, be))*1e3, timeit(lambda: Fn.conv_pool_fwd_train_tc(pk, tcp, ga, be))*1e3))
