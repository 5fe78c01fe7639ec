import sys, torch
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
prep = m._conv_prep_eval()
tcp = Fn.conv_tc_prepare(conv.weight.detach(), conv.bias.detach(), bn.weight.detach(), bn.bias.detach(), bn.running_mean, bn.running_var, bn.eps)
for _ in range(2):
    Fn.conv_pool_fwd(pk, prep); Fn.conv_pool_fwd_tc(pk, tcp)
torch.cuda.synchronize(); print("done")
