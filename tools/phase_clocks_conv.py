"""Per-phase cycles of conv_tc_kernel<INFER>, CTA 0 (library built with XB_NVCC_EXTRA=-DXB_PHASE_CLOCKS)."""
import ctypes, sys, torch
sys.path.insert(0, ".")
from explainn_b200 import networks, functional as Fn, _cabi
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
for _ in range(3):
    Fn.conv_pool_fwd_tc(pk, tcp)
torch.cuda.synchronize()
buf = (ctypes.c_longlong * 16)()
fn = _cabi.lib.xb_dbg_conv_clocks
fn.argtypes = [ctypes.c_void_p]
assert fn(buf) == 0
v = list(buf)
np_, ne = max(v[7], 1), max(v[15], 1)
print("issue warp: tiles %d | per tile: loop %d  wait-operand %d  wait-empty %d  issue %d" % (v[7], v[0] / np_, v[1] / np_, v[2] / np_, v[3] / np_))
print("epilogue group 0: tiles seen %d | per tile of the CTA: loop %d  wait-store %d  bar1 %d  wait-full %d  drain %d  bar2 %d  rest %d"
      % (v[15], v[8] / ne, v[9] / ne, v[10] / ne, v[11] / ne, v[12] / ne, v[13] / ne, v[14] / ne))
