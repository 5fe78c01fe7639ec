import sys, numpy as np, torch
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
from util import load_golden, onehot_from_codes, state_dict_from
from explainn_b200 import networks, functional as Fn
for name in ["eval_u40_f20", "eval_cfg1_ckpt", "eval_u7_f8", "eval_u33_f24"]:
    g = load_golden(name)
    m = networks.ExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]))
    m.load_state_dict(state_dict_from(g)); m = m.cuda().eval()
    x = torch.from_numpy(onehot_from_codes(g["codes"])).cuda()
    pk = m._pack(x)
    conv, bn = m.linears[0], m.linears[1]
    ref = Fn.conv_pool_fwd(pk, m._conv_prep_eval())
    tcp = Fn.conv_tc_prepare(conv.weight.detach(), conv.bias.detach(), bn.weight.detach(), bn.bias.detach(), bn.running_mean, bn.running_var, bn.eps)
    out = Fn.conv_pool_fwd_tc(pk, tcp)
    torch.cuda.synchronize()
    rel = ((out - ref).abs() / ref.abs().clamp_min(1e-30)).max().item()
    print(name, "max rel diff TC vs LUT:", rel, "shape", tuple(out.shape), "nan", torch.isnan(out).any().item())
    if rel > 1e-4:
        d = ((out - ref).abs() / ref.abs().clamp_min(1e-30))
        i = torch.nonzero(d > 1e-4)[:8]
        print(i.tolist()); print(out[tuple(i[0])].item(), ref[tuple(i[0])].item())
        print(out[0,0,:8], ref[0,0,:8])
