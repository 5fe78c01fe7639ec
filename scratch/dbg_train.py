import sys, numpy as np, torch
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
from util import load_golden, onehot_from_codes, state_dict_from
from explainn_b200 import networks
g = load_golden("train_u6_bce")
m = networks.ExplaiNN(int(g["U"]), int(g["L"]), int(g["C"]), int(g["F"]))
m.load_state_dict(state_dict_from(g, "sd0.")); m = m.cuda().train(); m.linears[9].eval()
x = torch.from_numpy(onehot_from_codes(g["codes"])).cuda(); y = torch.from_numpy(g["labels"]).cuda()
out = m(x); loss = torch.nn.BCEWithLogitsLoss()(out, y); loss.backward()
for k, p in m.named_parameters():
    ref = g["grad." + k]; got = p.grad.cpu().numpy()
    print(k, "maxerr/scale", np.abs(got - ref).max() / max(np.abs(ref).max(), 1e-30))
print("dbeta got", m.linears[1].bias.grad.cpu().numpy()); print("dbeta ref", g["grad.linears.1.bias"])
print("dgamma got", m.linears[1].weight.grad.cpu().numpy()); print("dgamma ref", g["grad.linears.1.weight"])
