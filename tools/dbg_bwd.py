import os, sys, numpy as np, torch
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
got = m.linears[6].weight.grad.cpu().numpy().reshape(6, 100, -1); ref = g["grad.linears.6.weight"].reshape(6, 100, -1)
print("dW1 got[0,:3,:6]\n", got[0, :3, :6], "\nref\n", ref[0, :3, :6])
gw = m.linears[0].weight.grad.cpu().numpy(); rw = g["grad.linears.0.weight"]
print("conv got", gw[0, 0, :5], "ref", rw[0, 0, :5])
