"""Debug aid: one training step (forward + backward) of a fixed problem after the on-chip memories were filled with
zeros / NaN / a large finite pattern; every gradient must come out bit-identical."""
import ctypes, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, torch
from util import onehot_from_codes, random_codes
from explainn_b200 import networks
from oracle import torch_port as tp
lib = ctypes.CDLL("tools/libdirty.so")      # nvcc -gencode arch=compute_100a,code=sm_100a -shared -Xcompiler -fPIC -o tools/libdirty.so tools/dirty_chip.cu
dev = "cuda:0"
for (U, L, C, F, B) in [(300, 251, 81, 19, 160), (300, 251, 81, 19, 512), (40, 200, 5, 20, 64)]:
    torch.manual_seed(7)
    port = tp.PortExplaiNN(U, L, C, F)
    x = torch.from_numpy(onehot_from_codes(random_codes(B, L, 10, n_frac=0.01))).to(dev)
    y = (torch.rand(B, C) < 0.5).float().to(dev)
    base = None
    for pat in (0x00000000, 0x7FC00000, 0x7F000000, 0xFFFFFFFF, 0x3F800000, 0x00000000):
        m = networks.ExplaiNN(U, L, C, F); m.load_state_dict(port.state_dict()); m = m.to(dev).train(); m.conv_backend = "tc"
        m.linears[9].eval()
        torch.cuda.synchronize()
        assert lib.dirty_chip(ctypes.c_uint32(pat), 148) == 0
        out = m(x)
        torch.nn.BCEWithLogitsLoss()(out, y).backward()
        cur = {"logits": out.detach().clone()}
        for k, p in m.named_parameters():
            cur[k] = p.grad.clone()
        if base is None:
            base = cur
        bad = [(k, float((cur[k] - base[k]).abs().max()), int((cur[k] != base[k]).sum())) for k in base
               if not torch.equal(cur[k], base[k]) and not k.startswith("final.")]
        print((U, L, C, F, B), "pattern %08x" % pat, "MISMATCH " + str(bad[:6]) if bad else "identical", flush=True)
