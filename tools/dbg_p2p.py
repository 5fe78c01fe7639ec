"""torchrun --nproc-per-node 2 tools/dbg_p2p.py : step-by-step check of the peer-memory plumbing and the fused kernel."""
import ctypes, os, sys, torch, torch.distributed as dist
sys.path.insert(0, ".")
from explainn_b200 import parallel, _cabi
from explainn_b200._cabi import call, ptr, stream
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device("cuda:%d" % rank)
dist.init_process_group("nccl", device_id=dev)
n = 1 << 20
pg = parallel.PeerGradients(n, dev)
g = pg.flat_grad
print(rank, "opened", [hex(p) for p in pg._opened], flush=True)
p = torch.zeros(n, device=dev); m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev)
step = torch.zeros(1, device=dev)
for s in range(1, 4):
    g.fill_(float(rank + 1) * s)
    call("explainn_p2p_adam_step", ptr(p), ptr(m), ptr(v), n, ptr(step), ctypes.cast(pg.g_array, ctypes.c_void_p),
         ctypes.cast(pg.f_array, ctypes.c_void_p), world, rank, s, 1e-3, 0.9, 0.999, 1e-8, stream())
    torch.cuda.synchronize()
    print(rank, "step", s, "p[0]", float(p[0]), "m[0]", float(m[0]), "flags", pg.flags[:20].tolist(), flush=True)
dist.barrier()
dist.destroy_process_group()
