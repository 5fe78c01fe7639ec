"""Per-pass timing of the interpretation sweep at cfg 4 (num_cnns=300, L=251): tensor-core vs table backend,
N-free and 1 % N sequences, sparse (BatchNorm gain 4) and dense (default init) hit regimes."""
import sys, torch
sys.path.insert(0, ".")
from explainn_b200 import networks, interpretation, functional as Fn
U, L, C, F, B = 300, 251, 81, 19, 32768
dev = "cuda:0"
g = torch.Generator(device=dev).manual_seed(1)
codes = torch.randint(0, 4, (B, L), generator=g, device=dev)
lut = torch.tensor(list(b"ACGT"), dtype=torch.uint8, device=dev)
text = lut[codes]
textn = text.clone(); textn[torch.rand((B, L), generator=g, device=dev) < 0.01] = ord("N")
def timeit(f, n=3):
    f(); torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): f()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / n
for regime in ("sparse", "dense"):
    torch.manual_seed(0)
    m = networks.ExplaiNN(U, L, C, F).to(dev).eval()
    if regime == "sparse":
        m.linears[1].weight.data.fill_(4.0)
    for name, t in (("N-free", text), ("1% N", textn)):
        pk = Fn.encode_dna_array(t, L, dev)
        for backend in ("tc", "lut"):
            acc = interpretation.PwmAccumulator(m, backend=backend)
            def pa():
                acc.zmax = None
                acc.scan_max(pk)
            ta = timeit(pa)
            def pa_warm():
                acc.scan_max(pk)             # zmax already holds the maximum: the steady state of a chunked sweep
            taw = timeit(pa_warm)
            acc.finish_max()
            def pb():
                acc.counts = acc.hits = None
                acc.scan_count(pk)
            tb = timeit(pb)
            extra = ""
            if backend == "tc":
                extra = " resolved(last B)=%d" % int(acc.tc.list[:4].view(torch.int32).item())
                acc.tc.check()
            print("%s %s %s: pass A cold %.2f ms, warm %.2f ms (%.1f M seq/s)  pass B %.2f ms (%.1f M seq/s)  both %.1f M seq/s  hits %d%s"
                  % (regime, name, backend, ta, taw, B / taw / 1e3, tb, B / tb / 1e3, B / (taw + tb) / 1e3, int(acc.hits.sum()), extra))
