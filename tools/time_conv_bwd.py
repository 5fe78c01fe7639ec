"""Times explainn_conv_pool_bwd at the cfg-2 training shape (batch 512, 300 units, L 251, filter 19) with the
tensor-core scatter (one `partial` slice per SM) and with the shared-memory scatter kernel (3 slices).
CUDA events around 5 replays of a CUDA graph of 20 launches (no launch gaps)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))


def main():
    from explainn_b200 import _cabi
    from explainn_b200 import functional as Fn
    from util import onehot_from_codes, random_codes
    dev = "cuda:0"
    shapes = [(512, 300, 251, 19), (128, 300, 251, 19), (2048, 300, 251, 19), (512, 100, 200, 19)]
    if len(sys.argv) > 1:
        shapes = shapes[:int(sys.argv[1])]
    for B, U, L, F in shapes:
        W = 4 * F
        S = L - F + 1
        P = (S - 7) // 7 + 1
        g = torch.Generator().manual_seed(0)
        x = onehot_from_codes(random_codes(B, L, 1, n_frac=0.01))
        pk = Fn.pack_onehot(torch.from_numpy(x).to(dev))
        gpool = torch.randn(B, U, P, generator=g).to(dev)
        pooled = torch.exp(torch.randn(B, U, P, generator=g)).to(dev)
        xhat = torch.randn(B, U, P, generator=g).to(dev)
        idx = torch.randint(0, 7, (B, U, P), generator=g, dtype=torch.uint8).to(dev)
        M = Fn.input_cooc_stats(pk, F)
        f32 = dict(dtype=torch.float32, device=dev)
        Mw = torch.zeros(U, W, **f32)
        z, o = torch.zeros(U, **f32), torch.ones(U, **f32)
        dW, db, dg, dbe = torch.empty(U, W, **f32), torch.empty(U, **f32), torch.empty(U, **f32), torch.empty(U, **f32)
        n_tc = _cabi.lib.explainn_conv_bwd_nsplit(U)
        sms = torch.cuda.get_device_properties(0).multi_processor_count
        for name, nsplit in (("tensor-core scatter", n_tc), ("shared-memory scatter", max(1, -(-5 * sms // U)))):
            partial = torch.empty(nsplit, U, W + 2, **f32)
            for phases, label in ((1, "scatter"), (2, "final"), (3, "both")):
                def run():
                    _cabi.call("explainn_conv_pool_bwd_phase", phases, 0, _cabi.ptr(pk.codes), _cabi.ptr(pk.nmask), B, L, U, F, _cabi.ptr(gpool),
                               _cabi.ptr(pooled), _cabi.ptr(xhat), _cabi.ptr(idx), _cabi.ptr(M), _cabi.ptr(Mw), _cabi.ptr(z), _cabi.ptr(z),
                               _cabi.ptr(o), _cabi.ptr(o), 0.0, _cabi.ptr(partial), nsplit, _cabi.ptr(dW), _cabi.ptr(db), _cabi.ptr(dg),
                               _cabi.ptr(dbe), _cabi.stream())
                for _ in range(3):
                    run()
                # 20 launches per CUDA graph: no launch gaps
                st = torch.cuda.Stream()
                st.wait_stream(torch.cuda.current_stream())
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.stream(st):
                    run()
                    with torch.cuda.graph(graph, stream=st):
                        for _ in range(20):
                            run()
                torch.cuda.synchronize()
                graph.replay()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda.synchronize()
                s.record()
                for _ in range(5):
                    graph.replay()
                e.record()
                torch.cuda.synchronize()
                print("B=%d U=%d L=%d F=%d  %-22s %-8s %8.2f us" % (B, U, L, F, name, label, s.elapsed_time(e) / 100 * 1e3), flush=True)


if __name__ == "__main__":
    main()
