#!/usr/bin/env python
"""bench.py -- throughput of the ExplaiNN hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A "step" is one optimisation step of `train.train_explainn`'s inner loop (train.py:55-73 of the reference:
forward, BCEWithLogits loss, backward, Adam) over one batch of 512 synthetic one-hot sequences PER GPU at
BASELINE.json configs[1] (AI-TAC style: num_cnns=300, L=251, 81 outputs).  One JSON line is printed by rank 0:

  value     whole-job training sequences/s with the fp32 one-hot dataset resident in HBM (dataset larger than L2,
            cycled), through the public FusedTrainer.step API (the path train_explainn takes)
  e2e       same metric with every batch starting in pinned HOST memory: H2D of the (B,4,L) fp32 batch and its
            labels, the step, and a D2H read of the loss, all inside the timed region
  roofline  fused conv+BN+exp+maxpool forward kernel: algorithmic bytes / CUDA-event time vs measured HBM peak
  cpu_baseline  the oracle's torch CPU port of the reference step, timed on this box's host cores
  infer / interp  eval-forward and PWM-extraction throughput at the same model shape (extra keys)

`--impl reference` times only the CPU port (the reference is pure Python/PyTorch; /root/reference is not
present on the GPU box), same metric/config keys.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = dict(U=300, L=251, C=81, F=19, B=512)      # BASELINE.json configs[1]
METRIC = "train_sequences_per_sec"
UNIT = "sequences/s"


def config_dict(n_gpus, extra=None):
    d = {"workload": "cfg2 AI-TAC: num_cnns=300 L=251 filter=19 classes=81, BCEWithLogits + Adam(lr=1e-3), "
                     "batch 512 per GPU, dropout 0.3 on",
         "global_batch": CFG["B"] * n_gpus, "per_gpu_batch": CFG["B"], "parallelism": "dp%d" % n_gpus,
         "l2_policy": "input dataset (40960 seq, 164 MB fp32 one-hot) larger than L2, cycled"}
    if extra:
        d.update(extra)
    return d


# ------------------------------------------------------------------------------------------------ CPU baseline
def cpu_port_step_throughput(batch, steps, warmup, threads=None):
    """Reference training step (oracle/torch_port.py = the reference's own torch.nn graph) on host cores."""
    import torch
    from oracle import torch_port as tp
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    torch.manual_seed(0)
    model = tp.PortExplaiNN(CFG["U"], CFG["L"], CFG["C"], CFG["F"])
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    crit = torch.nn.BCEWithLogitsLoss()
    x = tp.synthetic_onehot(batch, CFG["L"], seed=1714)
    y = (torch.rand(batch, CFG["C"], generator=torch.Generator().manual_seed(1)) < 0.5).float()
    for _ in range(warmup):
        tp.train_step(model, x, y, crit, opt)
    t0 = time.perf_counter()
    for _ in range(steps):
        loss = tp.train_step(model, x, y, crit, opt)
        float(loss)
    dt = time.perf_counter() - t0
    return batch * steps / dt, dt / steps, threads


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # bounded sample: a 128-sequence slice of the 512-sequence batch per step
    sample_b = 128
    v, sec, threads = cpu_port_step_throughput(sample_b, max(1, args.steps), max(1, args.warmup))
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(args.gpus, {"reference_sample": "%d-sequence slice of the 512 batch per step, CPU" % sample_b}),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": "%d steps x %d sequences (slice of the 512-sequence cfg2 batch), torch CPU, "
                                       "oracle/torch_port.py" % (args.steps, sample_b)},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [c.strip() for c in r.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-extras", action="store_true", help="skip the infer/interp/cpu_baseline/breakdown legs")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n_gpus = world

    import __graft_entry__
    if rank == 0:
        __graft_entry__.build()
    if world > 1:
        dist.barrier()
    from explainn_b200 import _cabi, interpretation, networks, train as xtrain
    from explainn_b200 import functional as Fn

    U, L, C, F, B = CFG["U"], CFG["L"], CFG["C"], CFG["F"], CFG["B"]
    W, K = max(3, args.warmup), max(1, args.steps)
    torch.manual_seed(1714 + rank)
    NB = 80                                            # 80 batches x 512 = 40960 sequences, 164 MB fp32 > L2
    gen = torch.Generator(device=dev).manual_seed(1714 + rank)
    codes = torch.randint(0, 4, (NB * B, L), generator=gen, device=dev)
    data = torch.zeros((NB * B, 4, L), dtype=torch.float32, device=dev)
    data.scatter_(1, codes.unsqueeze(1), 1.0)
    labels = (torch.rand((NB * B, C), generator=gen, device=dev) < 0.5).float()
    del codes

    torch.manual_seed(0)                                # identical initial weights on every rank
    model = networks.ExplaiNN(U, L, C, F).to(dev)
    model.train()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    crit = torch.nn.BCEWithLogitsLoss()
    assert xtrain.FusedTrainer.supports(model, crit, opt)
    trainer = xtrain.FusedTrainer(model, crit, opt, dropout_seed=1234 + rank)

    def batch(i):
        j = (i % NB) * B
        return data[j:j + B], labels[j:j + B]

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident timing ----------------------------------------------------------------------
    for i in range(W):
        trainer.step(*batch(i))
    sync_all()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = _cabi.LAUNCHES
    e0.record()
    for i in range(K):
        loss = trainer.step(*batch(W + i))
    e1.record()
    sync_all()
    launches = _cabi.LAUNCHES - l0
    ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    # keep the GPU busy a little longer so that nvidia-smi (100 ms period) gets samples under load; the count is
    # derived from the all-reduced time, so every rank issues the same number of collectives
    extra_steps = max(20, int(0.6 / max(ms / K * 1e-3, 1e-5)))
    for i in range(extra_steps):
        trainer.step(*batch(i))
    torch.cuda.synchronize()
    clocks = sampler.stop() if rank == 0 else None
    value = n_gpus * B * K / (ms * 1e-3)
    final_loss = float(loss.item())

    # ---- end-to-end timing: pinned host -> device every step, loss read back every step ---------------
    hx = torch.empty((8 * B, 4, L), dtype=torch.float32).pin_memory()
    hy = torch.empty((8 * B, C), dtype=torch.float32).pin_memory()
    hx.copy_(data[:8 * B]); hy.copy_(labels[:8 * B])
    Ke = max(10, K // 2)
    for i in range(3):
        float(trainer.step(hx[:B].to(dev, non_blocking=True), hy[:B].to(dev, non_blocking=True)).item())
    sync_all()
    e0.record()
    for i in range(Ke):
        j = (i % 8) * B
        x = hx[j:j + B].to(dev, non_blocking=True)
        y = hy[j:j + B].to(dev, non_blocking=True)
        lv = float(trainer.step(x, y).item())
    e1.record()
    sync_all()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t.item())
    e2e = {"value": n_gpus * B * Ke / (e2e_ms * 1e-3), "unit": UNIT,
           "h2d_bytes_per_step": B * 4 * L * 4 + B * C * 4, "d2h_bytes_per_step": 4, "steps": Ke}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config_dict(n_gpus), "clocks": clocks, "e2e": e2e,
            "gpu_launches": launches, "final_loss": final_loss, "cuda_graph": trainer._graph is not None}

    if not args.no_extras:
        # every rank takes part in the per-kernel breakdown (its eager steps contain the gradient all-reduce);
        # the remaining legs use no collective and run on rank 0 only
        line.update(extras(torch, dev, model, trainer, data, labels, batch, _cabi, Fn, interpretation, rank, world))
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def extras(torch, dev, model, trainer, data, labels, batch, _cabi, Fn, interpretation, rank=0, world=1):
    """Roofline of the fused conv kernel, per-kernel breakdown of the step, inference / interpretation
    throughput, CPU baseline.  Rank 0 only, outside the headline timed region."""
    out = {}
    U, L, C, F, B = CFG["U"], CFG["L"], CFG["C"], CFG["F"], CFG["B"]
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    P = model.pooled_length
    ev = lambda: torch.cuda.Event(enable_timing=True)

    # ---- per-kernel breakdown of one training step (eager, events around every C-ABI call) -----------
    real_call = _cabi.call
    from explainn_b200 import autograd_fn, functional, train as xtrain
    records = []

    def timed_call(name, *a):
        s, e = ev(), ev()
        s.record()
        rc = real_call(name, *a)
        e.record()
        records.append((name, s, e))
        return rc
    for mod in (_cabi, autograd_fn, functional, xtrain):
        mod.call = timed_call
    trainer.use_graph, g_saved = False, trainer._graph
    trainer._graph = None
    nprof = 10
    for i in range(nprof):
        trainer.step(*batch(i + 7))
    torch.cuda.synchronize()
    for mod in (_cabi, autograd_fn, functional, xtrain):
        mod.call = real_call
    trainer._graph, trainer.use_graph = g_saved, True
    if rank != 0:
        return out
    agg = {}
    for name, s, e in records:
        agg[name] = agg.get(name, 0.0) + s.elapsed_time(e)
    breakdown = {k.replace("explainn_", ""): round(v / nprof * 1e3, 2) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])}
    out["kernel_breakdown_us_per_step"] = breakdown
    conv_key = "explainn_conv_pool_fwd_train_tc" if "explainn_conv_pool_fwd_train_tc" in agg else "explainn_conv_pool_fwd_train"
    conv_ms = agg.get(conv_key, 0.0) / nprof

    # ---- roofline: fused conv+BN+exp+pool forward (training variant, the one inside the step) ---------
    bytes_train = B * ((L + 3) // 4 + (L + 7) // 8 + 9 * U * P)          # codes + mask + pooled,xhat fp32 + idx u8
    if conv_ms > 0:
        ach = bytes_train / (conv_ms * 1e-3) / 1e9
        kname = ("conv_tc_kernel<TRAIN> (tcgen05 implicit-GEMM fused conv+BN+exp+maxpool fwd, training variant)"
                 if conv_key.endswith("_tc") else "conv_scan_kernel<PoolTrainSink> (k-mer table fused conv+BN+exp+maxpool fwd)")
        out["roofline"] = {"kernel": kname,
                           "bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                           # dram__bytes_read.sum + dram__bytes_write.sum of one launch at this batch size, from the
                           # ncu --set full capture summarised in profiles/r1_ncu_full_summary_final.txt: the 45.7 MB the
                           # kernel writes are still resident in the 126 MB L2 when it ends
                           "traffic": 425728 if conv_key.endswith("_tc") and B == 512 and U == 300 else None,
                           "traffic_source": "profiles/r1_ncu_full_summary_final.txt",
                           "peak_source": peak_src, "avg_launch_us": conv_ms * 1e3,
                           "algorithmic_bytes_per_launch": bytes_train,
                           "note": "HBM fraction as BASELINE.json asks; the kernel's own ceiling is the tensor pipe "
                                   "(2*U*S*4F*2 fp16 flop per sequence), see DESIGN.md"}

    # ---- inference: eval forward on resident data, large batches --------------------------------------
    model.eval()
    IB = 8192
    with torch.no_grad():
        for i in range(2):
            model(data[:IB])
        torch.cuda.synchronize()
        s, e = ev(), ev()
        n_it = 5
        s.record()
        for i in range(n_it):
            o = model(data[i * IB:(i + 1) * IB])
        e.record()
        torch.cuda.synchronize()
        inf_ms = s.elapsed_time(e) / n_it
        # the inference conv kernel alone (roofline of the kernel BASELINE.json names)
        pk = model._pack(data[:IB])
        conv, bn = model.linears[0], model.linears[1]
        tcp = Fn.conv_tc_prepare(conv.weight.detach(), conv.bias.detach(), bn.weight.detach(), bn.bias.detach(),
                                 bn.running_mean, bn.running_var, bn.eps)
        Fn.conv_pool_fwd_tc(pk, tcp)
        torch.cuda.synchronize()
        s.record()
        for i in range(5):
            Fn.conv_pool_fwd_tc(pk, tcp)
        e.record()
        torch.cuda.synchronize()
        k_ms = s.elapsed_time(e) / 5
    bytes_inf = IB * ((L + 3) // 4 + (L + 7) // 8 + 4 * U * P)
    ach = bytes_inf / (k_ms * 1e-3) / 1e9
    out["infer"] = {"sequences_per_sec": IB / (inf_ms * 1e-3), "batch": IB, "ms_per_batch": inf_ms,
                    "conv_pool_kernel": {"name": "conv_tc_kernel<INFER>", "avg_launch_ms": k_ms, "sequences_per_sec": IB / (k_ms * 1e-3),
                                         "achieved_GBps": ach, "frac_of_hbm_peak": ach / hbm_peak,
                                         "algorithmic_bytes_per_launch": bytes_inf}}

    # ---- inference at the large-scale shape (BASELINE.json config 5: num_cnns=500, L=1000, 100 outputs) ----------
    try:
        from explainn_b200 import networks as xnet5
        torch.manual_seed(0)
        m5 = xnet5.ExplaiNN(500, 1000, 100, 19).to(dev).eval()
        g5 = torch.Generator(device=dev).manual_seed(7)
        c5 = torch.randint(0, 4, (2048, 1000), generator=g5, device=dev)
        x5 = torch.zeros((2048, 4, 1000), device=dev)
        x5.scatter_(1, c5.unsqueeze(1), 1.0)
        with torch.no_grad():
            m5(x5)
            torch.cuda.synchronize()
            s.record()
            for i in range(5):
                m5(x5)
            e.record()
            torch.cuda.synchronize()
        out["infer_cfg5"] = {"sequences_per_sec": 2048 / (s.elapsed_time(e) / 5 * 1e-3), "batch": 2048,
                             "workload": "num_cnns=500 L=1000 filter=19 classes=100, eval forward from fp32 one-hot on the device"}
        del m5, x5, c5
    except Exception as ex:       # pragma: no cover
        out["infer_cfg5"] = {"error": str(ex)}

    # ---- interpretation: fused two-pass PWM extraction -------------------------------------------------
    NI = 4 * IB
    with torch.no_grad():
        acc = interpretation.PwmAccumulator(model)
        acc.scan_max(data[:IB]); acc.finish_max(); acc.scan_count(data[:IB])
        torch.cuda.synchronize()
        s.record()
        acc = interpretation.PwmAccumulator(model)
        for i in range(0, NI, IB):
            acc.scan_max(data[i:i + IB])
        acc.finish_max()
        for i in range(0, NI, IB):
            acc.scan_count(data[i:i + IB])
        e.record()
        torch.cuda.synchronize()
    out["interp"] = {"sequences_per_sec": NI / (s.elapsed_time(e) * 1e-3), "sequences": NI, "passes": 2,
                     "total_hits": int(acc.hits.sum().item())}
    model.train()

    # ---- input encoding and PWM scoring ----------------------------------------------
    try:
        import numpy as np
        from explainn_b200 import networks as xnet, tools as xtools
        from oracle import canonical as oc
        NE = 4 * IB
        rng = np.random.default_rng(5)
        letters = np.frombuffer(b"ACGTN", dtype=np.uint8)
        cd = rng.integers(0, 4, size=(NE, L)).astype(np.uint8)
        cd[rng.random((NE, L)) < 0.01] = 4
        seqs = [letters[c].tobytes() for c in cd]
        xtools.encode_sequences(seqs[:IB], L, dev)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(0, NE, IB):
            pk_e = xtools.encode_sequences(seqs[i:i + IB], L, dev)     # host join + H2D of the bytes + encode kernel
        torch.cuda.synchronize()
        enc_s = time.perf_counter() - t0
        t0 = time.perf_counter()
        nref = 256
        for sq in seqs[:nref]:
            oc.dna_one_hot(sq.decode("ascii"), L)
        ref_s = time.perf_counter() - t0
        pwm = xnet.PWM(rng.normal(size=(U, 4, model._options["filter_size"])).astype(np.float32), L, "max").to(dev)
        pk_s = model._pack(data[:IB])          # the same N-free sequences the other legs use
        pwm.forward_packed(pk_s)
        torch.cuda.synchronize()
        s.record()
        for i in range(5):
            pwm.forward_packed(pk_s)
        e.record()
        torch.cuda.synchronize()
        out["io"] = {"encode_sequences_per_sec_e2e": NE / enc_s, "encode_h2d_bytes_per_seq": L,
                     "encode_cpu_port_sequences_per_sec": nref / ref_s, "encode_cpu_sample": "%d sequences, 1 core" % nref,
                     "pwm_score_sequences_per_sec": IB / (s.elapsed_time(e) / 5 * 1e-3), "pwm_score_groups": U,
                     "pwm_score_strands": 2}
    except Exception as ex:       # pragma: no cover
        out["io"] = {"error": str(ex)}

    # ---- CPU baseline: the reference step on this box's host cores (bounded sample; N = 1 only) -------
    if world > 1:
        return out
    try:
        v, sec, threads = cpu_port_step_throughput(B, steps=2, warmup=1)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                               "sample": "1 warm-up + 2 timed steps of the full 512-sequence cfg2 batch, torch CPU "
                                         "(oracle/torch_port.py restates the reference module graph)",
                               "ms_per_step": sec * 1e3}
    except Exception as ex:       # pragma: no cover
        out["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": "failed: %s" % ex}
    return out


if __name__ == "__main__":
    main()
