#!/usr/bin/env python
"""bench.py -- throughput of the ExplaiNN hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--no-extras]

A "step" is one optimisation step of `train.train_explainn`'s inner loop (train.py:55-73 of the reference:
forward, BCEWithLogits loss, backward, Adam) over one batch of 512 synthetic one-hot sequences PER GPU at
BASELINE.json configs[1] (AI-TAC style: num_cnns=300, L=251, 81 outputs).  One JSON line is printed by rank 0:

  value        whole-job training sequences/s with the fp32 one-hot dataset resident in HBM (dataset larger than L2,
               cycled), through the public FusedTrainer.step API (the path train_explainn takes)
  e2e          same metric with every batch starting in pinned HOST memory as ASCII bases (1 byte per base, encoded on
               the GPU by explainn_encode_dna): H2D of sequences and labels (next batch prefetched on a copy stream),
               the step, and a D2H read of the loss, all inside the timed region; `e2e.onehot_fp32` is the same with
               the reference's fp32 (B,4,L) batches (16 bytes per base)
  roofline     fused conv+BN+exp+maxpool forward kernel of the step: SURVEY 8(d) bytes / CUDA-event time vs the
               measured HBM peak, plus the tensor-pipe view (executed / useful flops vs the measured bf16 peak)
  cpu_baseline the oracle's torch CPU port of the reference step, timed on this box's host cores
  infer_cfg5 / interp_cfg4   configs[4] (num_cnns=500, L=1000, 100 outputs, >= 1 M sequences) eval forward and
               configs[3] (1 M sequences, num_cnns=300) PWM extraction, sharded over all ranks, device-resident and e2e
  dp_check     at N > 1: replicas bit-identical after the timed steps, sharded PWMs == unsharded PWMs

`--impl reference` times only the CPU port of the step on the same 512-sequence batch (the reference is pure
Python/PyTorch; /root/reference is not present on the GPU box), same metric/config keys.
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
    """The reference's own CPU path for the same step, config and metric: the full 512-sequence batch per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    B = CFG["B"]
    v, sec, threads = cpu_port_step_throughput(B, max(1, args.steps), max(1, args.warmup))
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_dict(args.gpus),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": "%d steps x the full %d-sequence cfg2 batch, torch CPU, oracle/torch_port.py "
                                       "(one process; the reference has no multi-device path)" % (args.steps, B)},
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


LETTERS = b"ACGT"


def ascii_from_codes(torch, codes):
    """(n, L) int64 base codes 0..3 on the device -> (n, L) uint8 ASCII letters (device)."""
    lut = torch.tensor(list(LETTERS), dtype=torch.uint8, device=codes.device)
    return lut[codes]


# ------------------------------------------------------------------------------------------------ GPU arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-extras", action="store_true", help="skip the roofline/infer/interp/cpu_baseline legs")
    ap.add_argument("--big-n", type=int, default=1 << 20, help="sequences of the infer_cfg5 / interp_cfg4 legs (whole job)")
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
    from explainn_b200 import _cabi, networks, train as xtrain

    U, L, C, F, B = CFG["U"], CFG["L"], CFG["C"], CFG["F"], CFG["B"]
    W, K = max(3, args.warmup), max(1, args.steps)
    torch.manual_seed(1714 + rank)
    NB = 80                                            # 80 batches x 512 = 40960 sequences, 164 MB fp32 > L2
    gen = torch.Generator(device=dev).manual_seed(1714 + rank)
    codes = torch.randint(0, 4, (NB * B, L), generator=gen, device=dev)
    data = torch.zeros((NB * B, 4, L), dtype=torch.float32, device=dev)
    data.scatter_(1, codes.unsqueeze(1), 1.0)
    labels = (torch.rand((NB * B, C), generator=gen, device=dev) < 0.5).float()
    text_dev = ascii_from_codes(torch, codes[:8 * B])
    del codes

    torch.manual_seed(rank)                             # different initial weights per rank on purpose: the trainer
    model = networks.ExplaiNN(U, L, C, F).to(dev)       # broadcasts rank 0's (dp_check below verifies the replicas)
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

    def max_over_ranks(ms):
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

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
    ms = max_over_ranks(e0.elapsed_time(e1))
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
    Ke = max(10, K // 2)

    def e2e_run(hx, hy):
        class Cycle:                                    # a "loader" over 8 pinned host batches, Ke + 3 steps long
            def __iter__(self):
                for i in range(Ke + 3):
                    j = (i % 8) * B
                    yield hx[j:j + B], hy[j:j + B]
        it = iter(xtrain.DevicePrefetcher(Cycle(), dev))
        for _ in range(3):
            float(trainer.step(*next(it)).item())
        sync_all()
        e0.record()
        for x, y in it:
            float(trainer.step(x, y).item())
        e1.record()
        sync_all()
        return max_over_ranks(e0.elapsed_time(e1))

    hy = torch.empty((8 * B, C), dtype=torch.float32).pin_memory()
    hy.copy_(labels[:8 * B])
    ht = torch.empty((8 * B, L), dtype=torch.uint8).pin_memory()
    ht.copy_(text_dev)
    e2e_ms = e2e_run(ht, hy)
    hx = torch.empty((8 * B, 4, L), dtype=torch.float32).pin_memory()
    hx.copy_(data[:8 * B])
    e2e_oh_ms = e2e_run(hx, hy)
    e2e = {"value": n_gpus * B * Ke / (e2e_ms * 1e-3), "unit": UNIT,
           "h2d_bytes_per_step": B * L + B * C * 4, "d2h_bytes_per_step": 4, "steps": Ke,
           "input": "pinned host uint8 ASCII bases (B,L) + fp32 labels, next batch prefetched on a copy stream "
                    "(train.DevicePrefetcher, the loop train_explainn runs), encoded on the GPU",
           "onehot_fp32": {"value": n_gpus * B * Ke / (e2e_oh_ms * 1e-3), "h2d_bytes_per_step": B * 4 * L * 4 + B * C * 4,
                           "input": "pinned host fp32 one-hot (B,4,L), the reference's DataLoader format"}}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": config_dict(n_gpus), "clocks": clocks, "e2e": e2e,
            "gpu_launches": launches, "final_loss": final_loss, "cuda_graph": trainer._graph is not None,
            "dp_optimizer": ("peer-memory reduce-scatter + Adam + all-gather kernel" if trainer.p2p is not None
                             else ("NCCL all-reduce + Adam" if world > 1 else "single GPU"))}

    if world > 1:
        line["dp_check"] = dp_check(torch, dist, dev, trainer, model, rank, world)
    if not args.no_extras:
        line.update(extras(torch, dev, model, trainer, data, labels, batch, _cabi, rank, world, args.big_n))
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def dp_check(torch, dist, dev, trainer, model, rank, world):
    """Self-check of the data-parallel paths on the hardware the bench runs on: (1) after all the steps above every
    rank holds bit-identical parameters and BatchNorm-independent state; (2) PWMs extracted from a dataset sharded over
    the ranks (MAX / SUM reductions) equal the unsharded extraction on rank 0, integer for integer."""
    from explainn_b200 import interpretation, parallel
    out = {}
    trainer.check_health()
    trainer.sync_buffers()                                           # every rank evaluates the same model below
    flat = trainer.flat_p.clone()
    ref = flat.clone()
    dist.broadcast(ref, src=0)
    same = torch.tensor([1 if torch.equal(ref.view(torch.int32), flat.view(torch.int32)) else 0], device=dev)
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    out["replicas_bit_identical"] = bool(int(same.item()))
    was_training = model.training
    model.eval()
    n = 4096 * world
    g = torch.Generator(device=dev).manual_seed(99)                  # same dataset on every rank
    codes = torch.randint(0, 5, (n, model._options["input_length"]), generator=g, device=dev)
    x = torch.zeros((n, 5, codes.shape[1]), device=dev).scatter_(1, codes.unsqueeze(1), 1.0)[:, :4].contiguous()
    pwm_s, counts_s, hits_s = parallel.sharded_pwms(x, model, dev)
    ok = True
    if rank == 0:
        pwm_1, counts_1, hits_1 = interpretation.get_pwms_explainn_fused(x, model, dev, return_counts=True)
        ok = bool((counts_s == counts_1).all() and (hits_s == hits_1).all() and (pwm_s == pwm_1).all())
        out["sharded_pwm_hits"] = int(hits_1.sum())
    t = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(t, src=0)
    out["sharded_pwms_equal_unsharded"] = bool(int(t.item()))
    model.train(was_training)
    out["status"] = "ok" if out["replicas_bit_identical"] and out["sharded_pwms_equal_unsharded"] else "FAILED"
    return out


def _peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def leg_breakdown(torch, trainer, batch, _cabi):
    """Eager CUDA events around every C-ABI call of ten training steps.  The events include the launch gaps of eager
    execution, so the entries add up to MORE than the CUDA-graph `ms_per_step`: use them as shares, not absolutes."""
    from explainn_b200 import autograd_fn, functional, train as xtrain
    ev = lambda: torch.cuda.Event(enable_timing=True)
    real_call = _cabi.call
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
    agg = {}
    for name, s, e in records:
        agg[name] = agg.get(name, 0.0) + s.elapsed_time(e)
    return {"timing": "eager CUDA events around each C-ABI call, launch gaps included (shares, not absolutes)",
            "us_per_step": {k.replace("explainn_", ""): round(v / nprof * 1e3, 2) for k, v in sorted(agg.items(), key=lambda kv: -kv[1])}}


def leg_conv_roofline(torch, dev, model, batch, hbm_peak, peak_src, peaks):
    """The fused conv+BN+exp+maxpool forward kernel exactly as the training step launches it (batch 512, batch
    statistics), timed alone with CUDA events on its launch stream: 50 back-to-back launches after 5 warm-ups."""
    from explainn_b200 import autograd_fn, functional as Fn
    U, L, C, F, B = CFG["U"], CFG["L"], CFG["C"], CFG["F"], CFG["B"]
    P = model.pooled_length
    S = L - F + 1
    x, _ = batch(3)
    cw, cb, g1, b1n = [p.detach() for p in autograd_fn._params(model)[:4]]
    bn1 = model.linears[1]
    pk = model._pack(x)
    M = Fn.input_cooc_stats(pk, F)
    rm, rv, nbt = bn1.running_mean.clone(), bn1.running_var.clone(), bn1.num_batches_tracked.clone()
    mean1, var1, _ = Fn.bn1_batch_stats(M, cw, cb, B * S, bn1.momentum, rm, rv, nbt)
    ones, zeros = torch.ones(U, device=dev), torch.zeros(U, device=dev)
    tcp = Fn.conv_tc_prepare(cw, cb, ones, zeros, mean1, var1, bn1.eps, sign_of=g1)
    for _ in range(5):
        Fn.conv_pool_fwd_train_tc(pk, tcp, g1, b1n)
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 50
    s.record()
    for _ in range(n):
        Fn.conv_pool_fwd_train_tc(pk, tcp, g1, b1n)
    e.record()
    torch.cuda.synchronize()
    us = s.elapsed_time(e) / n * 1e3
    # packed codes + N mask in, fp32 pooled + u8 argmax out; the kernel also writes x_hat at the maximum
    # (fp32, same shape) for the backward -- a design choice of this implementation, listed as overhead
    alg = B * ((L + 3) // 4 + (L + 7) // 8 + 4 * U * P + U * P)
    overhead = B * 4 * U * P
    ach = alg / (us * 1e-6) / 1e9
    n_mt, K16 = (U + 127) // 128, (4 * F + 15) // 16
    executed = B * n_mt * K16 * 2 * (2 * 128 * 240 * 16)         # hi and lo pass, M=128 N=240 K=16 per MMA
    useful = B * 2 * U * S * 4 * F                               # dense-equivalent flops of the conv itself
    tpeak = float(peaks.get("bf16_tflops", 0.0)) or None
    traffic, tsrc = None, None
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r2_conv_tc_train_traffic.json")))
        if t.get("B") == B and t.get("U") == U and t.get("L") == L:
            traffic, tsrc = t["dram_bytes_per_launch"], "profiles/r2_conv_tc_train_traffic.json (" + t.get("source", "ncu --set full") + ")"
    except Exception:
        pass
    return {"kernel": "conv_tc_kernel<TRAIN> (tcgen05 implicit-GEMM fused conv+BN+exp+maxpool forward, training variant)",
            "bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak, "traffic": traffic,
            "traffic_source": tsrc, "peak_source": peak_src, "avg_launch_us": us, "launches_timed": n,
            "algorithmic_bytes_per_launch": alg, "algorithmic_bytes_per_sequence": alg // B,
            "traffic_overhead_bytes_per_launch": overhead,
            "timing": "CUDA events around %d back-to-back launches of the kernel alone on the launch stream; the 45 MB "
                      "it writes stay in the 126 MB L2, inputs are 48 KB" % n,
            "tensor": {"executed_flops": executed, "useful_flops": useful, "achieved_tflops": executed / (us * 1e-6) / 1e12,
                       "useful_tflops": useful / (us * 1e-6) / 1e12, "peak": tpeak, "unit": "TFLOP/s",
                       "frac": (executed / (us * 1e-6) / 1e12 / tpeak) if tpeak else None,
                       "note": "fp16 MMAs (bf16-class rate); weights split hi+lo, M padded 300->384, N 233->240, K 76->80"},
            "note": "HBM fraction as BASELINE.json asks; the kernel's own limits are tensor-memory reads and "
                    "instruction issue of the pooling epilogue (DESIGN.md section 3)"}


def _ascii_pool(torch, dev, n, L, seed, n_frac=0.0):
    """n random sequences as ASCII text: (device uint8 (n,L), pinned host copy)."""
    g = torch.Generator(device=dev).manual_seed(seed)
    codes = torch.randint(0, 4, (n, L), generator=g, device=dev)
    text = ascii_from_codes(torch, codes)
    if n_frac > 0:
        text[torch.rand((n, L), generator=g, device=dev) < n_frac] = ord("N")
    host = torch.empty((n, L), dtype=torch.uint8).pin_memory()
    host.copy_(text)
    return text, host


def leg_infer_cfg5(torch, dist, dev, rank, world, n_total):
    """BASELINE.json configs[4]: num_cnns=500, L=1000, 100 outputs, n_total sequences sharded over the ranks
    (scripts/fasta2predictions.py:148-201 without the strand loop).  `device`: packed sequences resident, eval forward
    in chunks of 8192, logits kept on the device.  `e2e`: every chunk starts as ASCII text in pinned host memory (H2D on a
    copy stream one chunk ahead, encode on the GPU), the shards' logits are all-gathered in input order and rank 0
    copies the (n_total, 100) result to the host."""
    from explainn_b200 import functional as Fn, networks, parallel, train as xtrain
    U, L, C, F, CH = 500, 1000, 100, 19, 8192
    lo, hi = parallel.shard_range(n_total, rank, world)
    n = hi - lo
    torch.manual_seed(0)
    m5 = networks.ExplaiNN(U, L, C, F).to(dev).eval()
    pool_n = min(n, 4 * CH)
    text, host = _ascii_pool(torch, dev, pool_n, L, 7 + rank, n_frac=0.002)
    packs = [Fn.encode_dna_array(text[i:i + CH], L, dev) for i in range(0, pool_n, CH)]
    out = torch.empty((n, C), dtype=torch.float32, device=dev)
    ev = lambda: torch.cuda.Event(enable_timing=True)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def chunks():
        for i, o in enumerate(range(0, n, CH)):
            yield i % len(packs), o, min(CH, n - o)

    with torch.no_grad():
        m5.forward_packed(packs[0])
        sync_all()
        s, e = ev(), ev()
        s.record()
        for j, o, c in chunks():
            out[o:o + c] = m5.forward_packed(packs[j])[:c]
        e.record()
        sync_all()
        t = torch.tensor([s.elapsed_time(e)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms = float(t.item())

        class Loader:
            def __iter__(self):
                for j, o, c in chunks():
                    yield host[j * CH:j * CH + c], host[:1, :1]        # (sequences, dummy labels)
        result_host = torch.empty((n_total, C), dtype=torch.float32).pin_memory() if rank == 0 else None
        sync_all()
        s.record()
        o = 0
        for x, _ in xtrain.DevicePrefetcher(Loader(), dev):
            c = x.shape[0]
            out[o:o + c] = m5(x)
            o += c
        full = parallel.gather_rows(out, n_total)
        if rank == 0:
            result_host.copy_(full, non_blocking=True)
        e.record()
        sync_all()
        t = torch.tensor([s.elapsed_time(e)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    m5.check_inputs()
    finite = bool(torch.isfinite(out).all())
    del m5, out, packs, text
    torch.cuda.empty_cache()
    return {"workload": "cfg5 fasta2predictions: num_cnns=500 L=1000 filter=19 classes=100, eval forward, %d sequences "
                        "over %d GPU(s) in chunks of %d (pool of %d unique sequences per rank, cycled)" % (n_total, world, CH, pool_n),
            "sequences": n_total, "sequences_per_sec": n_total / (dev_ms * 1e-3), "ms": dev_ms,
            "e2e": {"sequences_per_sec": n_total / (e2e_ms * 1e-3), "ms": e2e_ms, "h2d_bytes_per_sequence": L,
                    "d2h_bytes_per_sequence": 4 * C, "gather": "all_gather of the logits shards in input order, D2H on rank 0"},
            "finite": finite}


def leg_interp_cfg4(torch, dist, dev, rank, world, n_total):
    """BASELINE.json configs[3]: unit max + PWM extraction for num_cnns=300 (L=251, filter 19) over n_total sequences
    sharded over the ranks (interpretation.py:126-149,176-227).  Pass A (per-unit max) -> all-reduce MAX -> thresholds ->
    pass B (counts) -> all-reduce SUM.  BatchNorm gain 4 puts the model in the sparse-hit regime of a trained model; the
    `dense` entry repeats the sweep on 1/16 of the data with default-initialised weights (99 % of the (sequence, unit)
    pairs hit)."""
    from explainn_b200 import functional as Fn, interpretation, networks, parallel, train as xtrain
    U, L, C, F, CH = 300, 251, 81, 19, 32768
    lo, hi = parallel.shard_range(n_total, rank, world)
    n = hi - lo
    pool_n = min(n, 4 * CH)
    text, host = _ascii_pool(torch, dev, pool_n, L, 11 + rank, n_frac=0.01)
    packs = [Fn.encode_dna_array(text[i:i + CH], L, dev) for i in range(0, pool_n, CH)]
    ev = lambda: torch.cuda.Event(enable_timing=True)

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def maxed(ms):
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def chunks(nn):
        for i, o in enumerate(range(0, nn, CH)):
            yield i % len(packs), min(CH, nn - o)

    def sub(pk, c):
        return pk if c == pk.B else Fn.Packed(pk.codes[:c], pk.nmask[:c], c, pk.L, pk.invalid)

    def sweep(model, nn, from_host):
        acc = interpretation.PwmAccumulator(model)
        kept = []
        if from_host:
            class Loader:
                def __iter__(self):
                    for j, c in chunks(nn):
                        yield host[j * CH:j * CH + c], host[:1, :1]
            for x, _ in xtrain.DevicePrefetcher(Loader(), dev):
                pk = model._pack(x)
                kept.append(pk)                                  # 96 bytes per sequence stay on the device for pass B
                acc.scan_max(pk)
        else:
            for j, c in chunks(nn):
                pk = sub(packs[j], c)
                kept.append(pk)
                acc.scan_max(pk)
        parallel.allreduce_unit_max(acc.zmax)
        acc.finish_max()
        for pk in kept:
            acc.scan_count(pk)
        parallel.allreduce_counts(acc.counts, acc.hits)
        pwm, counts, hits = acc.result()                         # D2H of counts / hits, PWMs on the host
        return pwm, hits

    def timed(model, nn, from_host):
        sync_all()
        s, e = ev(), ev()
        s.record()
        pwm, hits = sweep(model, nn, from_host)
        e.record()
        sync_all()
        return maxed(s.elapsed_time(e)), int(hits.sum())

    res = {}
    with torch.no_grad():
        torch.manual_seed(0)
        m4 = networks.ExplaiNN(U, L, C, F).to(dev).eval()
        m4.linears[1].weight.data.fill_(4.0)
        sweep(m4, min(n, CH), False)
        dev_ms, hits = timed(m4, n, False)
        e2e_ms, _ = timed(m4, n, True)
        res = {"workload": "cfg4 interpretation sweep: num_cnns=300 L=251 filter=19, %d sequences over %d GPU(s), two passes "
                           "(pool of %d unique sequences per rank, cycled), BatchNorm gain 4 (sparse hits)" % (n_total, world, pool_n),
               "sequences": n_total, "sequences_per_sec": n_total / (dev_ms * 1e-3), "ms": dev_ms, "total_hits": hits,
               "hits_per_sequence_unit": hits / (n_total * U),
               "e2e": {"sequences_per_sec": n_total / (e2e_ms * 1e-3), "ms": e2e_ms, "h2d_bytes_per_sequence": L,
                       "d2h_bytes": U * 4 * F * 8 + U * 8}}
        torch.manual_seed(0)
        md = networks.ExplaiNN(U, L, C, F).to(dev).eval()
        nd = max(CH // 8, n // 16)
        sweep(md, min(nd, 4096), False)
        d_ms, d_hits = timed(md, nd, False)
        res["dense"] = {"sequences": nd * world, "sequences_per_sec": nd * world / (d_ms * 1e-3), "ms": d_ms, "total_hits": d_hits,
                        "hits_per_sequence_unit": d_hits / (nd * world * U), "weights": "default initialisation"}
    del packs, text
    torch.cuda.empty_cache()
    return res


def extras(torch, dev, model, trainer, data, labels, batch, _cabi, rank=0, world=1, big_n=1 << 20):
    """Everything beside the headline: per-kernel shares of the step, roofline of the fused conv kernel, inference and
    interpretation throughput (cfg 5 / cfg 4 on ALL ranks), io leg, CPU baseline.  Outside the headline timed region."""
    import torch.distributed as dist
    from explainn_b200 import functional as Fn, interpretation
    out = {}
    U, L, C, F, B = CFG["U"], CFG["L"], CFG["C"], CFG["F"], CFG["B"]
    peaks = _peaks()
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    P = model.pooled_length
    ev = lambda: torch.cuda.Event(enable_timing=True)

    # every rank takes part in the legs that contain collectives
    bd = leg_breakdown(torch, trainer, batch, _cabi)
    try:
        cfg5 = leg_infer_cfg5(torch, dist, dev, rank, world, big_n)
    except Exception as ex:       # pragma: no cover
        cfg5 = {"error": repr(ex)}
    try:
        cfg4 = leg_interp_cfg4(torch, dist, dev, rank, world, big_n)
    except Exception as ex:       # pragma: no cover
        cfg4 = {"error": repr(ex)}
    if rank != 0:
        return out
    out["kernel_breakdown"] = bd
    out["infer_cfg5"] = cfg5
    out["interp_cfg4"] = cfg4
    out["roofline"] = leg_conv_roofline(torch, dev, model, batch, hbm_peak, peak_src, peaks)

    # ---- inference at cfg 2: eval forward on resident fp32 one-hot data, batches of 8192 ------------------
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
            model(data[i * IB:(i + 1) * IB])
        e.record()
        torch.cuda.synchronize()
        inf_ms = s.elapsed_time(e) / n_it
        # the inference conv kernel alone (the kernel BASELINE.json's 0.6-of-HBM target names)
        pk = model._pack(data[:IB])
        conv, bn = model.linears[0], model.linears[1]
        tcp = Fn.conv_tc_prepare(conv.weight.detach(), conv.bias.detach(), bn.weight.detach(), bn.bias.detach(),
                                 bn.running_mean, bn.running_var, bn.eps)
        Fn.conv_pool_fwd_tc(pk, tcp)
        torch.cuda.synchronize()
        s.record()
        for i in range(10):
            Fn.conv_pool_fwd_tc(pk, tcp)
        e.record()
        torch.cuda.synchronize()
        k_ms = s.elapsed_time(e) / 10
        # the stated-tolerance mode: one fp16 piece per weight (half the tensor-core work; tests/test_gpu_eval.py: 2e-3)
        Fn.conv_pool_fwd_tc(pk, tcp, fp16_weights=True)
        torch.cuda.synchronize()
        s.record()
        for i in range(10):
            Fn.conv_pool_fwd_tc(pk, tcp, fp16_weights=True)
        e.record()
        torch.cuda.synchronize()
        k16_ms = s.elapsed_time(e) / 10
    bytes_inf = IB * ((L + 3) // 4 + (L + 7) // 8 + 4 * U * P)
    ach = bytes_inf / (k_ms * 1e-3) / 1e9
    n_mt, K16 = (U + 127) // 128, (4 * F + 15) // 16
    executed = IB * n_mt * K16 * 2 * (2 * 128 * 240 * 16)
    tpeak = float(peaks.get("bf16_tflops", 0.0)) or None
    out["infer"] = {"sequences_per_sec": IB / (inf_ms * 1e-3), "batch": IB, "ms_per_batch": inf_ms,
                    "conv_pool_kernel": {"name": "conv_tc_kernel<INFER>", "avg_launch_ms": k_ms, "sequences_per_sec": IB / (k_ms * 1e-3),
                                         "achieved_GBps": ach, "frac_of_hbm_peak": ach / hbm_peak,
                                         "algorithmic_bytes_per_launch": bytes_inf, "executed_flops": executed,
                                         "achieved_tflops": executed / (k_ms * 1e-3) / 1e12,
                                         "frac_of_tensor_peak": (executed / (k_ms * 1e-3) / 1e12 / tpeak) if tpeak else None,
                                         "output": "325 MB of pooled activations per launch: larger than the 126 MB L2",
                                         "fp16_weight_mode": {"avg_launch_ms": k16_ms, "sequences_per_sec": IB / (k16_ms * 1e-3),
                                                              "frac_of_hbm_peak": bytes_inf / (k16_ms * 1e-3) / 1e9 / hbm_peak,
                                                              "tolerance": "2e-3 relative on pooled activations / logits (stated, tested)"}}}
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
            xtools.encode_sequences(seqs[i:i + IB], L, dev)     # host join + H2D of the bytes + encode kernel
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
        v, sec, threads = cpu_port_step_throughput(B, steps=3, warmup=1)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                               "sample": "1 warm-up + 3 timed steps of the full 512-sequence cfg2 batch, torch CPU "
                                         "(oracle/torch_port.py restates the reference module graph)",
                               "ms_per_step": sec * 1e3}
    except Exception as ex:       # pragma: no cover
        out["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": "failed: %s" % ex}
    return out


if __name__ == "__main__":
    main()
