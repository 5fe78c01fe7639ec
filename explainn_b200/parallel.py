"""Data-parallel helpers: one process per GPU (torchrun), torch.distributed for the plumbing.

The hot path shards by sequence.  Collectives exist only where the path has a real exchange step:
  training        one all-reduce (SUM) of the flat gradient buffer per step  (train.FusedTrainer.step)
  interpretation  all-reduce (MAX) of the per-unit maxima between the two passes,
                  all-reduce (SUM) of the PWM counts after the second pass
  inference       none (each rank keeps / writes its own shard)
Everything here works on CPU tensors with the gloo backend too, which is how the host logic is tested
(tests/test_parallel_gloo.py).
"""
import torch


def _dist():
    import torch.distributed as dist
    return dist if dist.is_available() and dist.is_initialized() else None


def world():
    d = _dist()
    return (d.get_rank(), d.get_world_size()) if d is not None else (0, 1)


def shard_range(n, rank=None, world_size=None):
    """Contiguous shard [lo, hi) of n items for this rank; sizes differ by at most one and the shards of
    ranks 0..W-1 concatenate to range(n) in order (so outputs gathered rank by rank keep the input order)."""
    if rank is None or world_size is None:
        rank, world_size = world()
    base, rem = divmod(n, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_unit_max(zmax):
    """MAX over ranks of the per-unit maximum pre-activation (pass A of the PWM extraction)."""
    d = _dist()
    if d is not None and d.get_world_size() > 1:
        d.all_reduce(zmax, op=d.ReduceOp.MAX)
    return zmax


def allreduce_counts(counts, hits):
    """SUM over ranks of the integer PWM counts and hit counters (pass B)."""
    d = _dist()
    if d is not None and d.get_world_size() > 1:
        d.all_reduce(counts, op=d.ReduceOp.SUM)
        d.all_reduce(hits, op=d.ReduceOp.SUM)
    return counts, hits


def allreduce_mean_gradients(flat_grad):
    """The one collective of a training step: SUM the flat gradient buffer; returns the factor by which the
    optimizer must scale it (1/world) so that the update uses the mean over ranks (DDP semantics)."""
    d = _dist()
    if d is not None and d.get_world_size() > 1:
        d.all_reduce(flat_grad, op=d.ReduceOp.SUM)
        return 1.0 / d.get_world_size()
    return 1.0


def gather_rows(local_rows, n_total):
    """All-gather variable-length row shards (made with shard_range) back into input order."""
    d = _dist()
    if d is None or d.get_world_size() == 1:
        return local_rows
    W = d.get_world_size()
    sizes = [shard_range(n_total, r, W) for r in range(W)]
    maxn = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((maxn,) + tuple(local_rows.shape[1:]), dtype=local_rows.dtype, device=local_rows.device)
    pad[:local_rows.shape[0]] = local_rows
    out = [torch.empty_like(pad) for _ in range(W)]
    d.all_gather(out, pad)
    return torch.cat([o[:hi - lo] for o, (lo, hi) in zip(out, sizes)], 0)


def sharded_pwms(sequences, model, device, batch=8192):
    """Fused PWM extraction over this rank's shard of `sequences` ((N,4,L) tensor, same on every rank) with the
    two reductions above; every rank returns the full-dataset PWMs."""
    from . import interpretation
    lo, hi = shard_range(sequences.shape[0])
    acc = interpretation.PwmAccumulator(model)
    with torch.no_grad():
        for i in range(lo, hi, batch):
            acc.scan_max(sequences[i:min(hi, i + batch)].to(device, non_blocking=True))
        if acc.zmax is None:
            acc.zmax = torch.full((model._options["num_cnns"],), float("-inf"), device=device)
        allreduce_unit_max(acc.zmax)
        acc.finish_max()
        for i in range(lo, hi, batch):
            acc.scan_count(sequences[i:min(hi, i + batch)].to(device, non_blocking=True))
        if acc.counts is None:
            U, F = model._options["num_cnns"], model._options["filter_size"]
            acc.counts = torch.zeros((U, 4, F), dtype=torch.int64, device=device)
            acc.hits = torch.zeros(U, dtype=torch.int64, device=device)
        allreduce_counts(acc.counts, acc.hits)
    return acc.result()
