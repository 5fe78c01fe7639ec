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


class _DevBuffer:
    """A cudaMalloc allocation of the C library exposed through __cuda_array_interface__ (torch.as_tensor wraps it
    without copying); freed when the last tensor view and this object are gone."""

    def __init__(self, nbytes, device):
        import ctypes
        from . import _cabi
        self._cabi, self.device = _cabi, device
        p = ctypes.c_void_p()
        with torch.cuda.device(device):
            _cabi.call("explainn_p2p_alloc", int(nbytes), ctypes.byref(p))
        self.ptr, self.nbytes = p.value, int(nbytes)
        self.__cuda_array_interface__ = {"shape": (self.nbytes,), "typestr": "|u1", "data": (self.ptr, False), "version": 2}

    def __del__(self):
        try:
            with torch.cuda.device(self.device):
                self._cabi.call("explainn_p2p_free", self.ptr)
        except Exception:
            pass


class PeerGradients:
    """Peer-memory plumbing of the fused all-reduce + Adam kernel (csrc/p2p.cu).  Every rank allocates its flat gradient
    buffer and a small flag buffer as plain cudaMalloc memory (CUDA IPC exports whole allocations), exchanges the IPC
    handles over the process group and maps the peers' buffers for its own GPU, so that `explainn_p2p_adam_step` reads
    all gradients over NVLink.  `self.flat_grad` is the torch view the backward kernels write.  One node only."""

    def __init__(self, n_floats, device):
        import ctypes
        from . import _cabi
        d = _dist()
        if d is None or d.get_world_size() < 2:
            raise RuntimeError("PeerGradients needs an initialised process group with at least two ranks")
        self.rank, self.world = d.get_rank(), d.get_world_size()
        if self.world > 8:
            raise RuntimeError("PeerGradients: at most 8 ranks (one NVSwitch box)")
        self._cabi, self.device = _cabi, torch.device(device)
        self._opened = []
        self.ok, self.error = True, ""
        # every rank takes part in both collectives below whatever happens locally, so a rank that cannot export or map
        # a buffer (no peer access, IPC not permitted) makes ALL ranks fall back together (self.ok False everywhere)
        payload = None
        try:
            self._gbuf = _DevBuffer(4 * n_floats, self.device)
            self._fbuf = _DevBuffer(_cabi.lib.explainn_p2p_flag_bytes(), self.device)
            self.flat_grad = torch.as_tensor(self._gbuf, device=self.device).view(torch.float32)
            self.flags = torch.as_tensor(self._fbuf, device=self.device).view(torch.int64)
            hg, hf = (ctypes.c_ubyte * 64)(), (ctypes.c_ubyte * 64)()
            with torch.cuda.device(self.device):
                _cabi.call("explainn_p2p_export", self._gbuf.ptr, ctypes.cast(hg, ctypes.c_void_p))
                _cabi.call("explainn_p2p_export", self._fbuf.ptr, ctypes.cast(hf, ctypes.c_void_p))
            payload = (n_floats, bytes(hg), bytes(hf))
        except Exception as ex:
            self.ok, self.error = False, "export: %s" % ex
        everyone = [None] * self.world
        d.all_gather_object(everyone, payload)
        if any(e is None for e in everyone):
            self.ok = False
            self.error = self.error or "a peer could not export its buffers"
        g_ptrs, f_ptrs = [], []
        if self.ok:
            try:
                for q, (qn, qg, qf) in enumerate(everyone):
                    if qn != n_floats:
                        raise RuntimeError("rank %d holds %d gradient elements, this rank %d" % (q, qn, n_floats))
                    if q == self.rank:
                        g_ptrs.append(self._gbuf.ptr)
                        f_ptrs.append(self._fbuf.ptr)
                        continue
                    pg, pf = ctypes.c_void_p(), ctypes.c_void_p()
                    with torch.cuda.device(self.device):
                        _cabi.call("explainn_p2p_open", ctypes.cast((ctypes.c_ubyte * 64).from_buffer_copy(qg), ctypes.c_void_p), ctypes.byref(pg))
                        self._opened.append(pg.value)
                        _cabi.call("explainn_p2p_open", ctypes.cast((ctypes.c_ubyte * 64).from_buffer_copy(qf), ctypes.c_void_p), ctypes.byref(pf))
                        self._opened.append(pf.value)
                    g_ptrs.append(pg.value)
                    f_ptrs.append(pf.value)
                self.g_array = (ctypes.c_void_p * 8)(*g_ptrs)
                self.f_array = (ctypes.c_void_p * 8)(*f_ptrs)
                torch.cuda.synchronize(self.device)
            except Exception as ex:
                self.ok, self.error = False, "open: %s" % ex
        agree = torch.tensor([1 if self.ok else 0], dtype=torch.int32, device=self.device)
        d.all_reduce(agree, op=d.ReduceOp.MIN)
        if int(agree.item()) == 0:
            self.ok = False
            self.error = self.error or "a peer could not map the buffers"

    def __del__(self):
        try:
            with torch.cuda.device(self.device):
                for p in self._opened:
                    self._cabi.call("explainn_p2p_close", p)
        except Exception:
            pass
