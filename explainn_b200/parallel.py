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
    """The one collective of a training step on the NCCL path (train.FusedTrainer._finish_allreduce): SUM the flat
    gradient buffer; returns the factor by which the optimizer must scale it (1/world) so that the update uses the
    mean over ranks (DDP semantics)."""
    d = _dist()
    if d is not None and d.get_world_size() > 1:
        d.all_reduce(flat_grad, op=d.ReduceOp.SUM)
        return 1.0 / d.get_world_size()
    return 1.0


def broadcast_model_state(model, src=0):
    """Make every replica start from rank `src`'s parameters and buffers (what DistributedDataParallel does at
    construction): without it, ranks that built their models from different RNG states would stay different."""
    d = _dist()
    if d is None or d.get_world_size() == 1:
        return
    for t in list(model.parameters()) + list(model.buffers()):
        d.broadcast(t.data, src=src)


def allreduce_mean_scalar(value, device):
    """Mean over ranks of a python float (epoch validation loss: every rank must take the same best-epoch and
    early-stopping decisions, or the replicas and the collectives of the next epoch fall apart)."""
    d = _dist()
    if d is None or d.get_world_size() == 1:
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device if d.get_backend() == "nccl" else "cpu")
    d.all_reduce(t, op=d.ReduceOp.SUM)
    return float(t.item()) / d.get_world_size()


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
    """Peer-memory plumbing of the fused reduce-scatter + Adam + all-gather kernel (csrc/p2p.cu).  Every rank allocates
    its flat gradient buffer, its flat parameter buffer and a small flag buffer as plain cudaMalloc memory (CUDA IPC
    exports whole allocations), exchanges the IPC handles over the process group and maps the peers' buffers for its own
    GPU, so that `explainn_p2p_adam_step` reads all gradients and writes all parameter replicas over NVLink.
    `self.flat_grad` / `self.flat_param` are the torch views the kernels and the nn.Parameters use.  One node only."""

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
            self._pbuf = _DevBuffer(4 * n_floats, self.device)
            self._fbuf = _DevBuffer(_cabi.lib.explainn_p2p_flag_bytes(), self.device)
            self.flat_grad = torch.as_tensor(self._gbuf, device=self.device).view(torch.float32)
            self.flat_param = torch.as_tensor(self._pbuf, device=self.device).view(torch.float32)
            self.flags = torch.as_tensor(self._fbuf, device=self.device).view(torch.int64)
            handles = []
            with torch.cuda.device(self.device):
                for buf in (self._gbuf, self._pbuf, self._fbuf):
                    h = (ctypes.c_ubyte * 64)()
                    _cabi.call("explainn_p2p_export", buf.ptr, ctypes.cast(h, ctypes.c_void_p))
                    handles.append(bytes(h))
            payload = (n_floats, handles)
        except Exception as ex:
            self.ok, self.error = False, "export: %s" % ex
        everyone = [None] * self.world
        d.all_gather_object(everyone, payload)
        if any(e is None for e in everyone):
            self.ok = False
            self.error = self.error or "a peer could not export its buffers"
        ptrs = [[], [], []]
        if self.ok:
            try:
                for q, (qn, qh) in enumerate(everyone):
                    if qn != n_floats:
                        raise RuntimeError("rank %d holds %d gradient elements, this rank %d" % (q, qn, n_floats))
                    if q == self.rank:
                        for lst, buf in zip(ptrs, (self._gbuf, self._pbuf, self._fbuf)):
                            lst.append(buf.ptr)
                        continue
                    with torch.cuda.device(self.device):
                        for lst, hb in zip(ptrs, qh):
                            pp = ctypes.c_void_p()
                            _cabi.call("explainn_p2p_open", ctypes.cast((ctypes.c_ubyte * 64).from_buffer_copy(hb), ctypes.c_void_p),
                                       ctypes.byref(pp))
                            self._opened.append(pp.value)
                            lst.append(pp.value)
                self.g_array = (ctypes.c_void_p * 8)(*ptrs[0])
                self.p_array = (ctypes.c_void_p * 8)(*ptrs[1])
                self.f_array = (ctypes.c_void_p * 8)(*ptrs[2])
                torch.cuda.synchronize(self.device)
            except Exception as ex:
                self.ok, self.error = False, "open: %s" % ex
        agree = torch.tensor([1 if self.ok else 0], dtype=torch.int32, device=self.device)
        d.all_reduce(agree, op=d.ReduceOp.MIN)
        if int(agree.item()) == 0:
            self.ok = False
            self.error = self.error or "a peer could not map the buffers"

    def slice_range(self, n_floats, offset=0):
        """[lo, hi) in floats of the slice this rank owns inside the range [offset, offset + n_floats) of the flat
        buffers (same split as csrc/p2p.cu)."""
        n4 = n_floats // 4
        return offset + 4 * (n4 * self.rank // self.world), offset + 4 * (n4 * (self.rank + 1) // self.world)

    def check(self):
        """Raise if a step of the peer-memory optimizer kernel gave up waiting for a rank (synchronises)."""
        import ctypes
        w = ctypes.c_int(0)
        with torch.cuda.device(self.device):
            self._cabi.call("explainn_p2p_error", self._fbuf.ptr, ctypes.byref(w))
        if w.value:
            raise RuntimeError("explainn_b200: the peer-memory optimizer step gave up waiting for rank %d "
                               "(XB_P2P_TIMEOUT_S); the replicas are no longer consistent" % (w.value - 1))

    def __del__(self):
        try:
            with torch.cuda.device(self.device):
                for p in self._opened:
                    self._cabi.call("explainn_p2p_close", p)
        except Exception:
            pass
