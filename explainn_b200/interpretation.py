"""interpretation -- drop-in for explainn/interpretation/interpretation.py of the reference (the hot-path
helpers: predictions, unit outputs, unit activations, PWM extraction, MEME export), plus the fused
two-pass PWM extraction that never materialises the (N,U,S) activation tensor.
"""
import numpy as np
import torch
import torch.nn as nn

from . import functional as Fn
from ._cabi import on_device_of as _on_device_of


def get_explainn_predictions(data_loader, model, device, isSigmoid=False):
    """Predictions and true labels for all sequences of the loader (interpretation.py:51-79).
    :return: tuple of numpy.arrays (predictions, labels) -- note the order, opposite to run_test."""
    outs, labels = [], []
    with torch.no_grad():
        for seq, lbl in data_loader:
            outs.append(model(seq.to(device, non_blocking=True)))
            labels.append(lbl)
    if hasattr(model, "check_inputs"):
        model.check_inputs()
    outputs = torch.cat(outs, 0).cpu()
    if isSigmoid:
        outputs = nn.Sigmoid()(outputs)
    return outputs.numpy(), torch.cat([l.cpu() for l in labels], 0).numpy()


def get_explainn_unit_outputs(data_loader, model, device):
    """Outputs of the individual units, shape (N, U) (interpretation.py:82-99)."""
    outs = []
    with torch.no_grad():
        for seq, lbl in data_loader:
            outs.append(model.linears(seq.to(device, non_blocking=True)))
    if hasattr(model, "check_inputs"):
        model.check_inputs()
    return torch.cat(outs, 0).cpu().numpy()


def get_explainn_unit_activations(data_loader, model, device):
    """exp(BN(conv)) activation map of every unit as float16, shape (N, U, S) (interpretation.py:126-149).
    The float16 rounding happens on the GPU, so half the bytes cross PCIe."""
    outs = []
    with torch.no_grad():
        for seq, lbl in data_loader:
            outs.append(model._activations(seq.to(device, non_blocking=True), torch.float16).cpu())
    model.check_inputs()
    return torch.cat(outs, 0).numpy()


def _as_onehot_tensor(sequences, device):
    if isinstance(sequences, torch.Tensor):
        return sequences.to(device)
    return torch.as_tensor(np.asarray(sequences)).to(device)


def pwm_from_counts(counts, hits):
    """Column-normalise integer counts the way interpretation.py:217-224 does for float32 one-hot input:
    float32 division widened to float64; units without any window above threshold keep 0.25."""
    counts = np.asarray(counts)
    U, four, F = counts.shape
    pwm = np.full((U, 4, F), 0.25)
    for u in range(U):
        if hits[u] == 0:
            continue
        c = counts[u].astype(np.float32)
        total = c.sum(axis=0)
        total[total == 0] = 1.0
        pwm[u] = np.nan_to_num(c / total)
    return pwm


def _chunks_on_device(array, device, budget_bytes=1 << 30):
    """Yield (lo, hi, device tensor) chunks of a host (or device) array along axis 0, each at most budget_bytes."""
    n = array.shape[0]
    if n == 0:
        return
    per = max(1, int(np.prod(array.shape[1:])) * array.dtype.itemsize if isinstance(array, np.ndarray)
              else array[0].numel() * array.element_size())
    step = max(1, budget_bytes // per)
    for lo in range(0, n, step):
        hi = min(n, lo + step)
        part = array[lo:hi]
        if isinstance(part, np.ndarray):
            part = torch.from_numpy(np.ascontiguousarray(part))
        yield lo, hi, part.to(device, non_blocking=True).contiguous()


def get_pwms_explainn(activations, sequences, filter_size, device=None, return_counts=False):
    """Convert unit activations to PWMs -- drop-in for explainn/interpretation/interpretation.py:176-227.

    activations: (N,U,S) array as returned by get_explainn_unit_activations (float16; float32 is accepted);
    sequences: (N,4,L) one-hot (torch tensor or numpy array); returns float64 (U,4,filter_size).
    The threshold 0.5 * max (in the activations' own dtype) and the window counts are computed on the GPU from the
    caller's array -- nothing is recomputed from the model, so the counts are the reference's integers for that
    array.  PWM values are float32 counts/total widened to float64 (the reference's result for float32 one-hot
    input; its float16-input path saturates sums at 2048 and is deliberately not reproduced).  Prefer
    get_pwms_explainn_fused when the activation array is not needed for anything else."""
    if isinstance(activations, torch.Tensor):
        act = activations
        if act.dtype not in (torch.float16, torch.float32):
            act = act.float()
    else:
        act = np.asarray(activations)
        if act.dtype not in (np.float16, np.float32):
            act = act.astype(np.float32)
    if act.ndim != 3:
        raise ValueError("activations must have shape (N, U, S), got %s" % (tuple(act.shape),))
    N, U, S = act.shape
    F = int(filter_size)
    if sequences.shape[0] != N or sequences.shape[1] != 4 or sequences.shape[2] != S + F - 1:
        raise ValueError("sequences of shape %s do not match activations %s and filter_size %d"
                         % (tuple(sequences.shape), (N, U, S), F))
    if device is None:
        if isinstance(sequences, torch.Tensor) and sequences.is_cuda:
            device = sequences.device
        elif isinstance(act, torch.Tensor) and act.is_cuda:
            device = act.device
        else:
            device = torch.device("cuda", torch.cuda.current_device())
    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("explainn_b200 runs only on CUDA devices (there is no CPU fallback)")
    tdtype = torch.float16 if (act.dtype in (np.float16, torch.float16)) else torch.float32
    invalid = torch.zeros(1, dtype=torch.int32, device=device)
    with torch.cuda.device(device):
        parts = list(_chunks_on_device(act, device))
        keep = len(parts) == 1                      # a single chunk stays on the device for the second pass
        state = None
        for lo, hi, a in parts:
            state = Fn.act_unit_max(a, state)
        if state is None:
            state = torch.zeros(2 * U, dtype=torch.int32, device=device)
        amax, thr = Fn.act_thresholds(state, U, tdtype)
        counts = torch.zeros((U, 4, F), dtype=torch.int64, device=device)
        hits = torch.zeros(U, dtype=torch.int64, device=device)
        for lo, hi, a in (parts if keep else _chunks_on_device(act, device)):
            x = sequences[lo:hi]
            if not isinstance(x, torch.Tensor):
                x = torch.from_numpy(np.ascontiguousarray(x))
            x = x.to(device, non_blocking=True)
            pk = Fn.pack_onehot(x, 1, invalid)
            Fn.act_pwm_count(a, pk, F, thr, counts, hits)
        Fn.check_invalid(invalid)
        counts_h, hits_h = counts.cpu().numpy(), hits.cpu().numpy()
    pwm = pwm_from_counts(counts_h, hits_h)
    if return_counts:
        return pwm, counts_h, hits_h
    return pwm


def get_explainn_unit_argmax(data_loader, model, device):
    """Per (sequence, unit): the maximum activation and the first position that attains it, computed inside the
    scan (no (N,U,S) tensor) -- `torch.max(activations, dim=2)` (networks.py:228,281) over `model.linears[:3]`.
    Returns (max (N,U) float32, argmax (N,U) int32).  `data_loader` yields (seq, label) batches or is an (N,4,L)
    tensor."""
    if model.linears[1].training:
        raise RuntimeError("the activation scan uses BatchNorm running statistics; call model.eval() first")
    prep = model._conv_prep_eval()
    vals, poss = [], []

    def batches():
        if isinstance(data_loader, torch.Tensor):
            for i in range(0, data_loader.shape[0], 8192):
                yield data_loader[i:i + 8192]
        else:
            for item in data_loader:
                yield item[0] if isinstance(item, (tuple, list)) else item

    with torch.no_grad():
        for seq in batches():
            a, p = Fn.interp_seq_argmax(model._pack(seq.to(device, non_blocking=True)), prep)
            vals.append(a)
            poss.append(p)
    model.check_inputs()
    return torch.cat(vals, 0).cpu().numpy(), torch.cat(poss, 0).cpu().numpy()


def unit_argmax_global(amax, apos):
    """From the per-sequence results of get_explainn_unit_argmax: for every unit the maximum activation over the whole
    dataset, the first sequence that attains it and the position inside it -> (max (U,), seq (U,), pos (U,))."""
    n_star = np.argmax(amax, axis=0)
    u = np.arange(amax.shape[1])
    return amax[n_star, u], n_star, apos[n_star, u]


class PwmAccumulator:
    """Two-pass PWM extraction over a stream of batches without the activation tensor.

    pass A (`scan_max`): per-unit maximum pre-activation -> float16 threshold 0.5*max (interpretation.py:187)
    pass B (`scan_count`): count the one-hot windows at every position above threshold (:203-214)
    In a multi-GPU run each rank scans its shard; `zmax` is all-reduced with MAX between the passes and
    `counts`/`hits` with SUM after pass B (see parallel.py).
    """

    def __init__(self, model, backend="tc"):
        """backend "tc": both passes on the tensor cores (positions near a decision re-evaluated in the canonical order:
        same integers as "lut"); "lut": the k-mer table kernel throughout."""
        if model.linears[1].training:
            raise RuntimeError("PWM extraction uses BatchNorm running statistics; call model.eval() first")
        if backend not in ("tc", "lut"):
            raise ValueError("backend must be 'tc' or 'lut'")
        self.model = model
        self.backend = backend
        conv, bn = model.linears[0], model.linears[1]
        with torch.cuda.device(conv.weight.device):
            self.prep = model._conv_prep_eval()
            self.tc = None
            if backend == "tc":
                tcp = Fn.conv_tc_prepare(conv.weight.detach(), conv.bias.detach(), bn.weight.detach(), bn.bias.detach(),
                                         bn.running_mean, bn.running_var, bn.eps)
                self.tc = Fn.InterpTc(conv.weight.detach(), self.prep, tcp)
        self.zmax = None
        self.thr = self.zlo = self.zhi = self.amax = None
        self.counts = self.hits = None
        self.keep_seq_hits = False      # set True to also collect, per batch, which sequences pass each unit's threshold
        self.seq_hits = []

    @_on_device_of(1)
    def scan_max(self, x):
        pk = self.model._pack(x)
        if self.tc is not None:
            if self.zmax is None:
                self.zmax = torch.full((self.prep.U,), float("-inf"), dtype=torch.float32, device=pk.codes.device)
            self.tc.unit_max(pk, self.zmax)
        else:
            self.zmax = Fn.interp_unit_max(pk, self.prep, self.zmax)

    def finish_max(self):
        if self.zmax is None:
            raise RuntimeError("scan_max was never called")
        with torch.cuda.device(self.zmax.device):
            self.thr, self.zlo, self.zhi, self.amax = Fn.interp_thresholds(self.zmax)

    @_on_device_of(1)
    def scan_count(self, x):
        pk = self.model._pack(x)
        seq_hit = None
        if self.keep_seq_hits:
            seq_hit = torch.zeros((pk.B, self.prep.U), dtype=torch.uint8, device=pk.codes.device)
            self.seq_hits.append(seq_hit)
        if self.tc is not None:
            if self.counts is None:
                self.counts = torch.zeros((self.prep.U, 4, self.prep.F), dtype=torch.int64, device=pk.codes.device)
                self.hits = torch.zeros(self.prep.U, dtype=torch.int64, device=pk.codes.device)
            self.tc.pwm_count(pk, self.thr, self.zlo, self.zhi, self.counts, self.hits, seq_hit)
        else:
            self.counts, self.hits = Fn.interp_pwm_count(pk, self.prep, self.thr, self.zlo, self.zhi, self.counts, self.hits,
                                                         seq_hit)

    def result(self):
        self.model.check_inputs()
        if self.tc is not None:
            self.tc.check()
        counts = self.counts.cpu().numpy()
        hits = self.hits.cpu().numpy()
        return pwm_from_counts(counts, hits), counts, hits


def get_pwms_explainn_fused(data_loader, model, device, return_counts=False, backend="tc"):
    """PWMs straight from the sequences: same result as
    get_pwms_explainn(get_explainn_unit_activations(loader, model, device), sequences, filter_size)
    but the (N,U,S) float16 tensor never exists.  `data_loader` yields (seq, label) batches (or is a
    single (N,4,L) tensor) and is iterated twice.  `backend`: see PwmAccumulator."""
    acc = PwmAccumulator(model, backend=backend)

    def batches():
        if isinstance(data_loader, torch.Tensor):
            for i in range(0, data_loader.shape[0], 8192):
                yield data_loader[i:i + 8192]
        else:
            for item in data_loader:
                yield item[0] if isinstance(item, (tuple, list)) else item

    with torch.no_grad():
        for seq in batches():
            acc.scan_max(seq.to(device, non_blocking=True))
        acc.finish_max()
        for seq in batches():
            acc.scan_count(seq.to(device, non_blocking=True))
    pwm, counts, hits = acc.result()
    if return_counts:
        return pwm, counts, hits
    return pwm


def pwm_to_meme(pwm, output_file_path, names=None, verbose=True):
    """Write PWMs as a MEME v4 file (interpretation.py:230-258): all-zero columns are skipped and `w=`
    counts the non-zero columns."""
    n_filters = len(pwm)
    if names is None:
        names = {i: f"filter{i}" for i in range(n_filters)}
    with open(output_file_path, "w") as meme_file:
        meme_file.write("MEME version 4\n")
        for i in range(n_filters):
            if np.sum(pwm[i][:, :]) > 0:
                meme_file.write("\n")
                meme_file.write("MOTIF %s \n" % names[i])
                meme_file.write("letter-probability matrix: alength= 4 w= %d \n"
                                % np.count_nonzero(np.sum(pwm[i][:, :], axis=0)))
            for j in range(pwm[i].shape[1]):
                if np.sum(pwm[i][:, j]) > 0:
                    meme_file.write("\t".join(str(pwm[i][c, j]) for c in range(4)) + "\n")
    if verbose:
        print('Saved PWM File as : {}'.format(output_file_path))


# ------------------------------------------------------------------------------------------ unit importance
def get_specific_unit_importance(activations, model, unit_outputs, filt, target_labels):
    """Importance (unit output x class weight) of unit `filt` over the sequences that activate it above its
    threshold -- drop-in for interpretation.py:303-340 (host arrays in, pandas.Series out)."""
    import pandas as pd
    thr = 0.5 * np.amax(activations, axis=(0, 2))
    hit = (activations[:, filt, :] > thr[filt]).any(axis=1)
    weights = model.final.weight.detach().cpu().numpy()
    res = {}
    for cl in range(len(target_labels)):
        res[target_labels[cl]] = np.multiply(unit_outputs, weights[cl])[:, filt][hit]
    return pd.Series(res)


def _hit_sequences(activations, chunk=1024):
    """(N,U) bool: does sequence n have any position above 0.5 * max for unit u (interpretation.py:269-279)."""
    activations = np.asarray(activations)
    thr = 0.5 * np.amax(activations, axis=(0, 2))
    out = np.empty(activations.shape[:2], bool)
    for i in range(0, activations.shape[0], chunk):
        out[i:i + chunk] = (activations[i:i + chunk] > thr[None, :, None]).any(axis=2)
    return out


def get_median_unit_importance(activations, model, unit_outputs, target_labels):
    """Median unit importance (unit output x class weight over the sequences that activate the unit) -- drop-in for
    interpretation.py:261-300.  The reference hard-codes 100 units and 50 classes; this returns
    (U, len(target_labels)), which is the same (100, 50) matrix for that case.  Units no sequence activates give NaN
    (numpy's median of an empty array, as in the reference)."""
    hit = _hit_sequences(activations)
    weights = model.final.weight.detach().cpu().numpy()
    unit_outputs = np.asarray(unit_outputs)
    U, O = hit.shape[1], len(target_labels)
    out = np.full((U, O), np.nan)
    for u in range(U):
        rows = hit[:, u]
        if rows.any():
            out[u] = np.median(unit_outputs[rows, u][:, None] * weights[None, :O, u], axis=0)
    return out


def unit_importance_fused(data_loader, model, device, target_labels=None):
    """Unit importances for ALL units without the activation tensor.

    The per-unit sequence sets ("has a position above 0.5 x max", interpretation.py:311-317) come out of the PWM
    count pass as a by-product; unit outputs come from the fused forward.  Returns
      importance: dict unit -> pandas.Series(label -> 1-D array over the sequences that activate the unit)
                  (the object get_specific_unit_importance returns, for every unit)
      median:     (U, C) array of medians (get_median_unit_importance, interpretation.py:261-300, without its
                  hard-coded 100 units / 50 classes); NaN where no sequence activates the unit
    """
    import pandas as pd

    def batches():
        if isinstance(data_loader, torch.Tensor):
            for i in range(0, data_loader.shape[0], 8192):
                yield data_loader[i:i + 8192]
        else:
            for item in data_loader:
                yield item[0] if isinstance(item, (tuple, list)) else item

    acc = PwmAccumulator(model)
    acc.keep_seq_hits = True
    units = []
    with torch.no_grad():
        for seq in batches():
            acc.scan_max(seq.to(device, non_blocking=True))
        acc.finish_max()
        for seq in batches():
            x = seq.to(device, non_blocking=True)
            acc.scan_count(x)
            units.append(model.linears(x))
    model.check_inputs()
    hit = torch.cat(acc.seq_hits, 0).bool().cpu().numpy()          # (N,U)
    unit_outputs = torch.cat(units, 0).cpu().numpy()               # (N,U)
    weights = model.final.weight.detach().cpu().numpy()            # (C,U)
    C, U = weights.shape
    if target_labels is None:
        target_labels = list(range(C))
    importance, median = {}, np.full((U, C), np.nan)
    for u in range(U):
        rows = hit[:, u]
        vals = unit_outputs[rows, u][:, None] * weights[None, :, u]      # (n_hit, C)
        importance[u] = pd.Series({target_labels[c]: vals[:, c] for c in range(C)})
        if vals.shape[0]:
            median[u] = np.median(vals, axis=0)
    return importance, median
