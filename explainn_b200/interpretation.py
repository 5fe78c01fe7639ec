"""interpretation -- drop-in for explainn/interpretation/interpretation.py of the reference (the hot-path
helpers: predictions, unit outputs, unit activations, PWM extraction, MEME export), plus the fused
two-pass PWM extraction that never materialises the (N,U,S) activation tensor.
"""
import numpy as np
import torch
import torch.nn as nn

from . import functional as Fn


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


class PwmAccumulator:
    """Two-pass PWM extraction over a stream of batches without the activation tensor.

    pass A (`scan_max`): per-unit maximum pre-activation -> float16 threshold 0.5*max (interpretation.py:187)
    pass B (`scan_count`): count the one-hot windows at every position above threshold (:203-214)
    In a multi-GPU run each rank scans its shard; `zmax` is all-reduced with MAX between the passes and
    `counts`/`hits` with SUM after pass B (see parallel.py).
    """

    def __init__(self, model):
        if model.linears[1].training:
            raise RuntimeError("PWM extraction uses BatchNorm running statistics; call model.eval() first")
        self.model = model
        self.prep = model._conv_prep_eval()
        self.zmax = None
        self.thr = self.zlo = self.zhi = self.amax = None
        self.counts = self.hits = None
        self.keep_seq_hits = False      # set True to also collect, per batch, which sequences pass each unit's threshold
        self.seq_hits = []

    def scan_max(self, x):
        pk = self.model._pack(x)
        self.zmax = Fn.interp_unit_max(pk, self.prep, self.zmax)

    def finish_max(self):
        if self.zmax is None:
            raise RuntimeError("scan_max was never called")
        self.thr, self.zlo, self.zhi, self.amax = Fn.interp_thresholds(self.zmax)

    def scan_count(self, x):
        pk = self.model._pack(x)
        seq_hit = None
        if self.keep_seq_hits:
            seq_hit = torch.zeros((pk.B, self.prep.U), dtype=torch.uint8, device=pk.codes.device)
            self.seq_hits.append(seq_hit)
        self.counts, self.hits = Fn.interp_pwm_count(pk, self.prep, self.thr, self.zlo, self.zhi, self.counts, self.hits,
                                                     seq_hit)

    def result(self):
        self.model.check_inputs()
        counts = self.counts.cpu().numpy()
        hits = self.hits.cpu().numpy()
        return pwm_from_counts(counts, hits), counts, hits


def get_pwms_explainn_fused(data_loader, model, device, return_counts=False):
    """PWMs straight from the sequences: same result as
    get_pwms_explainn(get_explainn_unit_activations(loader, model, device), sequences, filter_size)
    but the (N,U,S) float16 tensor never exists.  `data_loader` yields (seq, label) batches (or is a
    single (N,4,L) tensor) and is iterated twice."""
    acc = PwmAccumulator(model)

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
