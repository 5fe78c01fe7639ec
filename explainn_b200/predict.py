"""Strand-aware batched prediction -- the inner loop of scripts/fasta2predictions.py:148-201 and
scripts/tsv2predictions.py:144-197 of the reference (`_predict`): every sequence is scored on both strands
and [Fwd, Rev, Mean, Max] is reported per class.

The reverse strand is `s[::-1, ::-1]` on the one-hot matrix (fasta2predictions.py:160).  Here each batch is
packed once and reverse-complemented on the 2-bit form (explainn_revcomp_packed), so both strands share one
host-to-device copy and one packing pass.
"""
import numpy as np
import torch

from . import functional as Fn


def predict_both_strands(seqs, model, device, apply_sigmoid=False, batch_size=8192):
    """
    :param seqs: (N,4,L) one-hot array / tensor, or an iterable of such batches
    :param model: ExplaiNN model (set to eval() here, as _predict does at fasta2predictions.py:155)
    :param device: "cuda:0"
    :param apply_sigmoid: apply the logistic function to all four columns (fasta2predictions.py:186-187)
    :return: numpy.array (N, num_classes, 4), float64, columns Fwd, Rev, Mean, Max
    """
    model.to(device)
    model.eval()
    if isinstance(seqs, np.ndarray):
        seqs = torch.from_numpy(np.ascontiguousarray(seqs))
    batches = ([seqs[i:i + batch_size] for i in range(0, seqs.shape[0], batch_size)]
               if isinstance(seqs, torch.Tensor) else seqs)
    fwd_all, rev_all = [], []
    with torch.no_grad():
        for xb in batches:
            xb = torch.as_tensor(xb).to(device, non_blocking=True)
            pk = model._pack(xb)
            fwd_all.append(model.forward_packed(pk))
            rev_all.append(model.forward_packed(Fn.revcomp_packed(pk)))
    model.check_inputs()
    fwd = torch.cat(fwd_all, 0).cpu().numpy()
    rev = torch.cat(rev_all, 0).cpu().numpy()
    both = np.stack((fwd, rev), axis=2)                       # float32, as fasta2predictions.py:166-172
    preds = np.empty(fwd.shape + (4,))                        # float64, fasta2predictions.py:153
    preds[:, :, 0] = fwd
    preds[:, :, 1] = rev
    preds[:, :, 2] = np.mean(both, axis=2)
    preds[:, :, 3] = np.max(both, axis=2)
    if apply_sigmoid:
        preds = torch.sigmoid(torch.Tensor(preds)).numpy()
    return preds


def predictions_to_frame(predictions, ids):
    """Long-format table of fasta2predictions.py:190-200: SeqId, Class, Fwd, Rev, Mean, Max."""
    import pandas as pd
    dfs = []
    for i in range(predictions.shape[1]):
        df = pd.DataFrame(predictions[:, i, :], columns=["Fwd", "Rev", "Mean", "Max"])
        df["SeqId"] = ids
        df["Class"] = i
        dfs.append(df)
    df = pd.concat(dfs)[["SeqId", "Class", "Fwd", "Rev", "Mean", "Max"]]
    df.reset_index(drop=True, inplace=True)
    return df
