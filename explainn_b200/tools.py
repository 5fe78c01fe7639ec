"""Input encoding -- the on-device counterpart of explainn/utils/tools.py:136-181 (`dna_one_hot`).

`encode_sequences` is the product path: a batch of DNA strings goes to the GPU as raw bytes (1 byte per base)
and is turned into the packed 2-bit + N-mask form every kernel of this package reads.  `dna_one_hot` keeps the
reference's single-sequence signature and return type for callers that want the (4,L) float16 matrix back.
"""
import torch

from . import functional as Fn


def encode_sequences(seqs, seq_len, device="cuda"):
    """list of str/bytes -> functional.Packed on `device` (same trimming / padding / alphabet rules as `dna_one_hot`)."""
    return Fn.encode_dna(seqs, seq_len, device)


def one_hot_batch(seqs, seq_len, device="cuda", dtype=torch.float32):
    """list of str -> (B,4,L) one-hot tensor on the GPU."""
    return Fn.unpack_onehot(Fn.encode_dna(seqs, seq_len, device), dtype)


def dna_one_hot(seq, seq_len=None, flatten=False, device="cuda"):
    """Drop-in for tools.dna_one_hot (tools.py:136): numpy float16 (4,L), or (1,4L) with `flatten`."""
    if seq_len is None:
        seq_len = len(seq)
    if seq_len == 0:
        import numpy as np
        x = np.zeros((4, 0), dtype="float16")
    else:
        x = one_hot_batch([seq], seq_len, device, torch.float16)[0].cpu().numpy()
    if flatten:
        x = x.flatten()[None, :]
    return x
