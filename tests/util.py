"""Shared helpers for the test-suite: golden fixture loading and seeded synthetic inputs."""
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)
    return {k: z[k] for k in z.files}


def state_dict_from(g, prefix="sd."):
    return {k[len(prefix):]: torch.from_numpy(np.array(v)) for k, v in g.items() if k.startswith(prefix)}


def onehot_from_codes(codes):
    """uint8 (n,L) with 0..3 = A,C,G,T and 4 = N  ->  float32 (n,4,L)."""
    n, L = codes.shape
    x = np.zeros((n, 4, L), np.float32)
    for c in range(4):
        x[:, c, :] = (codes == c)
    return x


def random_codes(n, L, seed, n_frac=0.0):
    g = np.random.default_rng(seed)
    c = g.integers(0, 4, size=(n, L)).astype(np.uint8)
    if n_frac > 0:
        c[g.random((n, L)) < n_frac] = 4
    return c


EVAL_GOLDENS = ["eval_cfg1_ckpt", "eval_u40_f20", "eval_u7_f8", "eval_u33_f24"]
TRAIN_GOLDENS = ["train_u6_bce", "train_u40_mse", "train_u33_bce"]
