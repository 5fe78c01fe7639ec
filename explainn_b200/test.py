"""test.run_test -- drop-in for explainn/train/test.py:12-33 of the reference."""
import numpy as np
import torch
import torch.nn as nn


def run_test(model, data_loader, device, isSigmoid=False):
    """
    :param model: ExplaiNN model
    :param data_loader: torch DataLoader, data loader with the sequences of interest
    :param device: current available device ('cuda:0')
    :param isSigmoid: boolean, True if the model output is binary
    :return: (labels, outputs) numpy arrays -- same order as the reference (test.py:33)
    """
    outs, labels = [], []
    with torch.no_grad():
        for seq, lbl in data_loader:
            seq = seq.to(device, non_blocking=True)
            outs.append(model(seq))     # stays on the device: one D2H at the end instead of one per batch
            labels.append(lbl)
    if hasattr(model, "check_inputs"):
        model.check_inputs()
    if not outs:
        return np.array([]), np.array([])
    outputs = torch.cat(outs, 0).cpu()
    if isSigmoid:
        outputs = nn.Sigmoid()(outputs)     # on the host, as test.py:28-30 of the reference
    return torch.cat([l.cpu() for l in labels], 0).numpy(), outputs.numpy()
