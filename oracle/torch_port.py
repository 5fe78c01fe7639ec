"""Plain-PyTorch CPU restatement ("port") of the reference model, training step and
interpretation scan.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): it is the fp32 checker
for the floating-point kernels and the timed CPU baseline of bench.py (`cpu_baseline.kind =
"port"`, `--impl reference`); the product path never imports it.

The reference is pure Python on top of torch.nn, so the arithmetic below is executed by the very
same ATen CPU kernels the reference triggers; what is restated is the module graph:
  explainn/models/networks.py:73-145   unit stack for every num_fc and pool size / stride (module indices kept
                                       identical, so reference state_dicts load unchanged)
  explainn/models/networks.py:147      final Linear(U, C)
  explainn/models/networks.py:152-162  forward: tile the 4 input rows U times, stack, final
  explainn/train/train.py:54-73        one optimisation step
  explainn/interpretation/interpretation.py:126-149  activation scan (first three modules, float16)
Pinned against the reference by tests/golden/make_golden.py + tests/test_oracle_golden.py.
"""
import numpy as np
import torch
import torch.nn as nn

HIDDEN = 100  # networks.py:106 hard-codes 100 hidden channels per unit


class _Exp(nn.Module):
    def forward(self, x):
        return torch.exp(x)


class _AddTrailingDim(nn.Module):
    def forward(self, x):
        return x.unsqueeze(-1)


def pooled_len(L, F, pool_size=7, pool_stride=7):
    return int(((L - (F - 1)) - (pool_size - 1) - 1) / pool_stride + 1)


class PortExplaiNN(nn.Module):
    def __init__(self, num_cnns, input_length, num_classes, filter_size=19, pool_size=7, pool_stride=7, num_fc=2):
        super().__init__()
        U, P = num_cnns, pooled_len(input_length, filter_size, pool_size, pool_stride)
        self.U = U
        head = [nn.Conv1d(4 * U, U, filter_size, groups=U),  # 0
                nn.BatchNorm1d(U),                            # 1
                _Exp()]                                       # 2
        if num_fc == 0:          # networks.py:73-80
            tail = [nn.MaxPool1d(input_length - (filter_size - 1)), nn.Flatten()]
        elif num_fc == 1:        # networks.py:82-95
            tail = [nn.MaxPool1d(pool_size, pool_stride), nn.Flatten(), _AddTrailingDim(),
                    nn.Conv1d(P * U, U, 1, groups=U),                 # 6
                    nn.BatchNorm1d(U, 1e-05, 0.1, True),              # 7
                    nn.ReLU(), nn.Flatten()]
        elif num_fc == 2:        # networks.py:96-116
            tail = [nn.MaxPool1d(pool_size, pool_stride),             # 3
                    nn.Flatten(),                                     # 4
                    _AddTrailingDim(),                                # 5
                    nn.Conv1d(P * U, HIDDEN * U, 1, groups=U),        # 6
                    nn.BatchNorm1d(HIDDEN * U, 1e-05, 0.1, True),     # 7
                    nn.ReLU(),                                        # 8
                    nn.Dropout(0.3),                                  # 9
                    nn.Conv1d(HIDDEN * U, U, 1, groups=U),            # 10
                    nn.BatchNorm1d(U, 1e-05, 0.1, True),              # 11
                    nn.ReLU(),                                        # 12
                    nn.Flatten()]                                     # 13
        else:                    # networks.py:117-145
            tail = [nn.MaxPool1d(pool_size, pool_stride), nn.Flatten(), _AddTrailingDim(),
                    nn.Conv1d(P * U, HIDDEN * U, 1, groups=U),        # 6
                    nn.BatchNorm1d(HIDDEN * U, 1e-05, 0.1, True),     # 7
                    nn.ReLU()]                                        # 8
            self.linears_bg = nn.ModuleList([nn.Sequential(nn.Dropout(0.3), nn.Conv1d(HIDDEN * U, HIDDEN * U, 1, groups=U),
                                                           nn.BatchNorm1d(HIDDEN * U, 1e-05, 0.1, True), nn.ReLU())
                                             for _ in range(num_fc - 2)])
            self.last_linear = nn.Sequential(nn.Dropout(0.3), nn.Conv1d(HIDDEN * U, U, 1, groups=U),
                                             nn.BatchNorm1d(U, 1e-05, 0.1, True), nn.ReLU(), nn.Flatten())
        self.num_fc = num_fc
        self.linears = nn.Sequential(*(head + tail))
        self.final = nn.Linear(U, num_classes)

    def tile(self, x):
        return x.repeat(1, self.U, 1)

    def forward(self, x):
        outs = self.linears(self.tile(x))
        if self.num_fc > 2:      # networks.py:152-162
            for blk in self.linears_bg:
                outs = blk(outs)
            outs = self.last_linear(outs)
        return self.final(outs)

    # pieces used by the parity tests
    def conv_bn(self, x):
        return self.linears[:2](self.tile(x))

    def activations(self, x):
        return self.linears[:3](self.tile(x))

    def pooled(self, x):
        return self.linears[:4](self.tile(x))

    def unit_outputs(self, x):
        return self.linears(self.tile(x))


def train_step(model, x, y, criterion, optimizer):
    """train.py:59-67 for one batch; returns the loss tensor."""
    optimizer.zero_grad()
    out = model(x)
    loss = criterion(out, y)
    loss.backward()
    optimizer.step()
    return loss


def unit_activations_fp16(model, x, batch=128):
    """interpretation.py:143-147 on an in-memory tensor."""
    outs = []
    with torch.no_grad():
        for i in range(0, x.shape[0], batch):
            outs.append(model.activations(x[i:i + batch]).numpy().astype(np.half))
    return np.concatenate(outs, 0)


def synthetic_onehot(n, L, seed=1714, n_frac=0.0, device="cpu"):
    """i.i.d. uniform A/C/G/T one-hot (n,4,L) fp32; a fraction n_frac of columns zeroed (N)."""
    g = torch.Generator().manual_seed(seed)
    codes = torch.randint(0, 4, (n, L), generator=g)
    x = torch.zeros(n, 4, L)
    x.scatter_(1, codes.unsqueeze(1), 1.0)
    if n_frac > 0:
        keep = (torch.rand(n, 1, L, generator=g) >= n_frac).float()
        x = x * keep
    return x.to(device)
