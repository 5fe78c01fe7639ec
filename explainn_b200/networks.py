"""networks.ExplaiNN -- drop-in for explainn/models/networks.py:44-162 of the reference, running on the
hand-written sm_100a kernels of libexplainn_b200.so.

The module keeps the reference's attribute layout (`.linears` Sequential with the same module indices,
`.final`, `._options`), so reference state_dicts load unchanged (same 23 keys, same order, same shapes)
and callers that poke into the module (`model.linears[0].weight`, `model.final.weight.data.clamp_(0)`,
`model.linears[:3](x.repeat(1, U, 1))`, `model.linears(x.repeat(1, U, 1))`) keep working.  The contained
torch.nn modules are parameter containers: their own forward is never used on the fused paths.
"""
import torch
import torch.nn as nn

from . import functional as Fn
from ._cabi import on_device_of

HIDDEN = Fn.HIDDEN


class ExpActivation(nn.Module):
    """Exponential activation (networks.py:21-29 of the reference)."""

    def forward(self, x):
        return torch.exp(x)


class Unsqueeze(nn.Module):
    """Adds a trailing dimension (networks.py:32-37 of the reference)."""

    def forward(self, x):
        return x.unsqueeze(-1)


def _require_cuda(x):
    if not x.is_cuda:
        raise RuntimeError("explainn_b200 runs only on CUDA devices (sm_100a kernels; there is no CPU fallback). "
                           "Got input on %s." % x.device)


class _UnitStack(nn.Sequential):
    """`model.linears`: conv -> BN -> exp -> pool -> per-unit FC stack (networks.py:97-116).

    Calling it (or a `[:3]` / `[:4]` prefix of it) on the U-times tiled input the reference passes
    (`x.repeat(1, num_cnns, 1)`, interpretation.py:94-95,144-145) or on the plain (B,4,L) input runs
    the fused kernels.
    """

    def __init__(self, *mods):
        super().__init__(*mods)
        self._owner = None     # set by ExplaiNN (plain attribute: not a submodule, not in state_dict)

    def _bind(self, owner):
        object.__setattr__(self, "_owner", owner)

    def forward(self, x):
        return self._owner._unit_outputs(x)

    def __getitem__(self, idx):
        if isinstance(idx, slice):
            start, stop, step = idx.indices(len(self))
            if start == 0 and step == 1 and stop in (3, 4) and self._owner is not None:
                return _UnitStackPrefix(self._owner, stop)
            return nn.Sequential(*list(self._modules.values())[idx])
        return super().__getitem__(idx)


class _UnitStackPrefix(nn.Module):
    """`model.linears[:3]` (conv->BN->exp, (B,U,S)) and `model.linears[:4]` (+maxpool, (B,U,P))."""

    def __init__(self, owner, stop):
        super().__init__()
        object.__setattr__(self, "_owner", owner)
        self._stop = stop

    def forward(self, x):
        if self._stop == 3:
            return self._owner._activations(x, torch.float32)
        return self._owner._pooled(x)


class ExplaiNN(nn.Module):
    """The ExplaiNN model (reference: explainn/models/networks.py:44-162)."""

    def __init__(self, num_cnns, input_length, num_classes, filter_size=19, num_fc=2, pool_size=7, pool_stride=7,
                 weight_path=None):
        """Same positional / keyword arguments as the reference constructor (networks.py:48-49).

        num_cnns independent units, each one filter of width filter_size over a length-input_length one-hot
        sequence, max-pooling (pool_size, pool_stride), num_fc per-unit fully connected layers, and one Linear over
        the unit outputs to num_classes logits.  weight_path, if given, is a state_dict file of the reference (or of
        this class) and is loaded at the end of construction.  Limits of the kernels are checked here:
        filter_size <= 24, input_length <= 2048 (see the NotImplementedError messages)."""
        super().__init__()
        self._options = {
            "num_cnns": num_cnns,
            "input_length": input_length,
            "num_classes": num_classes,
            "filter_size": filter_size,
            "num_fc": num_fc,
            "pool_size": pool_size,
            "pool_stride": pool_stride,
            "weight_path": weight_path,
        }
        if num_fc < 0:
            raise ValueError("num_fc must be >= 0")
        if pool_size < 1 or pool_stride < 1:
            raise ValueError("pool_size and pool_stride must be >= 1")
        if filter_size > 24:
            raise NotImplementedError("explainn_b200: filter_size <= 24 (k-mer tables of the fused conv kernel)")
        if input_length > 2048:
            raise NotImplementedError("explainn_b200: input_length <= 2048 (a packed sequence is prefetched into the "
                                      "registers of one warp by the scan kernels)")
        U = num_cnns
        S = input_length - (filter_size - 1)
        P = 1 if num_fc == 0 else Fn.pooled_len(input_length, filter_size, pool_size, pool_stride)
        if S < 1 or P < 1:
            raise ValueError("input_length %d too short for filter_size %d and pool size %d" % (input_length, filter_size, pool_size))
        if num_fc == 0:          # networks.py:73-80: global max-pool, no FC
            self.linears = _UnitStack(
                nn.Conv1d(in_channels=4 * U, out_channels=U, kernel_size=filter_size, groups=U),   # 0
                nn.BatchNorm1d(U),                                                                  # 1
                ExpActivation(),                                                                    # 2
                nn.MaxPool1d(S),                                                                    # 3
                nn.Flatten())                                                                       # 4
        elif num_fc == 1:        # networks.py:82-95
            self.linears = _UnitStack(
                nn.Conv1d(in_channels=4 * U, out_channels=U, kernel_size=filter_size, groups=U),   # 0
                nn.BatchNorm1d(U),                                                                  # 1
                ExpActivation(),                                                                    # 2
                nn.MaxPool1d(pool_size, pool_stride),                                               # 3
                nn.Flatten(),                                                                       # 4
                Unsqueeze(),                                                                        # 5
                nn.Conv1d(in_channels=P * U, out_channels=U, kernel_size=1, groups=U),              # 6
                nn.BatchNorm1d(U, 1e-05, 0.1, True),                                                # 7
                nn.ReLU(),                                                                          # 8
                nn.Flatten())                                                                       # 9
        elif num_fc == 2:
            self.linears = _UnitStack(
                nn.Conv1d(in_channels=4 * U, out_channels=U, kernel_size=filter_size, groups=U),   # 0
                nn.BatchNorm1d(U),                                                                  # 1
                ExpActivation(),                                                                    # 2
                nn.MaxPool1d(pool_size, pool_stride),                                               # 3
                nn.Flatten(),                                                                       # 4
                Unsqueeze(),                                                                        # 5
                nn.Conv1d(in_channels=P * U, out_channels=HIDDEN * U, kernel_size=1, groups=U),     # 6
                nn.BatchNorm1d(HIDDEN * U, 1e-05, 0.1, True),                                       # 7
                nn.ReLU(),                                                                          # 8
                nn.Dropout(0.3),                                                                    # 9
                nn.Conv1d(in_channels=HIDDEN * U, out_channels=U, kernel_size=1, groups=U),         # 10
                nn.BatchNorm1d(U, 1e-05, 0.1, True),                                                # 11
                nn.ReLU(),                                                                          # 12
                nn.Flatten())                                                                       # 13
        else:                    # networks.py:117-145: num_fc - 2 extra hidden layers between FC1 and the 100 -> 1 layer
            self.linears = _UnitStack(
                nn.Conv1d(in_channels=4 * U, out_channels=U, kernel_size=filter_size, groups=U),   # 0
                nn.BatchNorm1d(U),                                                                  # 1
                ExpActivation(),                                                                    # 2
                nn.MaxPool1d(pool_size, pool_stride),                                               # 3
                nn.Flatten(),                                                                       # 4
                Unsqueeze(),                                                                        # 5
                nn.Conv1d(in_channels=P * U, out_channels=HIDDEN * U, kernel_size=1, groups=U),     # 6
                nn.BatchNorm1d(HIDDEN * U, 1e-05, 0.1, True),                                       # 7
                nn.ReLU())                                                                          # 8
            self.linears_bg = nn.ModuleList([nn.Sequential(
                nn.Dropout(0.3),
                nn.Conv1d(in_channels=HIDDEN * U, out_channels=HIDDEN * U, kernel_size=1, groups=U),
                nn.BatchNorm1d(HIDDEN * U, 1e-05, 0.1, True),
                nn.ReLU()) for _ in range(num_fc - 2)])
            self.last_linear = nn.Sequential(
                nn.Dropout(0.3),
                nn.Conv1d(in_channels=HIDDEN * U, out_channels=U, kernel_size=1, groups=U),
                nn.BatchNorm1d(U, 1e-05, 0.1, True),
                nn.ReLU(),
                nn.Flatten())
        self.linears._bind(self)
        self.final = nn.Linear(U, num_classes)
        self._invalid = None      # lazily read counter of non-one-hot input columns
        # "tc": tcgen05/TMEM implicit-GEMM conv for inference and the training forward (fp32-rounding accurate);
        # "tc_fp16": inference only, one fp16 piece per weight (stated tolerance 2e-3, half the tensor-core work);
        # "lut": the k-mer table kernel (bit-identical to the CPU oracle).  The interpretation pass always uses "lut".
        self.conv_backend = "tc"
        # data parallel only: True = BatchNorm batch statistics over the batches of ALL ranks (the reference's single-device
        # maths on the concatenated batch; default unit only, eager execution); False = each rank's own batch (DDP default)
        self.sync_bn = False
        if weight_path:
            self.load_state_dict(torch.load(weight_path))

    # ------------------------------------------------------------------ helpers
    @property
    def pooled_length(self):
        o = self._options
        if o["num_fc"] == 0:
            return 1
        return Fn.pooled_len(o["input_length"], o["filter_size"], o["pool_size"], o["pool_stride"])

    def _pack(self, x):
        if isinstance(x, Fn.Packed):
            return x
        _require_cuda(x)
        o = self._options
        if x.dim() == 2 and x.dtype == torch.uint8:
            # (B, L) ASCII bases ("ACGT", anything else = N): encoded on the device, 16x fewer bytes than fp32 one-hot
            if x.shape[1] != o["input_length"]:
                raise ValueError("expected text input of shape (B, %d), got %s" % (o["input_length"], tuple(x.shape)))
            return Fn.encode_dna_array(x, o["input_length"], x.device)
        if x.dim() != 3 or x.shape[2] != o["input_length"]:
            raise ValueError("expected input of shape (B, 4, %d), got %s" % (o["input_length"], tuple(x.shape)))
        if self._invalid is None or self._invalid.device != x.device:
            self._invalid = torch.zeros(1, dtype=torch.int32, device=x.device)
        return Fn.pack_onehot(x, o["num_cnns"], self._invalid)

    def check_inputs(self):
        """Raise ValueError if any batch seen so far contained a column that is neither one-hot nor
        all-zero (synchronises the device; the drivers call it once per loop, not per batch)."""
        if self._invalid is not None:
            Fn.check_invalid(self._invalid)

    def _conv_prep_eval(self):
        conv, bn = self.linears[0], self.linears[1]
        return Fn.conv_prepare(conv.weight.detach(), conv.bias.detach(), bn.weight.detach(), bn.bias.detach(),
                               bn.running_mean, bn.running_var, bn.eps)

    def _bn_uses_batch_stats(self):
        return self.linears[1].training

    # ------------------------------------------------------------------ fused pieces (eval statistics)
    @on_device_of(1)
    def _activations(self, x, dtype=torch.float16):
        """exp(BN(conv(x))) as (B,U,S) -- interpretation.py:143-145."""
        if self._bn_uses_batch_stats():
            raise RuntimeError("the activation scan uses BatchNorm running statistics; call model.eval() first "
                               "(scripts/interpret.py:199 of the reference does)")
        pk = self._pack(x)
        return Fn.unit_activations(pk, self._conv_prep_eval(), dtype)

    @on_device_of(1)
    def _pooled(self, x):
        """max-pooled activations (B,U,P)."""
        return self._pooled_packed(self._pack(x))

    def _pooled_packed(self, pk):
        if self._bn_uses_batch_stats():
            raise RuntimeError("model.linears[:4] on the fused path needs eval() mode")
        o = self._options
        if o["num_fc"] == 0:       # MaxPool1d(S): exp of the maximum over all positions, (B,U,1)
            return Fn.pwm_score(pk, self._conv_prep_eval(), 2).unsqueeze(-1)
        if o["pool_size"] != Fn.POOL or o["pool_stride"] != Fn.POOL:
            # any other MaxPool1d (networks.py:102): pre-activation map from the table kernel, then the general pool
            z = Fn.unit_activations(pk, self._conv_prep_eval(), preact=True)
            return Fn.pool_general(z, o["pool_size"], o["pool_stride"])
        if self.conv_backend in ("tc", "tc_fp16"):
            conv, bn = self.linears[0], self.linears[1]
            tcp = Fn.conv_tc_prepare(conv.weight.detach(), conv.bias.detach(), bn.weight.detach(), bn.bias.detach(),
                                     bn.running_mean, bn.running_var, bn.eps)
            return Fn.conv_pool_fwd_tc(pk, tcp, fp16_weights=self.conv_backend == "tc_fp16")
        return Fn.conv_pool_fwd(pk, self._conv_prep_eval())

    @on_device_of(1)
    def _unit_outputs_eval(self, x):
        return self._unit_outputs_eval_packed(self._pack(x))

    def forward_packed(self, pk):
        """Eval-mode logits from already packed sequences (functional.Packed), e.g. the reverse strand."""
        unit = self._unit_outputs_eval_packed(pk)
        return Fn.linear_fwd(unit, self.final.weight.detach(), self.final.bias.detach())

    def _unit_outputs_eval_packed(self, pk):
        L = self.linears
        nfc = self._options["num_fc"]
        if nfc == 0:
            return self._pooled_packed(pk).squeeze(-1)
        pooled = self._pooled_packed(pk)
        if nfc == 1:
            bn = L[7]
            return Fn.fc1_stack_fwd_eval(pooled, L[6].weight.detach(), L[6].bias.detach(),
                                         (bn.weight.detach(), bn.bias.detach(), bn.running_mean, bn.running_var), bn.eps)
        if nfc > 2:
            return self._unit_outputs_eval_deep(pooled)
        bn2, bn3 = L[7], L[11]
        return Fn.fc_stack_fwd_eval(
            pooled, L[6].weight.detach(), L[6].bias.detach(),
            (bn2.weight.detach(), bn2.bias.detach(), bn2.running_mean, bn2.running_var),
            L[10].weight.detach(), L[10].bias.detach(),
            (bn3.weight.detach(), bn3.bias.detach(), bn3.running_mean, bn3.running_var), bn2.eps)

    @on_device_of(1)
    def _hidden_eval(self, X, fc, bn):
        return Fn.hidden_fwd_eval(X, fc.weight.detach(), fc.bias.detach(),
                                  (bn.weight.detach(), bn.bias.detach(), bn.running_mean, bn.running_var), bn.eps)

    def _unit_outputs_eval_deep(self, pooled):
        """num_fc > 2 in eval mode: hidden blocks, then the last hidden layer + last_linear in the FC-stack kernel."""
        from . import variants
        conv, bn1, fcs, bns, drops, last_fc, last_bn = variants.layout(self)
        X = pooled
        for fc, bn in zip(fcs[:-1], bns[:-1]):
            X = self._hidden_eval(X, fc, bn)
        fc, bn = fcs[-1], bns[-1]
        return Fn.fc_stack_fwd_eval(
            X, fc.weight.detach(), fc.bias.detach(), (bn.weight.detach(), bn.bias.detach(), bn.running_mean, bn.running_var),
            last_fc.weight.detach(), last_fc.bias.detach(),
            (last_bn.weight.detach(), last_bn.bias.detach(), last_bn.running_mean, last_bn.running_var), bn.eps)

    def _unit_outputs(self, x):
        if self._options["num_fc"] > 2 and not self.training:
            # the reference's `linears` ends behind the first hidden layer for num_fc > 2 (networks.py:117-129): calling it
            # returns those activations, shape (B, 100 * num_cnns, 1)
            X = self._hidden_eval(self._pooled(x), self.linears[6], self.linears[7])
            return X.reshape(X.shape[0], -1, 1)
        if self.training:
            from . import autograd_fn
            return autograd_fn.unit_outputs_train(self, x)
        return self._unit_outputs_eval(x)

    # ------------------------------------------------------------------ forward
    @on_device_of(1)
    def forward(self, x):
        """x: (B,4,L) one-hot float tensor on the GPU -> (B,C) logits (networks.py:152-162)."""
        if self.training:
            from . import autograd_fn
            return autograd_fn.forward_train(self, x)
        unit = self._unit_outputs_eval(x)
        return Fn.linear_fwd(unit, self.final.weight.detach(), self.final.bias.detach())


class PWM(nn.Module):
    """PWM (Position Weight Matrix) scoring -- drop-in for explainn/models/networks.py:393-437.

    Same constructor, `._options` and frozen `.conv1d` (weight (G,4,F) = the PWMs, bias 0) as the reference.  The
    forward scores both strands with the fused one-hot conv kernel (the reverse strand is the packed reverse
    complement, `_flip(_flip(x, 1), 2)` in the reference) and reduces over positions inside the kernel; neither the
    G-times tiled input nor the (B,G,2S) score tensor is built.
    """

    def __init__(self, pwms, input_length, scoring="sum"):
        super().__init__()
        pwms = torch.as_tensor(pwms, dtype=torch.float32)
        groups, _, kernel_size = pwms.shape
        self._options = {
            "groups": groups,
            "kernel_size": kernel_size,
            "sequence_length": input_length,
            "scoring": scoring,
        }
        if kernel_size > 24:
            raise NotImplementedError("explainn_b200: kernel_size <= 24 (k-mer tables of the fused conv kernel)")
        self.conv1d = nn.Conv1d(in_channels=4 * groups, out_channels=1 * groups, kernel_size=kernel_size, groups=groups)
        self.conv1d.bias.data = torch.zeros(groups)
        self.conv1d.weight.data = pwms.clone()
        for p in self.conv1d.parameters():
            p.requires_grad = False
        self._invalid = None

    def _prep(self):
        w = self.conv1d.weight.detach()
        G = w.shape[0]
        one = torch.ones(G, dtype=torch.float32, device=w.device)
        zero = torch.zeros(G, dtype=torch.float32, device=w.device)
        # identity BatchNorm (gamma 1, beta 0, mean 0, var 1, eps 0): the tables hold the PWM weights unchanged
        return Fn.conv_prepare(w, self.conv1d.bias.detach(), one, zero, zero, one, 0.0)

    def forward_packed(self, pk):
        prep = self._prep()
        use_max = self._options["scoring"] == "max"
        out = Fn.pwm_score(pk, prep, use_max)
        return Fn.pwm_score(Fn.revcomp_packed(pk), prep, use_max, out=out)

    @on_device_of(1)
    def forward(self, x):
        """x: (B,4,L) one-hot float tensor on the GPU -> (B,G) scores."""
        _require_cuda(x)
        if x.dim() != 3 or x.shape[1] != 4:
            raise ValueError("expected input of shape (B, 4, L), got %s" % (tuple(x.shape),))
        if x.shape[2] < self._options["kernel_size"]:
            raise RuntimeError("Calculated padded input size per channel: (%d). Kernel size: (%d). Kernel size can't be "
                               "greater than actual input size" % (x.shape[2], self._options["kernel_size"]))
        if self._invalid is None or self._invalid.device != x.device:
            self._invalid = torch.zeros(1, dtype=torch.int32, device=x.device)
        return self.forward_packed(Fn.pack_onehot(x, 1, self._invalid))
