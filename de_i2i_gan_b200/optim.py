"""Flat-buffer Adam on the fused kernel (dei2i_adam): the drop-in for ``optim.Adam(network.parameters(), lr,
betas=(0.5, 0.999))`` built at trainers/base_trainer.py:68-89.

All parameters of a network are re-homed into ONE contiguous FP32 vector (``p.data`` become views), their gradients
into a second one (``p.grad`` are views, zeroed by one memset, accumulated into by autograd), Adam state into two
more.  ``step()`` is then one kernel launch per parameter group, and the gradient vector is exactly what the
data-parallel bucket all-reduce (de_i2i_gan_b200/parallel.py) sends over NCCL - no gather/scatter copies.

Semantics vs torch.optim.Adam: identical update rule (bias correction, eps outside the sqrt, optional L2 weight decay, no
amsgrad).  torch skips parameters whose ``grad is None``; here such parameters see a zero gradient, for which the Adam
update is exactly zero (m = v = 0  =>  step = 0 / (0 + eps)), so unused parameters (the generator's never-executed
``norm_s`` / ``conv_s``) stay bit-identical as well.
"""
from typing import List

import torch

from . import ops


def _align(n: int, a: int = 64) -> int:
    return (n + a - 1) // a * a


class FlatAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, decoupled=False):
        """decoupled=True: torch.optim.AdamW semantics (p *= 1 - lr * weight_decay before the Adam update) instead of Adam's L2 term."""
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, decoupled=decoupled)
        super().__init__(params, defaults)
        self._flat = []  # per group: dict(p=, g=, m=, v=, step=)
        for group in self.param_groups:
            self._flat.append(self._flatten(group["params"]))

    def add_param_group(self, param_group):
        """torch.optim.Optimizer.add_param_group + a flat buffer for the new group (MAETrainer adds the mask token's parameters to
        the generator's optimizer, mae_trainer.py:28).  A group without parameters (non-learnable token types) is ignored."""
        params = param_group["params"]
        params = [params] if isinstance(params, torch.Tensor) else list(params)
        if not params:
            return
        param_group = dict(param_group, params=params)
        super().add_param_group(param_group)
        if hasattr(self, "_flat") and len(self._flat) < len(self.param_groups):
            self._flat.append(self._flatten(self.param_groups[-1]["params"]))

    @staticmethod
    def _flatten(params: List[torch.Tensor]):
        assert len(params) > 0
        dev = params[0].device
        assert all(p.dtype == torch.float32 and p.device == dev for p in params), "FlatAdam: FP32 parameters on one device"
        offs, total = [], 0
        for p in params:
            offs.append(total)
            total += _align(p.numel())  # 256-byte aligned slices: vector loads and NCCL like it
        flat_p = torch.zeros(total, dtype=torch.float32, device=dev)
        flat_g = torch.zeros(total, dtype=torch.float32, device=dev)
        with torch.no_grad():
            for p, o in zip(params, offs):
                flat_p[o:o + p.numel()].copy_(p.data.reshape(-1))
                p.data = flat_p[o:o + p.numel()].view(p.shape)
                p.grad = flat_g[o:o + p.numel()].view(p.shape)
        return dict(p=flat_p, g=flat_g, m=torch.zeros_like(flat_p), v=torch.zeros_like(flat_p), step=0, offsets=offs,
                    params=params)

    # -- introspection used by the data-parallel reducer and by the tests
    @property
    def flat_grads(self):
        return [f["g"] for f in self._flat]

    @property
    def flat_params(self):
        return [f["p"] for f in self._flat]

    def slices(self, gi=0):
        f = self._flat[gi]
        return [(p, o, p.numel()) for p, o in zip(f["params"], f["offsets"])]

    def _rebind(self):
        """Re-attach gradient views if something replaced them (e.g. a foreign ``zero_grad(set_to_none=True)``)."""
        for f in self._flat:
            for p, o in zip(f["params"], f["offsets"]):
                want = f["g"][o:o + p.numel()]
                if p.grad is None:
                    p.grad = want.view(p.shape)
                elif p.grad.data_ptr() != want.data_ptr():
                    want.copy_(p.grad.reshape(-1))
                    p.grad = want.view(p.shape)
                if p.data.data_ptr() != f["p"][o:o + p.numel()].data_ptr():  # e.g. load_state_dict keeps storage; .to() does not
                    f["p"][o:o + p.numel()].copy_(p.data.reshape(-1))
                    p.data = f["p"][o:o + p.numel()].view(p.shape)

    def zero_grad(self, set_to_none: bool = False):
        """One memset per group.  ``set_to_none`` is accepted for API compatibility and ignored: the gradient views must
        stay attached so that autograd accumulates straight into the flat vector."""
        self._rebind()
        for f in self._flat:
            f["g"].zero_()

    def advance(self):
        """Host half of a step: counts the step and refreshes the two per-step scalars (lr / (1 - beta1^t) and
        1 / sqrt(1 - beta2^t), computed in double like torch.optim.Adam) in DEVICE memory.  ``step()`` calls it itself, except while a CUDA graph is being captured: a graphed trainer step calls
        ``advance()`` before every replay, the captured ``dei2i_adam_dev`` launch then reads the fresh values."""
        for group, f in zip(self.param_groups, self._flat):
            f["step"] += 1
            b1, b2 = group["betas"]
            t = f["step"]
            if "hyper" not in f:
                f["hyper"] = torch.zeros(3, dtype=torch.float32, device=f["p"].device)
            decay = 1.0 - float(group["lr"]) * float(group.get("weight_decay", 0.0)) if group.get("decoupled", False) else 1.0
            # pageable source: the runtime stages the 12 bytes before returning, so the host may run ahead safely
            f["hyper"].copy_(torch.tensor([float(group["lr"]) / (1.0 - b1 ** t), 1.0 / (1.0 - b2 ** t) ** 0.5, decay], dtype=torch.float32))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        capturing = torch.cuda.is_available() and torch.cuda.is_current_stream_capturing()
        if not capturing:
            self._rebind()
            self.advance()
        for group, f in zip(self.param_groups, self._flat):
            b1, b2 = group["betas"]
            if group.get("weight_decay", 0.0) and not group.get("decoupled", False):  # torch.optim.Adam: grad += wd * param (StarGAN-v2 solver.py:72-80)
                f["g"].add_(f["p"], alpha=float(group["weight_decay"]))
            ops.adam_dev(f["p"], f["g"], f["m"], f["v"], f["hyper"], float(b1), float(b2), float(group["eps"]))
        return loss


class FlatRMSprop(FlatAdam):
    """``optim.RMSprop(params, lr)`` (alpha 0.99, eps 1e-8, no momentum) on the fused kernel - the WGAN trainer's
    optimizer (options/wgan_options.py:49-57).  ``clamp_(lo, hi)`` is WGanModel.weight_clipping (wgan_model.py:18-21)
    as one launch over the flat parameter vector."""

    def __init__(self, params, lr=1e-2, alpha=0.99, eps=1e-8):
        torch.optim.Optimizer.__init__(self, params, dict(lr=lr, alpha=alpha, eps=eps))
        self._flat = [self._flatten(g["params"]) for g in self.param_groups]

    def advance(self):
        """Host half of a step (see FlatAdam.advance): the learning rate goes to device memory so that a captured step sees the
        scheduler's updates."""
        for group, f in zip(self.param_groups, self._flat):
            f["step"] += 1
            if "hyper" not in f:
                f["hyper"] = torch.zeros(1, dtype=torch.float32, device=f["p"].device)
            f["hyper"].copy_(torch.tensor([float(group["lr"])], dtype=torch.float32))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        capturing = torch.cuda.is_available() and torch.cuda.is_current_stream_capturing()
        if not capturing:
            self._rebind()
            self.advance()
        for group, f in zip(self.param_groups, self._flat):
            ops.rmsprop_dev(f["p"], f["g"], f["v"], f["hyper"], float(group["alpha"]), float(group["eps"]))
        return loss

    @torch.no_grad()
    def clamp_(self, lo, hi):
        for f in self._flat:
            ops.clamp_(f["p"], lo, hi)
