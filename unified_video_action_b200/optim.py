"""Fused AdamW + EMA step for the head parameters, backed by ``uva_adamw_ema_step``.

Mirrors what the reference does to these parameters every training step — ``torch.optim.AdamW`` with
the parameter groups of ``add_weight_decay`` (policy/unified_video_action_policy.py:325-360: no decay on
1-D tensors and biases) followed by ``EMAModel.step`` (model/autoregressive/ema_model.py:40-89) — in one
HBM pass instead of ~10.  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes
from typing import Iterable, List, Optional

import torch
import torch.nn as nn

from . import _lib


class _Hyper(ctypes.Structure):
    _fields_ = [(n, ctypes.c_float) for n in ("lr", "beta1", "beta2", "eps", "ema_decay", "grad_scale")] + [("step", ctypes.c_int32)]


def ema_decay_at(optimization_step: int, update_after_step=0, inv_gamma=1.0, power=2 / 3, min_value=0.0, max_value=0.9999) -> float:
    """EMAModel.get_decay (ema_model.py:46-56)."""
    step = max(0, optimization_step - update_after_step - 1)
    if step <= 0:
        return 0.0
    value = 1 - (1 + step / inv_gamma) ** -power
    return max(min_value, min(value, max_value))


class HeadAdamW:
    """AdamW (+ optional EMA) over the parameters of one or more heads.

    ``weight_decay`` applies to tensors with more than one dimension whose name does not end in
    ``.bias`` (add_weight_decay, POL:325-341).  ``ema_modules`` are deep copies of ``modules`` that
    receive the moving average (``EMAModel.averaged_model``)."""

    def __init__(self, modules: Iterable[nn.Module], lr: float, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-2, ema_modules: Optional[Iterable[nn.Module]] = None, **ema_kwargs):
        self.modules = list(modules)
        self.ema_modules = list(ema_modules) if ema_modules is not None else None
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        self.ema_kwargs = ema_kwargs
        self.params: List[torch.nn.Parameter] = []
        self.wd: List[float] = []
        self.ema: List[Optional[torch.Tensor]] = []
        for mi, m in enumerate(self.modules):
            ema_named = dict(self.ema_modules[mi].named_parameters()) if self.ema_modules is not None else {}
            for name, p in m.named_parameters():
                if not p.requires_grad:
                    continue
                if p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("HeadAdamW needs contiguous fp32 parameters")
                self.params.append(p)
                self.wd.append(0.0 if (p.dim() == 1 or name.endswith(".bias")) else weight_decay)
                self.ema.append(ema_named.get(name))
        self.exp_avg = [torch.zeros_like(p, memory_format=torch.contiguous_format) for p in self.params]
        self.exp_avg_sq = [torch.zeros_like(p, memory_format=torch.contiguous_format) for p in self.params]
        self.step_count = 0          # optimizer steps taken (torch's state["step"])
        self.optimization_step = 0   # EMAModel.optimization_step

    def zero_grad(self, set_to_none: bool = True):
        for p in self.params:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()

    @torch.no_grad()
    def step(self, grad_scale: float = 1.0):
        live = [i for i, p in enumerate(self.params) if p.grad is not None]
        if not live:
            return
        dev = self.params[live[0]].device
        if dev.type != "cuda":
            raise RuntimeError("HeadAdamW runs on CUDA (sm_100a) only; there is no CPU fallback")
        self.step_count += 1
        decay = ema_decay_at(self.optimization_step, **self.ema_kwargs) if self.ema_modules is not None else 0.0
        grads = []
        for i in live:
            g = self.params[i].grad
            if g.dtype != torch.float32 or not g.is_contiguous():
                g = g.float().contiguous()
            grads.append(g)
        n = len(live)
        vp = ctypes.c_void_p
        arr = lambda ts: (vp * n)(*[None if t is None else t.data_ptr() for t in ts])
        hp = _Hyper(self.lr, self.betas[0], self.betas[1], self.eps, decay, grad_scale, self.step_count)
        numel = (ctypes.c_int64 * n)(*[self.params[i].numel() for i in live])
        wd = (ctypes.c_float * n)(*[self.wd[i] for i in live])
        with torch.cuda.device(dev):
            L = _lib.lib()
            rc = L.uva_adamw_ema_step(n, arr([self.params[i] for i in live]), arr(grads),
                                      arr([self.exp_avg[i] for i in live]), arr([self.exp_avg_sq[i] for i in live]),
                                      arr([self.ema[i] for i in live]) if self.ema_modules is not None else None,
                                      numel, wd, ctypes.byref(hp), torch.cuda.current_stream().cuda_stream)
        if rc != 0:
            raise RuntimeError(f"uva_adamw_ema_step failed ({rc}): {L.uva_last_error().decode()}")
        self.optimization_step += 1
        # the parameters changed behind torch's version counters: make the heads re-pack their operands
        for m in self.modules + (self.ema_modules or []):
            for sub in m.modules():
                if hasattr(sub, "_packed_key"):
                    sub._packed_key = None
                if "_pn_key" in sub.__dict__:  # packed operands of the fused DiffActLoss pre-net
                    sub.__dict__["_pn_key"] = None
