"""Step tail: global gradient-norm clip + AdamW on bf16 parameters in two multi-tensor kernels.

Mirrors what the reference's step does after backward -- ``accelerator.clip_grad_norm_(params,
max_grad_norm=1.0)`` (training.py:264 -> SP/transformers/trainer.py:2489-2499 ->
``torch.nn.utils.clip_grad_norm_``) and ``torch.optim.AdamW.step`` -- with the same interface as a
torch optimizer, plus :meth:`B200AdamW.clip_and_step` which fuses the clip coefficient into the
update so the gradients are read twice and never rewritten."""
from __future__ import annotations

import ctypes
from typing import Iterable, Optional

import torch

from . import _cabi, ops
from ._cabi import check

__all__ = ["B200AdamW"]


class B200AdamW(torch.optim.Optimizer):
    """AdamW for bf16 CUDA parameters (bf16 moments, fp32 arithmetic -- torch's fused AdamW math).

    ``step()`` is a plain AdamW step; ``clip_and_step(max_norm)`` is ``clip_grad_norm_`` + ``step()``
    and returns the total gradient norm (fp32 device scalar), leaving ``.grad`` untouched."""

    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 1e-2):
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._t = 0

    def _tensors(self, group):
        ps = [p for p in group["params"] if p.grad is not None]
        for p in ps:
            if not p.is_cuda or p.dtype != torch.bfloat16 or p.grad.dtype != torch.bfloat16:
                raise _cabi.B200SftError("B200AdamW handles bf16 CUDA parameters with bf16 gradients only")
            if not p.is_contiguous() or not p.grad.is_contiguous():
                raise _cabi.B200SftError("B200AdamW needs contiguous parameters and gradients")
            st = self.state[p]
            if not st:
                st["exp_avg"] = torch.zeros_like(p)
                st["exp_avg_sq"] = torch.zeros_like(p)
        return ps

    @staticmethod
    def _ptr_array(tensors):
        return (ctypes.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])

    @torch.no_grad()
    def grad_norm(self) -> torch.Tensor:
        """sqrt(sum of squares of every gradient) as an fp32 device scalar (fixed summation order)."""
        grads = [p.grad for g in self.param_groups for p in g["params"] if p.grad is not None]
        dev = grads[0].device
        lib = _cabi.load()
        numels = (ctypes.c_size_t * len(grads))(*[g.numel() for g in grads])
        with torch.cuda.device(dev):
            ws_bytes = lib.b200sft_grad_norm_workspace_bytes(numels, len(grads))
            ws = torch.empty(max(ws_bytes, 8), dtype=torch.uint8, device=dev)
            total = torch.empty((), dtype=torch.float32, device=dev)
            check(lib.b200sft_grad_norm_bf16(self._ptr_array(grads), numels, len(grads), total.data_ptr(), ws.data_ptr(),
                                             ws_bytes, torch.cuda.current_stream(dev).cuda_stream), "b200sft_grad_norm_bf16")
            ops.STATS.kernel_launches += 2
        return total

    @torch.no_grad()
    def clip_and_step(self, max_norm: Optional[float] = None) -> Optional[torch.Tensor]:
        total = self.grad_norm() if (max_norm is not None and max_norm > 0) else None
        self._t += 1
        self._opt_called = True  # lets torch LR schedulers see that an optimizer step happened
        lib = _cabi.load()
        for group in self.param_groups:
            ps = self._tensors(group)
            if not ps:
                continue
            dev = ps[0].device
            numels = (ctypes.c_size_t * len(ps))(*[p.numel() for p in ps])
            b1, b2 = group["betas"]
            with torch.cuda.device(dev):
                check(lib.b200sft_adamw_clip_bf16(
                    self._ptr_array(ps), self._ptr_array([p.grad for p in ps]),
                    self._ptr_array([self.state[p]["exp_avg"] for p in ps]),
                    self._ptr_array([self.state[p]["exp_avg_sq"] for p in ps]), numels, len(ps), float(group["lr"]),
                    float(b1), float(b2), float(group["eps"]), float(group["weight_decay"]), self._t,
                    float(max_norm) if total is not None else 0.0, None if total is None else total.data_ptr(),
                    torch.cuda.current_stream(dev).cuda_stream), "b200sft_adamw_clip_bf16")
                ops.STATS.kernel_launches += 1
        return total

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        self.clip_and_step(None)
        return loss
