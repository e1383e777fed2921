"""Torch-facing operators over the C ABI (include/b200sft.h).

Each function checks shapes/dtypes, takes device memory and the current stream from torch, and
calls libb200sft.so through ctypes with raw pointers.  They are registered as torch custom ops
(namespace ``b200sft``) so they compose with autograd and show up by name in profiler traces.

There is no fallback: CPU tensors are rejected and a missing library raises.
"""
from __future__ import annotations

from typing import NamedTuple, Optional, Tuple

import torch

from . import _cabi
from ._cabi import DENOM_DEVICE, DENOM_NONE, DENOM_VALID, FLAG_ROUND_LOGITS_BF16, check

IGNORE_INDEX = -100  # torch.nn.functional.cross_entropy default, SP/transformers/loss/loss_utils.py:28


class _Stats:
    """Bookkeeping for bench.py: how many of this library's kernels were launched, and (when
    ``events`` is a dict) CUDA-event pairs around the fused calls on their launch stream."""

    def __init__(self):
        self.kernel_launches = 0
        self.events = None  # None = off; else {"flce_fwd": [(start, end), ...], ...}

    def reset(self, with_events: bool = False):
        self.kernel_launches = 0
        self.events = {} if with_events else None

    def begin(self, name: str, device):
        if self.events is None:
            return None
        ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        ev[0].record(torch.cuda.current_stream(device))
        self.events.setdefault(name, []).append(ev)
        return ev

    def end(self, ev, device):
        if ev is not None:
            ev[1].record(torch.cuda.current_stream(device))

    def mean_ms(self, name: str) -> float:
        evs = (self.events or {}).get(name, [])
        return sum(a.elapsed_time(b) for a, b in evs) / len(evs) if evs else float("nan")


STATS = _Stats()


__all__ = [
    "label_prep", "label_frontend", "gather_rows", "compaction_list", "rmsnorm_gather", "rmsnorm_scatter_bwd",
    "LabelPlan", "flce_forward", "flce_backward", "compact_rows", "scatter_rows", "gemm_bf16", "scale_",
    "scale_cast_f32_to_bf16", "cast_bf16_to_f32", "fused_linear_cross_entropy", "IGNORE_INDEX",
    "FLAG_ROUND_LOGITS_BF16",
]


def _stream_ptr(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise _cabi.B200SftError("b200sft operators run on CUDA tensors only; there is no CPU fallback")


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


# --------------------------------------------------------------------------------------------
# label front-end
# --------------------------------------------------------------------------------------------
def label_prep(labels: torch.Tensor, ignore_index: int = IGNORE_INDEX) -> Tuple[torch.Tensor, torch.Tensor]:
    """Shift labels left by one and count valid tokens, on the device.

    Mirrors ``ForCausalLMLoss`` (SP/transformers/loss/loss_utils.py:57-60) and
    ``Trainer._get_num_items_in_batch`` (SP/transformers/trainer.py:2136).
    Returns ``(shifted int32 [B*T], counts int64 [2] = (unshifted valid, shifted valid))``.
    """
    _require_cuda(labels)
    if labels.dtype != torch.int64 or labels.dim() != 2:
        raise ValueError("labels must be an int64 [B, T] tensor")
    labels = labels.contiguous()
    B, T = labels.shape
    with torch.cuda.device(labels.device):
        shifted = torch.empty(B * T, dtype=torch.int32, device=labels.device)
        counts = torch.empty(2, dtype=torch.int64, device=labels.device)
        lib = _cabi.load()
        check(lib.b200sft_label_prep(_ptr(labels), B, T, ignore_index, _ptr(shifted), _ptr(counts),
                                     _stream_ptr(labels)), "b200sft_label_prep")
        STATS.kernel_launches += 1 if B * T > 0 else 0
    return shifted, counts


# --------------------------------------------------------------------------------------------
# fused LM-head + cross-entropy
# --------------------------------------------------------------------------------------------
class LabelPlan(NamedTuple):
    """Device outputs of :func:`label_frontend`."""
    shifted: torch.Tensor            # [B*T] int32 shifted labels
    counts: torch.Tensor             # [2] int64: non-ignored labels before / after the shift
    labels: Optional[torch.Tensor]   # [B,T] int64 unshifted labels (when asked for)
    row_index: Optional[torch.Tensor]  # compaction list (when asked for): original row of compact row j
    labels_c: Optional[torch.Tensor]   # compacted labels, ignore_index past the live count
    pos: Optional[torch.Tensor]        # compact position of each row, -1 if ignored
    n_live: Optional[torch.Tensor]     # int32 scalar


def label_frontend(input_ids: Optional[torch.Tensor] = None, labels: Optional[torch.Tensor] = None,
                   attention_mask: Optional[torch.Tensor] = None, completion_mask: Optional[torch.Tensor] = None,
                   position_ids: Optional[torch.Tensor] = None, ignore_index: int = IGNORE_INDEX,
                   compact: bool = True, return_labels: bool = False) -> LabelPlan:
    """Everything between the raw batch and the fused loss in ONE launch: the collator's masking
    (padding, completion-only, packed-sequence separators -- SP/transformers/data/data_collator.py:1426-1428),
    the ForCausalLMLoss shift (loss_utils.py:57-60), both valid counts (trainer.py:2136) and the
    live-row list the fused head skips masked rows with.  All inputs int64 ``[B,T]`` on the device;
    ``labels`` defaults to ``input_ids``; masks may be None."""
    src = labels if labels is not None else input_ids
    if src is None:
        raise ValueError("need labels or input_ids")
    given = [t for t in (input_ids, labels, attention_mask, completion_mask, position_ids) if t is not None]
    _require_cuda(*given)
    if src.dim() != 2:
        raise ValueError("expected [B, T] inputs")
    for t in given:
        if t.dtype != torch.int64 or t.shape != src.shape or not t.is_contiguous():
            raise ValueError("label front-end inputs must be contiguous int64 tensors of one [B, T] shape")
    B, T = src.shape
    lib = _cabi.load()
    dev = src.device
    with torch.cuda.device(dev):
        shifted = torch.empty(B * T, dtype=torch.int32, device=dev)
        counts = torch.empty(2, dtype=torch.int64, device=dev)
        labels_out = torch.empty(B, T, dtype=torch.int64, device=dev) if return_labels else None
        row_index = labels_c = pos = n_live = None
        if compact:
            row_index, labels_c, pos = (torch.empty(B * T, dtype=torch.int32, device=dev) for _ in range(3))
            n_live = torch.empty((), dtype=torch.int32, device=dev)
        check(lib.b200sft_label_frontend(_ptr(input_ids), _ptr(labels), _ptr(attention_mask), _ptr(completion_mask),
                                         _ptr(position_ids), B, T, ignore_index, _ptr(labels_out), _ptr(shifted),
                                         _ptr(counts), _ptr(row_index), _ptr(labels_c), _ptr(pos), _ptr(n_live),
                                         _stream_ptr(src)), "b200sft_label_frontend")
        if B * T:
            STATS.kernel_launches += 1
    return LabelPlan(shifted, counts, labels_out, row_index, labels_c, pos, n_live)


def gather_rows(x: torch.Tensor, plan: LabelPlan) -> torch.Tensor:
    """``xc[j] = x[plan.row_index[j]]`` for the live rows of a :class:`LabelPlan` (zero rows up to the
    next 128-row boundary): the gather half of :func:`compact_rows`."""
    _require_cuda(x, plan.row_index, plan.n_live)
    if x.dtype != torch.bfloat16 or x.dim() != 2 or not x.is_contiguous() or plan.row_index.shape != (x.shape[0],):
        raise ValueError("x must be contiguous bf16 [N,H] with one plan entry per row")
    N, H = x.shape
    if N == 0:
        return x
    lib = _cabi.load()
    with torch.cuda.device(x.device):
        xc = torch.empty_like(x)
        check(lib.b200sft_gather_rows(_ptr(x), _ptr(plan.row_index), _ptr(plan.n_live), N, H, _ptr(xc),
                                      _stream_ptr(x)), "b200sft_gather_rows")
        STATS.kernel_launches += 1
    return xc


def compaction_list(labels: torch.Tensor, ignore_index: int = IGNORE_INDEX) -> LabelPlan:
    """The live-row list of already-shifted int32 ``labels`` as a :class:`LabelPlan` (no counts, no gather)."""
    _require_cuda(labels)
    if labels.dtype != torch.int32 or labels.dim() != 1 or not labels.is_contiguous():
        raise ValueError("labels must be contiguous int32 [N]")
    N = labels.shape[0]
    dev = labels.device
    with torch.cuda.device(dev):
        row_index, labels_c, pos = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
        n_live = torch.zeros((), dtype=torch.int32, device=dev)
        if N:
            check(_cabi.load().b200sft_compact_labels(_ptr(labels), N, ignore_index, _ptr(row_index), _ptr(labels_c),
                                                      _ptr(pos), _ptr(n_live), _stream_ptr(labels)),
                  "b200sft_compact_labels")
            STATS.kernel_launches += 1
    return LabelPlan(labels, None, None, row_index, labels_c, pos, n_live)


def rmsnorm_gather(h: torch.Tensor, gamma: torch.Tensor, eps: float, plan: Optional[LabelPlan] = None
                   ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Final RMSNorm (SP/transformers/models/smollm3/modeling_smollm3.py:271-276) of the rows the loss
    scores, written compacted: ``xc[j] = gamma * bf16(h[row_index[j]] * rstd[j])``.  Without a plan (or
    a plan without a list) every row is normalised in place order.  Returns ``(xc [N,H] bf16, rstd [N] fp32)``."""
    _require_cuda(h, gamma)
    if h.dtype != torch.bfloat16 or gamma.dtype != torch.bfloat16 or h.dim() != 2 or gamma.shape != (h.shape[1],):
        raise ValueError("h must be bf16 [N,H] and gamma bf16 [H]")
    if not (h.is_contiguous() and gamma.is_contiguous()):
        raise ValueError("h and gamma must be contiguous")
    N, H = h.shape
    listed = plan is not None and plan.row_index is not None
    if listed and plan.row_index.shape != (N,):
        raise ValueError("label plan does not belong to these rows")
    with torch.cuda.device(h.device):
        xc = torch.empty_like(h)
        rstd = torch.empty(N, dtype=torch.float32, device=h.device)
        if N:
            check(_cabi.load().b200sft_rmsnorm_gather(_ptr(h), _ptr(gamma), float(eps),
                                                      _ptr(plan.row_index) if listed else None,
                                                      _ptr(plan.n_live) if listed else None, N, H, _ptr(xc), _ptr(rstd),
                                                      _stream_ptr(h)), "b200sft_rmsnorm_gather")
            STATS.kernel_launches += 1
    return xc, rstd


def rmsnorm_scatter_bwd(dyc: torch.Tensor, h: torch.Tensor, gamma: torch.Tensor, rstd: torch.Tensor,
                        plan: Optional[LabelPlan] = None, need_dh: bool = True, need_dgamma: bool = False
                        ) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
    """Backward of :func:`rmsnorm_gather`: ``dyc`` is the gradient w.r.t. the normalised compact rows
    (``flce_backward``'s dX); returns ``(dh [N,H] bf16 with zero rows where ignored, dgamma [H] bf16)``."""
    _require_cuda(dyc, h, gamma, rstd)
    if dyc.shape != h.shape or dyc.dtype != torch.bfloat16 or not dyc.is_contiguous():
        raise ValueError("dyc must be contiguous bf16 of h's shape")
    N, H = h.shape
    listed = plan is not None and plan.row_index is not None
    lib = _cabi.load()
    with torch.cuda.device(h.device):
        dh = torch.empty_like(h) if need_dh else None
        dgamma = torch.zeros_like(gamma) if need_dgamma else None
        if N and (need_dh or need_dgamma):
            ws_bytes = lib.b200sft_rmsnorm_bwd_workspace_bytes(H) if need_dgamma else 0
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=h.device) if ws_bytes else None
            check(lib.b200sft_rmsnorm_scatter_bwd(_ptr(dyc), _ptr(h), _ptr(gamma), _ptr(rstd),
                                                  _ptr(plan.pos) if listed else None,
                                                  _ptr(plan.row_index) if listed else None,
                                                  _ptr(plan.n_live) if listed else None, N, H, _ptr(dh), _ptr(dgamma),
                                                  _ptr(ws), ws_bytes, _stream_ptr(h)), "b200sft_rmsnorm_scatter_bwd")
            STATS.kernel_launches += (1 if need_dh else 0) + (2 if need_dgamma else 0)
    return dh, dgamma


def _check_flce_inputs(x: torch.Tensor, w: torch.Tensor, labels: torch.Tensor) -> Tuple[int, int, int]:
    _require_cuda(x, w, labels)
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16:
        raise ValueError("x and w must be bfloat16")
    if labels.dtype != torch.int32:
        raise ValueError("labels must be int32 (use label_prep)")
    if x.dim() != 2 or w.dim() != 2 or x.shape[-1] != w.shape[-1]:
        raise ValueError(f"x [N,H] / w [V,H] shape mismatch: {tuple(x.shape)} vs {tuple(w.shape)}")
    if labels.dim() != 1 or labels.shape[0] != x.shape[0]:
        raise ValueError("labels must be [N]")
    if not (x.is_contiguous() and w.is_contiguous() and labels.is_contiguous()):
        raise ValueError("x, w and labels must be contiguous")
    return x.shape[0], x.shape[1], w.shape[0]


def flce_forward(x: torch.Tensor, w: torch.Tensor, labels: torch.Tensor, scale: float = 1.0,
                 denom: Optional[torch.Tensor] = None, reduction_mean: bool = False,
                 ignore_index: int = IGNORE_INDEX, flags: int = 0, n_rows: Optional[torch.Tensor] = None,
                 with_stats: bool = False):
    """loss, lse, n_valid[, n_correct, entropy_sum] = fused lm_head + cross-entropy forward; no
    [N,V] tensor is allocated.

    ``loss = scale * sum_valid(lse - z_target) / D`` with D = 1, ``*denom`` (int64 device scalar,
    the trainer's num_items_in_batch) or the local valid-row count (``reduction_mean``).
    ``n_rows`` (int32 device scalar) = live rows when ``x``/``labels`` come from :func:`compact_rows`.
    ``with_stats`` adds the logits-free token-accuracy count and entropy sum.
    """
    N, H, V = _check_flce_inputs(x, w, labels)
    if denom is not None and reduction_mean:
        raise ValueError("pass either denom or reduction_mean")
    if denom is not None:
        _require_cuda(denom)
        if denom.dtype != torch.int64 or denom.numel() != 1:
            raise ValueError("denom must be a single int64 on the device")
    if n_rows is not None:
        _require_cuda(n_rows)
        if n_rows.dtype != torch.int32 or n_rows.numel() != 1:
            raise ValueError("n_rows must be a single int32 on the device")
    mode = DENOM_DEVICE if denom is not None else (DENOM_VALID if reduction_mean else DENOM_NONE)
    if N == 0:  # empty batch: nothing to launch.  torch: sum-reduction 0, mean-reduction nan (0/0)
        loss = torch.full((), float("nan") if reduction_mean else 0.0, dtype=torch.float32, device=x.device)
        out = (loss, torch.empty(0, dtype=torch.float32, device=x.device), torch.zeros((), dtype=torch.int64, device=x.device))
        return (*out, torch.zeros((), dtype=torch.int64, device=x.device),
                torch.zeros((), dtype=torch.float32, device=x.device)) if with_stats else out
    lib = _cabi.load()
    with torch.cuda.device(x.device):
        loss = torch.empty((), dtype=torch.float32, device=x.device)
        lse = torch.empty(N, dtype=torch.float32, device=x.device)
        n_valid = torch.empty((), dtype=torch.int64, device=x.device)
        ev = STATS.begin("flce_fwd", x.device)
        if with_stats:
            ws_bytes = lib.b200sft_flce_fwd_stats_workspace_bytes(N, H, V)
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
            n_correct = torch.empty((), dtype=torch.int64, device=x.device)
            entropy_sum = torch.empty((), dtype=torch.float32, device=x.device)
            check(lib.b200sft_flce_fwd_stats(_ptr(x), _ptr(w), _ptr(labels), N, H, V, ignore_index, float(scale), mode,
                                             _ptr(denom), flags, _ptr(n_rows), _ptr(loss), _ptr(lse), _ptr(n_valid),
                                             _ptr(n_correct), _ptr(entropy_sum), _ptr(ws), ws_bytes, _stream_ptr(x)),
                  "b200sft_flce_fwd_stats")
        else:
            ws_bytes = lib.b200sft_flce_fwd_workspace_bytes(N, H, V)
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
            check(lib.b200sft_flce_fwd(_ptr(x), _ptr(w), _ptr(labels), N, H, V, ignore_index, float(scale), mode,
                                       _ptr(denom), flags, _ptr(n_rows), _ptr(loss), _ptr(lse), _ptr(n_valid),
                                       _ptr(ws), ws_bytes, _stream_ptr(x)), "b200sft_flce_fwd")
        STATS.end(ev, x.device)
        STATS.kernel_launches += 3  # GEMM+softmax epilogue, row merge, loss reduce
    if with_stats:
        return loss, lse, n_valid, n_correct, entropy_sum
    return loss, lse, n_valid


def flce_backward(x: torch.Tensor, w: torch.Tensor, labels: torch.Tensor, lse: torch.Tensor,
                  grad_out: torch.Tensor, scale: float = 1.0, denom: Optional[torch.Tensor] = None,
                  ignore_index: int = IGNORE_INDEX, flags: int = 0, chunk_cols: int = 0,
                  n_rows: Optional[torch.Tensor] = None, need_dx: bool = True, need_dw: bool = True
                  ) -> Tuple[Optional[torch.Tensor], Optional[torch.Tensor]]:
    """dx [N,H] bf16, dw [V,H] bf16 for the loss produced by :func:`flce_forward` (with compacted
    inputs and ``n_rows``, dx is compact too: expand it with :func:`scatter_rows`).  ``need_dx`` /
    ``need_dw`` False skip the corresponding GEMM and return None in its place."""
    N, H, V = _check_flce_inputs(x, w, labels)
    _require_cuda(lse, grad_out)
    if lse.dtype != torch.float32 or lse.shape != (N,):
        raise ValueError("lse must be float32 [N]")
    grad_out = grad_out.to(torch.float32).reshape(()).contiguous()
    mode = DENOM_DEVICE if denom is not None else DENOM_NONE
    if N == 0:  # empty batch: no rows, zero weight gradient
        return (torch.empty_like(x) if need_dx else None), (torch.zeros_like(w) if need_dw else None)
    lib = _cabi.load()
    with torch.cuda.device(x.device):
        ws_bytes = lib.b200sft_flce_bwd_workspace_bytes(N, H, V, chunk_cols)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
        if not (need_dx or need_dw):
            return None, None
        dx = torch.empty_like(x) if need_dx else None
        dw = torch.empty_like(w) if need_dw else None
        ev = STATS.begin("flce_bwd", x.device)
        check(lib.b200sft_flce_bwd(_ptr(x), _ptr(w), _ptr(labels), _ptr(lse), _ptr(grad_out), N, H, V,
                                   ignore_index, float(scale), mode, _ptr(denom), flags, chunk_cols, _ptr(n_rows),
                                   _ptr(dx), _ptr(dw), _ptr(ws), ws_bytes, _stream_ptr(x)), "b200sft_flce_bwd")
        STATS.end(ev, x.device)
        STATS.kernel_launches += lib.b200sft_flce_bwd_launch_count(N, H, V, chunk_cols, int(need_dx), int(need_dw))
    return dx, dw


def compact_rows(x: torch.Tensor, labels: torch.Tensor, ignore_index: int = IGNORE_INDEX):
    """Gather the rows whose label is not ``ignore_index`` (stable order).  Returns
    ``(xc [N,H], labels_c [N], row_index [N], pos [N], n_live int32 scalar)`` -- all on the device;
    the live count is never read by the host."""
    _require_cuda(x, labels)
    if x.dtype != torch.bfloat16 or labels.dtype != torch.int32 or x.dim() != 2 or labels.shape != (x.shape[0],):
        raise ValueError("x must be bf16 [N,H] and labels int32 [N]")
    if not (x.is_contiguous() and labels.is_contiguous()):
        raise ValueError("x and labels must be contiguous")
    N, H = x.shape
    if N == 0:
        z = torch.empty(0, dtype=torch.int32, device=x.device)
        return x, labels, z, z.clone(), torch.zeros((), dtype=torch.int32, device=x.device)
    lib = _cabi.load()
    with torch.cuda.device(x.device):
        xc = torch.empty_like(x)
        labels_c = torch.empty_like(labels)
        row_index = torch.empty_like(labels)
        pos = torch.empty_like(labels)
        n_live = torch.empty((), dtype=torch.int32, device=x.device)
        check(lib.b200sft_compact_rows(_ptr(x), _ptr(labels), N, H, ignore_index, _ptr(xc), _ptr(labels_c),
                                       _ptr(row_index), _ptr(pos), _ptr(n_live), _stream_ptr(x)), "b200sft_compact_rows")
        STATS.kernel_launches += 2
    return xc, labels_c, row_index, pos, n_live


def scatter_rows(dxc: torch.Tensor, pos: torch.Tensor) -> torch.Tensor:
    """dx[r] = dxc[pos[r]] if pos[r] >= 0 else 0 (inverse of :func:`compact_rows` for gradients)."""
    _require_cuda(dxc, pos)
    N, H = dxc.shape
    if N == 0:
        return dxc
    lib = _cabi.load()
    with torch.cuda.device(dxc.device):
        dx = torch.empty_like(dxc)
        check(lib.b200sft_scatter_rows(_ptr(dxc), _ptr(pos), N, H, _ptr(dx), _stream_ptr(dxc)), "b200sft_scatter_rows")
        STATS.kernel_launches += 1
    return dx


def gemm_bf16(a: torch.Tensor, b: torch.Tensor, M: int, N: int, K: int, a_mn_major: bool = False,
              b_mn_major: bool = False) -> torch.Tensor:
    """C[M,N] = A . B^T on the tcgen05 engine.  A is [M,K] (or [K,M] when ``a_mn_major``), B is
    [N,K] (or [K,N] when ``b_mn_major``).  Test hook for the engine's operand layouts."""
    _require_cuda(a, b)
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16 or not a.is_contiguous() or not b.is_contiguous():
        raise ValueError("a and b must be contiguous bfloat16")
    exp_a = (K, M) if a_mn_major else (M, K)
    exp_b = (K, N) if b_mn_major else (N, K)
    if tuple(a.shape) != exp_a or tuple(b.shape) != exp_b:
        raise ValueError(f"expected a {exp_a}, b {exp_b}; got {tuple(a.shape)}, {tuple(b.shape)}")
    ldc = (N + 7) // 8 * 8
    lib = _cabi.load()
    with torch.cuda.device(a.device):
        c = torch.empty(M, ldc, dtype=torch.bfloat16, device=a.device)
        check(lib.b200sft_gemm_bf16(_ptr(a), _ptr(b), _ptr(c), M, N, K, a.shape[1], b.shape[1], ldc,
                                    int(a_mn_major), int(b_mn_major), _stream_ptr(a)), "b200sft_gemm_bf16")
    return c[:, :N]


# --------------------------------------------------------------------------------------------
# gradient-bucket scale / cast
# --------------------------------------------------------------------------------------------
def scale_(buf: torch.Tensor, scale: float) -> torch.Tensor:
    """In-place ``buf *= scale`` for a contiguous bf16 or fp32 bucket (default_hooks.py:18-32's div_)."""
    _require_cuda(buf)
    if not buf.is_contiguous():
        raise ValueError("bucket must be contiguous")
    lib = _cabi.load()
    with torch.cuda.device(buf.device):
        if buf.dtype == torch.bfloat16:
            check(lib.b200sft_scale_bf16_inplace(_ptr(buf), buf.numel(), float(scale), _stream_ptr(buf)),
                  "b200sft_scale_bf16_inplace")
        elif buf.dtype == torch.float32:
            check(lib.b200sft_scale_f32_inplace(_ptr(buf), buf.numel(), float(scale), _stream_ptr(buf)),
                  "b200sft_scale_f32_inplace")
        else:
            raise ValueError(f"unsupported bucket dtype {buf.dtype}")
        STATS.kernel_launches += 1 if buf.numel() > 0 else 0
    return buf


def scale_cast_f32_to_bf16(src: torch.Tensor, scale: float, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """bf16(src * scale): the compress half of bf16_compress_hook (default_hooks.py:57-93)."""
    _require_cuda(src)
    if src.dtype != torch.float32 or not src.is_contiguous():
        raise ValueError("src must be contiguous float32")
    lib = _cabi.load()
    with torch.cuda.device(src.device):
        if out is None:
            out = torch.empty(src.shape, dtype=torch.bfloat16, device=src.device)
        check(lib.b200sft_scale_cast_f32_to_bf16(_ptr(src), _ptr(out), src.numel(), float(scale), _stream_ptr(src)),
              "b200sft_scale_cast_f32_to_bf16")
        STATS.kernel_launches += 1 if src.numel() > 0 else 0
    return out


def cast_bf16_to_f32(src: torch.Tensor, out: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
    """out = fp32(src) * scale: the decompress half of bf16_compress_hook."""
    _require_cuda(src, out)
    if src.dtype != torch.bfloat16 or out.dtype != torch.float32 or src.numel() != out.numel():
        raise ValueError("src bf16 / out fp32 of equal size expected")
    lib = _cabi.load()
    with torch.cuda.device(src.device):
        check(lib.b200sft_cast_bf16_to_f32(_ptr(src), _ptr(out), src.numel(), float(scale), _stream_ptr(src)),
              "b200sft_cast_bf16_to_f32")
        STATS.kernel_launches += 1 if src.numel() > 0 else 0
    return out


# --------------------------------------------------------------------------------------------
# torch custom ops + autograd
# --------------------------------------------------------------------------------------------
@torch.library.custom_op("b200sft::flce_fwd", mutates_args=())
def _flce_fwd_op(x: torch.Tensor, w: torch.Tensor, labels: torch.Tensor, scale: float,
                 denom: Optional[torch.Tensor], reduction_mean: bool, ignore_index: int, flags: int,
                 n_rows: Optional[torch.Tensor], with_stats: bool
                 ) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor, torch.Tensor]:
    out = flce_forward(x, w, labels, scale, denom, reduction_mean, ignore_index, flags, n_rows, with_stats)
    if with_stats:
        return out
    return (*out, x.new_zeros((), dtype=torch.int64), x.new_zeros((), dtype=torch.float32))


@_flce_fwd_op.register_fake
def _(x, w, labels, scale, denom, reduction_mean, ignore_index, flags, n_rows, with_stats):
    return (x.new_empty((), dtype=torch.float32), x.new_empty((x.shape[0],), dtype=torch.float32),
            x.new_empty((), dtype=torch.int64), x.new_empty((), dtype=torch.int64),
            x.new_empty((), dtype=torch.float32))


@torch.library.custom_op("b200sft::flce_bwd", mutates_args=())
def _flce_bwd_op(x: torch.Tensor, w: torch.Tensor, labels: torch.Tensor, lse: torch.Tensor,
                 grad_out: torch.Tensor, scale: float, denom: Optional[torch.Tensor], ignore_index: int,
                 flags: int, chunk_cols: int, n_rows: Optional[torch.Tensor], need_dx: bool, need_dw: bool
                 ) -> Tuple[torch.Tensor, torch.Tensor]:
    dx, dw = flce_backward(x, w, labels, lse, grad_out, scale, denom, ignore_index, flags, chunk_cols, n_rows,
                           need_dx, need_dw)
    # custom ops return tensors: an empty tensor stands for "not computed"
    return (dx if dx is not None else x.new_empty(0)), (dw if dw is not None else w.new_empty(0))


@_flce_bwd_op.register_fake
def _(x, w, labels, lse, grad_out, scale, denom, ignore_index, flags, chunk_cols, n_rows, need_dx, need_dw):
    return (torch.empty_like(x) if need_dx else x.new_empty(0)), (torch.empty_like(w) if need_dw else w.new_empty(0))


class _FusedLinearCrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, labels, scale, denom, reduction_mean, ignore_index, flags, chunk_cols, skip_ignored,
                with_stats, plan, norm_weight, norm_eps):
        pos = n_rows = rstd = None
        xk, lk = x, labels
        listed = plan is not None and plan.row_index is not None
        if norm_weight is not None:  # x holds PRE-norm rows: the final RMSNorm rides on the gather
            if skip_ignored and not listed:
                plan, listed = compaction_list(labels, ignore_index), True
            if not skip_ignored:
                plan, listed = None, False
            xk, rstd = rmsnorm_gather(x, norm_weight, norm_eps, plan)
            if listed:
                lk, pos, n_rows = plan.labels_c, plan.pos, plan.n_live
        elif skip_ignored and listed:
            xk, lk, pos, n_rows = gather_rows(x, plan), plan.labels_c, plan.pos, plan.n_live  # list already made
        elif skip_ignored:  # run every GEMM on the live rows only
            xk, lk, _row_index, pos, n_rows = compact_rows(x, labels, ignore_index)
        loss, lse, n_valid, n_correct, entropy_sum = torch.ops.b200sft.flce_fwd(
            xk, w, lk, scale, denom, reduction_mean, ignore_index, flags, n_rows, with_stats)
        # reduction="mean" divides by the local valid count; backward takes it as a device scalar
        bwd_denom = n_valid if reduction_mean else denom
        fused_norm = norm_weight is not None  # then also: pre-norm rows, weight, 1/rms and the list, for its backward
        ctx.save_for_backward(xk, w, lk, lse, bwd_denom, pos, n_rows, x if fused_norm else None, norm_weight, rstd,
                              plan.row_index if (fused_norm and listed) else None)
        ctx.scale, ctx.ignore_index, ctx.flags, ctx.chunk_cols = scale, ignore_index, flags, chunk_cols
        ctx.fused_norm = fused_norm
        ctx.mark_non_differentiable(n_valid, n_correct, entropy_sum)
        return loss, n_valid, n_correct, entropy_sum

    @staticmethod
    def backward(ctx, grad_loss, *_unused):
        xk, w, lk, lse, bwd_denom, pos, n_rows, h, gamma, rstd, row_index = ctx.saved_tensors
        need_dx, need_dw = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        need_dgamma = ctx.fused_norm and ctx.needs_input_grad[12]
        dx, dw = torch.ops.b200sft.flce_bwd(xk, w, lk, lse, grad_loss, ctx.scale, bwd_denom, ctx.ignore_index,
                                            ctx.flags, ctx.chunk_cols, n_rows, need_dx or need_dgamma, need_dw)
        dgamma = None
        if ctx.fused_norm:
            plan = LabelPlan(None, None, None, row_index, lk, pos, n_rows) if row_index is not None else None
            dx, dgamma = rmsnorm_scatter_bwd(dx, h, gamma, rstd, plan, need_dx, need_dgamma)
        else:
            dx = None if not need_dx else (scatter_rows(dx, pos) if pos is not None else dx)
        return (dx, (dw if need_dw else None), None, None, None, None, None, None, None, None, None, None,
                dgamma, None)


def fused_linear_cross_entropy(hidden: torch.Tensor, weight: torch.Tensor, shifted_labels: torch.Tensor,
                               num_items_in_batch: Optional[torch.Tensor] = None, scale: float = 1.0,
                               ignore_index: int = IGNORE_INDEX, flags: int = 0, chunk_cols: int = 0,
                               skip_ignored_rows: bool = False, return_stats: bool = False,
                               plan: Optional[LabelPlan] = None, norm_weight: Optional[torch.Tensor] = None,
                               norm_eps: float = 1e-6):
    """Differentiable ``cross_entropy(hidden @ weight.T, shifted_labels)`` without the logits.

    ``num_items_in_batch`` None -> mean over local valid rows (``reduction="mean"``), else the
    sum divided by that device scalar -- SP/transformers/loss/loss_utils.py:28-42.
    ``hidden`` is [N,H] bf16, ``weight`` [V,H] bf16, ``shifted_labels`` int32 [N] from
    :func:`label_prep`.  ``skip_ignored_rows`` compacts away rows labelled ``ignore_index`` first
    (same result; less work when many rows are masked).  ``return_stats`` -> ``(loss, stats)`` with
    ``stats = {"n_valid", "n_correct", "entropy_sum"}`` device scalars (token accuracy and mean
    entropy are n_correct / n_valid and entropy_sum / n_valid).  ``plan``: the :class:`LabelPlan`
    ``shifted_labels`` came from; its compaction list is then reused instead of being rebuilt.
    ``norm_weight`` (bf16 [H]) / ``norm_eps``: ``hidden`` holds the rows BEFORE the model's final RMSNorm
    (modeling_smollm3.py:438); the norm is applied to the scored rows while they are gathered, and its
    backward (dX and the weight's gradient) is part of this op's backward.
    """
    denom = None
    if num_items_in_batch is not None:
        if not torch.is_tensor(num_items_in_batch):
            num_items_in_batch = torch.tensor(int(num_items_in_batch), dtype=torch.int64, device=hidden.device)
        denom = num_items_in_batch.to(device=hidden.device, dtype=torch.int64).reshape(())
    loss, n_valid, n_correct, entropy_sum = _FusedLinearCrossEntropy.apply(
        hidden, weight, shifted_labels, float(scale), denom, num_items_in_batch is None, ignore_index, flags,
        chunk_cols, bool(skip_ignored_rows), bool(return_stats), plan, norm_weight, float(norm_eps))
    if return_stats:
        return loss, {"n_valid": n_valid, "n_correct": n_correct, "entropy_sum": entropy_sum}
    return loss
