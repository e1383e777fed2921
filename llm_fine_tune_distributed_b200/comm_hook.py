"""Seam #2: DDP gradient communication hook for one 8xB200 NVSwitch box.

Replaces DDP's built-in bucket allreduce / ``default_hooks.allreduce_hook``
(SP/torch/distributed/algorithms/ddp_comm_hooks/default_hooks.py:18-54), which the reference gets
from ``DistributedDataParallel(bucket_cap_mb=50, find_unused_parameters=False)``
(training.py:249-256), and ``bf16_compress_hook`` (default_hooks.py:57-93,116-134) for fp32
gradients.  Contract (SP/torch/nn/parallel/distributed.py:1987-2067): called on the autograd thread
once per full bucket of a synchronising backward, receives the flat *unscaled* bucket, returns a
Future of a tensor of the same shape.

What changes: the 1/world scale (and the fp32<->bf16 casts of the compress variant) run in this
library's vectorised kernels on a dedicated side stream, fenced against the autograd stream with
events, so neither the scale nor the NCCL launch sits on the backward stream; NCCL then reduces
over NVLink 5 / NVSwitch (NVLS when available -- none of the reference's socket-only NCCL_*
overrides, deploy/pytorchjob.yaml:53-64, are set).  The collective itself is ``torch.distributed``
plumbing (ProcessGroupNCCL).
"""
# NB: no `from __future__ import annotations` here -- DDP checks the hook's annotations as objects
# (bucket: dist.GradBucket, return: torch.futures.Future[torch.Tensor]), distributed.py:2262-2290.
from typing import Callable, Optional

import torch
import torch.distributed as dist

from . import ops

__all__ = ["B200HookState", "b200_allreduce_hook", "b200_bf16_compress_hook", "register_b200_comm_hook"]


class B200HookState:
    """State object handed to ``register_comm_hook``.

    ``scale_fn(buf, s)`` scales a bucket in place; it defaults to the CUDA kernel and exists so that
    the CPU/gloo plumbing tests can exercise the hook's control flow (there is no CPU kernel).
    """

    def __init__(self, process_group: Optional[dist.ProcessGroup] = None,
                 scale_fn: Optional[Callable[[torch.Tensor, float], torch.Tensor]] = None,
                 use_side_stream: bool = True):
        self.process_group = process_group
        self.scale_fn = scale_fn or ops.scale_
        self.use_side_stream = use_side_stream
        self._streams = {}
        self.buckets_seen = 0
        self.bytes_reduced = 0

    def group(self) -> dist.ProcessGroup:
        return self.process_group if self.process_group is not None else dist.group.WORLD

    def side_stream(self, device: torch.device) -> "torch.cuda.Stream":
        s = self._streams.get(device)
        if s is None:
            s = torch.cuda.Stream(device=device, priority=-1)
            self._streams[device] = s
        return s


def _launch(state: B200HookState, buf: torch.Tensor, body: Callable[[], "dist.Work"]) -> "dist.Work":
    """Run ``body`` (kernels + async collective) on the side stream, ordered after everything the
    autograd stream has written into the bucket so far."""
    if buf.is_cuda and state.use_side_stream:
        cur = torch.cuda.current_stream(buf.device)
        side = state.side_stream(buf.device)
        side.wait_stream(cur)
        buf.record_stream(side)
        with torch.cuda.stream(side):
            return body()
    return body()


def b200_allreduce_hook(state: B200HookState, bucket: dist.GradBucket) -> torch.futures.Future[torch.Tensor]:
    """Average a gradient bucket across ranks: scale by 1/world in place, then allreduce(SUM)."""
    group = state.group()
    buf = bucket.buffer()
    world = group.size()
    state.buckets_seen += 1
    state.bytes_reduced += buf.numel() * buf.element_size()

    def body():
        state.scale_fn(buf, 1.0 / world)
        return dist.all_reduce(buf, group=group, async_op=True)

    work = _launch(state, buf, body)
    return work.get_future().then(lambda fut: fut.value()[0])


def b200_bf16_compress_hook(state: B200HookState, bucket: dist.GradBucket) -> torch.futures.Future[torch.Tensor]:
    """fp32 bucket -> bf16 (scaled) -> allreduce -> fp32 in the original buffer; bf16 buckets take
    the plain path.  Halves NVLink bytes for fp32-gradient training."""
    buf = bucket.buffer()
    if buf.dtype != torch.float32 or not buf.is_cuda:
        return b200_allreduce_hook(state, bucket)
    group = state.group()
    world = group.size()
    state.buckets_seen += 1
    state.bytes_reduced += buf.numel() * 2
    holder = {}

    def body():
        holder["c"] = ops.scale_cast_f32_to_bf16(buf, 1.0 / world)
        return dist.all_reduce(holder["c"], group=group, async_op=True)

    work = _launch(state, buf, body)

    def decompress(fut):
        comp = fut.value()[0]
        ops.cast_bf16_to_f32(comp, buf)  # runs on the stream the future's callback is bound to
        return buf

    return work.get_future().then(decompress)


def register_b200_comm_hook(ddp_model: torch.nn.parallel.DistributedDataParallel,
                            state: Optional[B200HookState] = None, compress_fp32: bool = False) -> B200HookState:
    """``ddp_model.register_comm_hook(state, hook)`` with this library's hook; returns the state."""
    state = state or B200HookState(getattr(ddp_model, "process_group", None))
    ddp_model.register_comm_hook(state, b200_bf16_compress_hook if compress_fp32 else b200_allreduce_hook)
    return state
