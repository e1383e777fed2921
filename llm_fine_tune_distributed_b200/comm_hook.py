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

__all__ = ["B200HookState", "B200NvlinkState", "b200_allreduce_hook", "b200_bf16_compress_hook",
           "b200_nvlink_allreduce_hook", "register_b200_comm_hook"]


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
        # measurement aid: when `timeline` is a list, every hook call appends
        # (event recorded on the autograd stream when the bucket was handed over, event recorded on the side
        # stream when its reduction finished, bucket bytes)
        self.timeline = None

    def group(self) -> dist.ProcessGroup:
        return self.process_group if self.process_group is not None else dist.group.WORLD

    def side_stream(self, device: torch.device) -> "torch.cuda.Stream":
        s = self._streams.get(device)
        if s is None:
            s = torch.cuda.Stream(device=device, priority=-1)
            self._streams[device] = s
        return s


def _mark_ready(state: B200HookState, buf: torch.Tensor):
    if state.timeline is None or not buf.is_cuda:
        return None
    ev = torch.cuda.Event(enable_timing=True)
    ev.record(torch.cuda.current_stream(buf.device))
    return ev


def _mark_done(state: B200HookState, buf: torch.Tensor, ready, stream) -> None:
    if ready is None:
        return
    ev = torch.cuda.Event(enable_timing=True)
    ev.record(stream)
    state.timeline.append((ready, ev, buf.numel() * buf.element_size()))


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
        # default_hooks.py:25 divides (tensor.div_(world_size)); on CUDA that division by a host scalar IS a multiply by
        # the fp32 reciprocal (ATen BinaryDivTrueKernel), so scaling by fp32(1/world) gives the reference's bits for
        # every world size, powers of two or not (tests/test_gpu_allreduce_kernel_one_gpu.py, worlds 2..8)
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


class B200NvlinkState(B200HookState):
    """State for :func:`b200_nvlink_allreduce_hook`: ONE symmetric-memory arena
    (``torch.distributed._symmetric_memory``: every rank maps every peer's copy, plus an NVLS
    multicast address when the fabric supports it) sized for all gradients of the model, allocated
    and exchanged here -- a collective, so construct the state on every rank at the same point,
    before training.  Buckets are carved from the arena in arrival order during each backward, so
    the hook itself performs no allocation and no host-side collective."""

    def __init__(self, total_grad_elements: int, device: torch.device,
                 process_group: Optional[dist.ProcessGroup] = None, use_multimem: Optional[bool] = None,
                 max_buckets: int = 512):
        super().__init__(process_group)
        import torch.distributed._symmetric_memory as symm_mem

        world = self.group().size()
        self.align = 8 * world                       # elements: 16-byte vectors, one slice per rank
        self.capacity = (int(total_grad_elements) + max_buckets * self.align + self.align - 1) // self.align * self.align
        with torch.cuda.device(device):
            self.arena = symm_mem.empty(self.capacity, dtype=torch.bfloat16, device=device)
            self.handle = symm_mem.rendezvous(self.arena, self.group())
        # measured on 8xB200 (profiles/r01_allreduce_microbench.jsonl): the in-switch (NVLS) reduction wins from
        # 4 ranks up; with 2 ranks plain peer loads are faster (each rank pulls half a bucket over one hop)
        auto = bool(self.handle.multicast_ptr) and world > 2
        self.use_multimem = auto if use_multimem is None else bool(use_multimem)
        self.path_used = "multimem" if self.use_multimem else "p2p"
        self._offset = 0
        self._seen = set()
        self.trace = None  # set to a list to record (bucket index, numel, offset) per hook call (debugging)
        # with `timeline` on: per hook call (ready, side-stream-free, staged, peers-arrived, kernel-done, done, bytes)
        self.phases = []

    @classmethod
    def for_model(cls, model: torch.nn.Module, process_group: Optional[dist.ProcessGroup] = None, **kw):
        params = [p for p in model.parameters() if p.requires_grad]
        return cls(sum(p.numel() for p in params), params[0].device, process_group, **kw)

    def carve(self, bucket_index: int, numel: int):
        """(element offset, padded length) of this bucket inside the arena for the current backward."""
        if bucket_index in self._seen:  # an index repeats: a new backward pass has started
            self._seen.clear()
            self._offset = 0
        self._seen.add(bucket_index)
        padded = (numel + self.align - 1) // self.align * self.align
        off = self._offset
        if off + padded > self.capacity:
            raise RuntimeError(f"symmetric arena too small: need {off + padded} elements, have {self.capacity}")
        self._offset = off + padded
        if self.trace is not None:
            self.trace.append((bucket_index, numel, off))
        return off, padded


class _DevicePtr:
    """Minimal __cuda_array_interface__ carrier: lets torch alias `nbytes` of device memory at `ptr`
    as a tensor whose storage starts exactly there."""

    def __init__(self, ptr: int, n_int16: int):
        self.__cuda_array_interface__ = {"shape": (n_int16,), "typestr": "<i2", "data": (int(ptr), False),
                                         "version": 3, "strides": None}


def _alias_bf16(ptr: int, numel: int, device: torch.device) -> torch.Tensor:
    # DDP builds its per-parameter views of a hook's result with as_strided(sizes, strides, offset)
    # where `offset` is relative to the bucket start, i.e. it assumes storage_offset == 0
    # (Reducer::initialize_bucket_views).  A plain slice of the arena would therefore be read from
    # the arena's start; alias the slice as a tensor with its own zero-offset storage instead.
    return torch.as_tensor(_DevicePtr(ptr, numel), device=device).view(torch.bfloat16)


def b200_nvlink_allreduce_hook(state: B200NvlinkState, bucket: dist.GradBucket) -> torch.futures.Future[torch.Tensor]:
    """Average a bf16 gradient bucket with this library's own NVLink kernel: copy into the symmetric
    arena, then ONE kernel reduces across ranks (in the NVSwitch via multimem when available, else
    by peer loads), applies 1/world in registers and broadcasts the result to every rank.  The
    future's value is the arena slice; DDP copies it back into the gradients.  Non-bf16 or CPU
    buckets take the NCCL hook."""
    buf = bucket.buffer()
    if not buf.is_cuda or buf.dtype != torch.bfloat16:
        return b200_allreduce_hook(state, bucket)
    import ctypes

    from . import _cabi

    group = state.group()
    world, rank, n = group.size(), group.rank(), buf.numel()
    state.buckets_seen += 1
    state.bytes_reduced += n * 2
    off, padded = state.carve(bucket.index(), n)
    ready = _mark_ready(state, buf)
    hdl = state.handle
    view = _alias_bf16(state.arena.data_ptr() + 2 * off, n, buf.device)  # same memory as arena[off:off+n]
    lib = _cabi.load()
    cur = torch.cuda.current_stream(buf.device)
    side = state.side_stream(buf.device)
    side.wait_stream(cur)
    buf.record_stream(side)
    marks = [] if ready is not None else None       # measurement aid: events between the phases, on the side stream

    def mark():
        if marks is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(side)
            marks.append(ev)

    with torch.cuda.device(buf.device), torch.cuda.stream(side):
        mark()                                      # side stream free (earlier buckets done) and the bucket is ready
        view.copy_(buf)
        mark()                                      # staged in the arena
        hdl.barrier(channel=0)                      # every rank's copy is in place
        mark()                                      # ... = how long this rank waited for its peers
        if state.use_multimem:
            _cabi.check(lib.b200sft_allreduce_scale_bf16_multimem(int(hdl.multicast_ptr) + 2 * off, padded, rank, world,
                                                                  1.0 / world, side.cuda_stream),
                        "b200sft_allreduce_scale_bf16_multimem")
        else:
            ptrs = (ctypes.c_uint64 * world)(*[int(p) + 2 * off for p in hdl.buffer_ptrs])
            _cabi.check(lib.b200sft_allreduce_scale_bf16_p2p(ptrs, padded, rank, world, 1.0 / world, side.cuda_stream),
                        "b200sft_allreduce_scale_bf16_p2p")
        ops.STATS.kernel_launches += 1
        mark()                                      # reduce + scale + broadcast kernel
        hdl.barrier(channel=0)                      # every rank's slice of the result has landed everywhere
        _mark_done(state, buf, ready, side)
        if marks is not None:
            state.phases.append((ready, *marks, state.timeline[-1][1], n * 2))
        fut = torch.futures.Future(devices=[buf.device])
        fut.set_result(view)
    return fut


def register_b200_comm_hook(ddp_model: torch.nn.parallel.DistributedDataParallel,
                            state: Optional[B200HookState] = None, compress_fp32: bool = False,
                            nvlink_kernel: bool = False) -> B200HookState:
    """``ddp_model.register_comm_hook(state, hook)`` with this library's hook; returns the state.
    ``nvlink_kernel`` selects the fused NVLink reduce+scale kernel instead of scale kernel + NCCL."""
    pg = getattr(ddp_model, "process_group", None)
    if nvlink_kernel:
        state = state or B200NvlinkState.for_model(ddp_model.module, pg)
        ddp_model.register_comm_hook(state, b200_nvlink_allreduce_hook)
        return state
    state = state or B200HookState(pg)
    ddp_model.register_comm_hook(state, b200_bf16_compress_hook if compress_fp32 else b200_allreduce_hook)
    return state
