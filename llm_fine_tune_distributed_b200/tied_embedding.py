"""Early reduction of the tied ``lm_head`` / ``embed_tokens`` gradient under DDP (second half).

``lm_head.weight`` is ``embed_tokens.weight`` (SP/transformers/models/smollm3/modeling_smollm3.py:367,447) and the
reference trains it (training.py:136-141).  Autograd completes that parameter's gradient only when the embedding's
backward -- the last op of ``backward()`` -- has added its rows, so stock DDP (training.py:249-256) cannot start
reducing its 521 MiB bucket (62 % of the step's gradient bytes) until all backward compute is over.

Averaging over ranks is linear, so on a *synchronising* backward this module splits the gradient in two:

* the embedding output is handed to the decoder as a leaf (the embedding op is left out of the autograd graph), so the
  tied weight's gradient is the head's dW alone and is complete at the START of backward; DDP's reducer marks it
  ready then and the ordinary bucket hook (``comm_hook.py``) averages it -- together with whatever earlier
  ``no_sync`` micro-batches accumulated -- while the 36 decoder layers are still running their backward;
* the gradient w.r.t. the embedding output (dY, ``[tokens, H]``) is captured when backward reaches it, published with
  the token ids in symmetric memory, and after DDP has finalised the backward every rank adds
  ``grad[v] += (1/world) * sum_{ranks, positions with id v} dY`` for the touched rows with one kernel that reads the
  peers' rows over NVLink (``b200sft_embed_grad_rows``, csrc/embed_grad.cu) -- at most world x tokens rows of 4 KB
  instead of 525 MB, in one fixed summation order so the replicas stay bit-identical.

Micro-batches under ``no_sync`` (gradient accumulation) and everything outside training keep the stock graph.
The result equals stock DDP's average up to bf16 rounding (the sum is split differently); tests state the error.
"""
from __future__ import annotations

import ctypes
from typing import Callable, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from . import _cabi

__all__ = ["TiedEmbeddingExchange", "SymmetricRowTransport"]


class SymmetricRowTransport:
    """Symmetric-memory arena ``[n | ids | dY]`` per rank plus the kernel's scratch, allocated once (a collective:
    construct on every rank at the same point).  ``publish`` stages this rank's rows, ``apply`` runs the exchange."""

    def __init__(self, cap: int, hidden: int, vocab: int, device: torch.device, group: Optional[dist.ProcessGroup] = None):
        import torch.distributed._symmetric_memory as symm_mem

        self.group = group if group is not None else dist.group.WORLD
        self.world, self.rank = self.group.size(), self.group.rank()
        self.cap, self.hidden, self.vocab, self.device = int(cap), int(hidden), int(vocab), device
        self.ids_off = 256
        self.dy_off = self.ids_off + (self.cap * 4 + 255) // 256 * 256
        nbytes = self.dy_off + self.cap * self.hidden * 2
        with torch.cuda.device(device):
            self.arena = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
            self.handle = symm_mem.rendezvous(self.arena, self.group)
            lib = _cabi.load()
            self.ws_bytes = lib.b200sft_embed_grad_workspace_bytes(self.world, self.cap, self.vocab, self.hidden)
            self.workspace = torch.empty(self.ws_bytes, dtype=torch.uint8, device=device)
        self.n_view = self.arena[:4].view(torch.int32)
        self.ids_view = self.arena[self.ids_off:self.ids_off + self.cap * 4].view(torch.int32)
        self.dy_view = self.arena[self.dy_off:].view(torch.bfloat16).view(self.cap, self.hidden)
        peers = [int(p) for p in self.handle.buffer_ptrs]
        self._n_ptrs = (ctypes.c_uint64 * self.world)(*peers)
        self._ids_ptrs = (ctypes.c_uint64 * self.world)(*[p + self.ids_off for p in peers])
        self._dy_ptrs = (ctypes.c_uint64 * self.world)(*[p + self.dy_off for p in peers])

    def publish(self, ids32: torch.Tensor, dy: torch.Tensor) -> None:
        n = ids32.numel()
        self.ids_view[:n].copy_(ids32)
        self.dy_view[:n].copy_(dy.reshape(n, self.hidden))
        self.n_view.fill_(n)

    def apply(self, grad: torch.Tensor, scale: float, mark: Optional[Callable[[], None]] = None) -> None:
        """``mark`` (measurement aid) is called after the first barrier and after the kernels."""
        from . import ops

        stream = torch.cuda.current_stream(self.device)
        with torch.cuda.device(self.device):
            self.handle.barrier(channel=1)            # every rank has published its rows
            if mark is not None:
                mark()
            _cabi.check(_cabi.load().b200sft_embed_grad_rows(self._ids_ptrs, self._dy_ptrs, self._n_ptrs, self.world,
                                                             self.cap, self.vocab, self.hidden, float(scale),
                                                             grad.data_ptr(), self.workspace.data_ptr(), self.ws_bytes,
                                                             stream.cuda_stream), "b200sft_embed_grad_rows")
            ops.STATS.kernel_launches += 6
            if mark is not None:
                mark()
            self.handle.barrier(channel=1)            # every rank has read them: the arena may be rewritten


class TiedEmbeddingExchange:
    """Installs the split described in the module docstring on a DDP-wrapped causal LM whose input embedding and
    output head share one weight.  ``max_tokens`` bounds the tokens of one micro-batch on any rank (per-device batch x
    sequence length).  ``transport`` is the row exchange (default: :class:`SymmetricRowTransport`; the CPU tests pass a
    gloo stand-in, there being no CPU kernel)."""

    def __init__(self, ddp_model: nn.parallel.DistributedDataParallel, max_tokens: int,
                 transport: Optional[object] = None, transport_factory: Optional[Callable] = None):
        inner = ddp_model.module
        emb = inner.get_input_embeddings()
        head = getattr(inner, "lm_head", None)
        if emb is None or head is None or emb.weight is not head.weight:
            raise ValueError("the tied-embedding exchange needs lm_head.weight to BE embed_tokens.weight")
        self.ddp, self.emb, self.weight = ddp_model, emb, emb.weight
        self.padding_idx = emb.padding_idx
        self.group = ddp_model.process_group
        self.world = self.group.size()
        self.max_tokens = int(max_tokens)
        vocab, hidden = self.weight.shape
        if transport is None:
            make = transport_factory or SymmetricRowTransport
            transport = make(self.max_tokens, hidden, vocab, self.weight.device, self.group)
        self.transport = transport
        self.enabled = True
        self.exchanges = 0            # synchronising backwards that went through the row exchange
        self._attention_mask = None   # set per forward by the loss seam (padding positions carry no gradient)
        self._pending = None
        # measurement aid: when a list, every exchange appends six CUDA events recorded on the current stream: dY captured
        # (= backward compute is over), rows published, DDP finalised, peers arrived, rows added, all ranks done
        self.events = None
        self._marks = None
        self._handle = emb.register_forward_hook(self._forward_hook)

    def remove(self) -> None:
        self._handle.remove()

    def set_attention_mask(self, mask: Optional[torch.Tensor]) -> None:
        self._attention_mask = mask

    # ---- forward: cut the embedding out of the graph on synchronising micro-batches ----------------------
    def _active(self, output: torch.Tensor) -> bool:
        return (self.enabled and torch.is_grad_enabled() and output.requires_grad and self.emb.training
                and self.ddp.require_backward_grad_sync and self.weight.requires_grad)

    def _forward_hook(self, module, args, output):
        mask, self._attention_mask = self._attention_mask, None
        if not self._active(output):
            return output
        ids = args[0]
        if ids.numel() > self.max_tokens:
            raise RuntimeError(f"micro-batch of {ids.numel()} tokens exceeds the exchange's capacity ({self.max_tokens})")
        ids32 = ids.reshape(-1).to(torch.int32)
        skip = None
        if self.padding_idx is not None:        # nn.Embedding gives its padding_idx row no gradient
            skip = ids32 == int(self.padding_idx)
        if mask is not None and mask.numel() == ids32.numel():   # padded positions: their dY is exactly zero
            pad = mask.reshape(-1) == 0
            skip = pad if skip is None else (skip | pad)
        if skip is not None:
            ids32 = torch.where(skip, torch.full_like(ids32, -1), ids32)
        leaf = output.detach().requires_grad_(True)
        leaf.register_hook(lambda g, ids32=ids32, leaf=leaf: self._capture(ids32, g, leaf))
        return leaf

    # ---- backward: publish the rows when they exist, apply them after DDP has finalised -------------------
    def _capture(self, ids32: torch.Tensor, grad_out: torch.Tensor, leaf: torch.Tensor) -> None:
        dy = grad_out if grad_out.dtype == torch.bfloat16 else grad_out.to(torch.bfloat16)
        self._marks = [self._mark()] if self.events is not None and dy.is_cuda else None
        self.transport.publish(ids32, dy.contiguous())
        if self._marks is not None:
            self._marks.append(self._mark())
        self._pending = leaf
        torch.autograd.Variable._execution_engine.queue_callback(self._requeue)
        return None

    def _requeue(self) -> None:
        # End-of-backward callbacks run in queueing order, and DDP's reducer queues the one that copies the reduced
        # buckets back into .grad when its LAST bucket becomes ready -- which, when the first decoder layer trains
        # too, can be after this hook ran.  By the time any callback runs every node of the backward has executed, so
        # DDP's is in the list: queueing the real work now puts it behind DDP's in every case.
        torch.autograd.Variable._execution_engine.queue_callback(self._finish)

    def _finish(self) -> None:
        leaf, self._pending = self._pending, None
        if leaf is not None:
            leaf.grad = None                     # the captured rows live in the transport; drop autograd's copy
        grad = self.weight.grad
        if grad is None:
            raise RuntimeError("tied weight has no gradient after backward: was the head's backward skipped?")
        if self._marks is not None:   # (captured, published, DDP finalised, peers arrived, rows added, all ranks done)
            self._marks.append(self._mark())
            self.transport.apply(grad, 1.0 / self.world, mark=lambda: self._marks.append(self._mark()))
            self.events.append((*self._marks, self._mark()))
            self._marks = None
        else:
            self.transport.apply(grad, 1.0 / self.world)
        self.exchanges += 1

    @staticmethod
    def _mark() -> "torch.cuda.Event":
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        return ev
