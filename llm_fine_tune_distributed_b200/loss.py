"""Seam #1: the ``Trainer.compute_loss`` override that routes the LM head + cross-entropy of a HF
causal-LM through the fused sm_100a kernels.

The reference trains with ``trl.SFTTrainer`` (training.py:289-300); its loss is produced inside
``SmolLM3ForCausalLM.forward`` (SP/transformers/models/smollm3/modeling_smollm3.py:501-508):
``logits = self.lm_head(hidden)`` then ``self.loss_function(logits, labels, vocab_size, **kwargs)``.
Both calls are attribute look-ups on the model, so the drop-in is done without touching
transformers:

* ``model.lm_head`` becomes a :class:`FusedLMHead` that owns the *same* weight Parameter under the
  same name (state-dict key ``lm_head.weight`` and the tie to ``embed_tokens.weight`` are
  unchanged) and, while the fused path is armed, returns the hidden states untouched;
* ``model.loss_function`` becomes :func:`fused_causal_lm_loss`, which receives those hidden states
  in place of logits and calls the fused forward/backward.

The model is still invoked as ``model(**inputs)`` -- through ``DistributedDataParallel.forward``
when wrapped -- so DDP's reducer is armed exactly as in the reference.  Outside the armed context
(``generate``, plain ``model(input_ids)``) the head is an ordinary linear layer.
"""
from __future__ import annotations

import contextlib
from typing import Any, Optional, Tuple, Union

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops

__all__ = ["FusedLMHead", "FusedFinalNorm", "install_fused_head", "fused_head_armed", "fused_causal_lm_loss", "FusedCELossMixin"]


class FusedLMHead(nn.Module):
    """Stand-in for ``nn.Linear(hidden, vocab, bias=False)`` that can be bypassed.

    ``weight`` is the original Parameter object, so optimizer state, ``tie_weights`` and
    ``save_pretrained`` see no difference.
    """

    def __init__(self, linear: nn.Linear):
        super().__init__()
        if linear.bias is not None:
            raise ValueError("the fused head supports bias-free lm_head only (SmolLM3: modeling_smollm3.py:455)")
        self.in_features = linear.in_features
        self.out_features = linear.out_features
        self.weight = linear.weight
        self.armed = False
        self.last_stats = None  # {"n_valid", "n_correct", "entropy_sum"} of the latest fused call (token_stats)
        self.plan = None  # ops.LabelPlan made ahead of the forward (labels built on the device), else None

    def forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
        if self.armed:
            return hidden_states  # consumed by fused_causal_lm_loss; no [B,T,V] tensor is created
        return F.linear(hidden_states, self.weight)

    def extra_repr(self) -> str:
        return f"in_features={self.in_features}, out_features={self.out_features}, bias=False, fused=True"


class FusedFinalNorm(nn.Module):
    """Stand-in for the decoder's final RMSNorm (``model.model.norm``, modeling_smollm3.py:438) that can be
    bypassed: armed, it hands the pre-norm hidden states on and the fused loss normalises only the rows it
    scores (``ops.rmsnorm_gather``).  ``weight`` is the original Parameter under the original name."""

    def __init__(self, norm: nn.Module):
        super().__init__()
        self.weight = norm.weight
        self.variance_epsilon = float(norm.variance_epsilon)
        self.armed = False

    def forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
        if self.armed:
            return hidden_states
        dtype = hidden_states.dtype   # unarmed: the stock arithmetic (modeling_smollm3.py:271-276)
        h = hidden_states.to(torch.float32)
        h = h * torch.rsqrt(h.pow(2).mean(-1, keepdim=True) + self.variance_epsilon)
        return self.weight * h.to(dtype)

    def extra_repr(self) -> str:
        return f"{tuple(self.weight.shape)}, eps={self.variance_epsilon}, fused=True"


def _final_norm(inner: nn.Module) -> Optional[nn.Module]:
    return getattr(getattr(inner, "model", None), "norm", None)


def _norm_is_fusable(norm: Optional[nn.Module]) -> bool:
    """True when ``norm`` computes exactly what the fused kernels restate (weight * bf16(x * rsqrt(mean(x^2) + eps)),
    modeling_smollm3.py:271-276): checked on a probe input instead of trusting the class name."""
    if isinstance(norm, FusedFinalNorm):
        return True
    w = getattr(norm, "weight", None)
    if norm is None or w is None or not hasattr(norm, "variance_epsilon") or w.dim() != 1 or w.dtype != torch.bfloat16:
        return False
    try:
        with torch.no_grad():
            g = torch.Generator(device="cpu").manual_seed(0)
            x = (torch.randn(4, w.shape[0], generator=g) * 2).to(device=w.device, dtype=w.dtype)
            return bool(torch.equal(norm(x), FusedFinalNorm(norm)(x)))
    except Exception:
        return False


def _unwrap(model: nn.Module) -> nn.Module:
    while hasattr(model, "module") and isinstance(getattr(model, "module"), nn.Module):
        model = model.module
    return model


def install_fused_head(model: nn.Module, flags: int = 0, chunk_cols: int = 0, skip_ignored_rows: bool = True,
                       token_stats: bool = False, fuse_final_norm: Optional[bool] = None) -> nn.Module:
    """Patch a HF causal-LM in place (idempotent).  Returns the (unwrapped) model.

    ``skip_ignored_rows``: compact away positions labelled -100 (padding, prompt, packed-sample
    boundaries) before the GEMMs.  ``token_stats``: also produce the logits-free token-accuracy
    count and entropy sum; the latest values are kept in ``model.lm_head.last_stats``.
    ``fuse_final_norm``: also take over the decoder's final RMSNorm (``model.model.norm``) so that only
    the scored rows are normalised, inside the gather that feeds the head.  None (default) = when that
    module is a bf16 RMSNorm computing exactly the reference's formula; True = required (raises otherwise)."""
    inner = _unwrap(model)
    head = getattr(inner, "lm_head", None)
    if head is None:
        raise ValueError("model has no lm_head")
    if not isinstance(head, FusedLMHead):
        head = FusedLMHead(head)
        inner.lm_head = head
    if fuse_final_norm is not False:
        norm = _final_norm(inner)
        if _norm_is_fusable(norm):
            if not isinstance(norm, FusedFinalNorm):
                inner.model.norm = FusedFinalNorm(norm)
        elif fuse_final_norm:
            raise ValueError("fuse_final_norm needs model.model.norm to be a bf16 RMSNorm computing "
                             "weight * (x * rsqrt(mean(x^2) + variance_epsilon)).to(bf16)")
    # the model's own loss; kept on the model so that a deep copy (whose patched function still points at the
    # original's head) can be re-patched without stacking closures
    stock_loss = getattr(inner, "_b200sft_stock_loss", None) or inner.loss_function

    def loss_function(logits, labels, vocab_size=None, **kwargs):
        if head.armed:
            norm = _final_norm(inner)
            fused_norm = isinstance(norm, FusedFinalNorm) and norm.armed  # then `logits` are PRE-norm hidden states
            out = fused_causal_lm_loss(logits, head.weight, labels, flags=flags, chunk_cols=chunk_cols,
                                       skip_ignored_rows=skip_ignored_rows, return_stats=token_stats,
                                       plan=head.plan, norm_weight=norm.weight if fused_norm else None,
                                       norm_eps=norm.variance_epsilon if fused_norm else 1e-6, **kwargs)
            if token_stats:
                out, head.last_stats = out
            return out
        return stock_loss(logits=logits, labels=labels, vocab_size=vocab_size, **kwargs)

    if getattr(inner, "_b200sft_patched", None) != id(inner):
        inner._b200sft_stock_loss = stock_loss
        inner.loss_function = loss_function
        inner._b200sft_patched = id(inner)
    return inner


@contextlib.contextmanager
def fused_head_armed(model: nn.Module, plan: Optional["ops.LabelPlan"] = None):
    """While active, ``model(**inputs, labels=...)`` computes its loss with the fused kernels.
    ``plan``: an :class:`ops.LabelPlan` already made for this batch's labels (reused, not rebuilt)."""
    head = _unwrap(model).lm_head
    if not isinstance(head, FusedLMHead):
        raise RuntimeError("call install_fused_head(model) first")
    norm = _final_norm(_unwrap(model))
    norm = norm if isinstance(norm, FusedFinalNorm) else None
    prev = head.armed, head.plan, (norm.armed if norm is not None else False)
    head.armed, head.plan = True, plan
    if norm is not None:
        norm.armed = True
    try:
        yield
    finally:
        head.armed, head.plan = prev[0], prev[1]
        if norm is not None:
            norm.armed = prev[2]


def fused_causal_lm_loss(hidden_states: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor,
                         num_items_in_batch: Optional[Union[torch.Tensor, int]] = None,
                         ignore_index: int = -100, shift_labels: Optional[torch.Tensor] = None,
                         flags: int = 0, chunk_cols: int = 0, skip_ignored_rows: bool = True,
                         return_stats: bool = False, plan: Optional["ops.LabelPlan"] = None,
                         norm_weight: Optional[torch.Tensor] = None, norm_eps: float = 1e-6, **_unused: Any):
    """``ForCausalLMLoss`` (SP/transformers/loss/loss_utils.py:45-67) on hidden states.

    Same arguments and meaning, except that the first argument holds ``[B,T,H]`` hidden states:
    labels are shifted left (or ``shift_labels`` used as given), flattened, and the loss is the
    sum over non-ignored positions divided by ``num_items_in_batch`` -- or the mean over local
    non-ignored positions when that is None.  ``plan``: the :class:`ops.LabelPlan` of ``labels`` when the
    caller already ran the label front-end (labels built on the device); used as is.  ``norm_weight`` /
    ``norm_eps``: ``hidden_states`` are the decoder's output BEFORE its final RMSNorm, which is then applied
    to the scored rows only (modeling_smollm3.py:438,271-276).
    """
    if hidden_states.dim() == 2:
        hidden_states = hidden_states.unsqueeze(0)
    B, T, H = hidden_states.shape
    if shift_labels is not None:
        plan = None
    if plan is not None:
        if plan.shifted.numel() != B * T or plan.shifted.device != hidden_states.device:
            raise ValueError("label plan does not belong to this batch")
        if skip_ignored_rows and plan.row_index is None:
            raise ValueError("label plan was made without the compaction list")
        shifted = plan.shifted
    elif shift_labels is None:  # one launch: shift, counts and the live-row list the head compacts with
        plan = ops.label_frontend(labels=labels.to(hidden_states.device).reshape(B, T).contiguous(),
                                  ignore_index=ignore_index, compact=skip_ignored_rows)
        shifted = plan.shifted
    else:
        shifted = shift_labels.to(hidden_states.device).reshape(-1).to(torch.int32).contiguous()
    h2 = hidden_states.reshape(B * T, H)
    if h2.dtype != torch.bfloat16:
        h2 = h2.to(torch.bfloat16)
    w = weight if weight.dtype == torch.bfloat16 else weight.to(torch.bfloat16)
    return ops.fused_linear_cross_entropy(h2.contiguous(), w.contiguous(), shifted, num_items_in_batch,
                                          ignore_index=ignore_index, flags=flags, chunk_cols=chunk_cols,
                                          skip_ignored_rows=skip_ignored_rows, return_stats=return_stats, plan=plan,
                                          norm_weight=None if norm_weight is None else norm_weight.to(torch.bfloat16),
                                          norm_eps=norm_eps)


class FusedCELossMixin:
    """Mix into a ``transformers.Trainer`` / ``trl.SFTTrainer`` subclass (before the trainer class):

        class FusedSFTTrainer(FusedCELossMixin, SFTTrainer): ...

    ``compute_loss`` keeps the signature and return convention of
    ``Trainer.compute_loss`` (SP/transformers/trainer.py:1938-2020) for the default path the
    reference takes (no label smoother, no custom loss function): the model is called with the
    inputs plus ``num_items_in_batch``, the loss is read from its outputs and multiplied by the
    number of processes to cancel DDP's gradient averaging (:2013-2019).  On the first call with a
    DDP-wrapped model it also registers the gradient communication hook (seam #2), because
    ``accelerator.prepare`` wraps the model only inside ``train()`` (trainer.py:1463,1607).
    """

    b200sft_flags: int = 0
    b200sft_chunk_cols: int = 0
    b200sft_skip_ignored_rows: bool = True
    b200sft_token_stats: bool = False  # True: log mean_token_accuracy / entropy like TRL, without logits
    b200sft_fuse_final_norm: Optional[bool] = None  # final RMSNorm inside the head's row gather: None = when it applies
    b200sft_register_comm_hook: bool = True
    b200sft_nvlink_hook: bool = False  # True (or env B200SFT_COMM=nvlink): the fused NVLink reduce+scale kernel
    # > 0: under DDP, split the tied lm_head / embed_tokens gradient (tied_embedding.py) so that its bucket is reduced
    # at the start of backward; the value bounds the tokens of one micro-batch (per-device batch x sequence length)
    b200sft_tied_exchange_tokens: int = 0

    def _b200sft_prepare(self, model: nn.Module) -> None:
        inner = _unwrap(model)
        if (not isinstance(getattr(inner, "lm_head", None), FusedLMHead)
                or getattr(inner, "_b200sft_patched", None) != id(inner)   # e.g. a deep copy of a patched model
                or (self.b200sft_fuse_final_norm and not isinstance(_final_norm(inner), FusedFinalNorm))):  # required
            install_fused_head(inner, self.b200sft_flags, self.b200sft_chunk_cols, self.b200sft_skip_ignored_rows,
                               self.b200sft_token_stats, self.b200sft_fuse_final_norm)
        if (self.b200sft_register_comm_hook and isinstance(model, nn.parallel.DistributedDataParallel)
                and not getattr(model, "_b200sft_hook", False)):
            import os

            from .comm_hook import register_b200_comm_hook

            nvlink = self.b200sft_nvlink_hook or os.environ.get("B200SFT_COMM") == "nvlink"
            # the first compute_loss call happens at the same point on every rank, which the symmetric-memory
            # rendezvous of the NVLink hook needs
            register_b200_comm_hook(model, nvlink_kernel=nvlink)
            model._b200sft_hook = True
        if (self.b200sft_tied_exchange_tokens > 0 and isinstance(model, nn.parallel.DistributedDataParallel)
                and getattr(model, "_b200sft_tied_exchange", None) is None):
            from .tied_embedding import TiedEmbeddingExchange

            # allocates and exchanges symmetric memory: a collective, reached at the same point on every rank
            model._b200sft_tied_exchange = TiedEmbeddingExchange(model, self.b200sft_tied_exchange_tokens)

    def _b200sft_num_processes(self) -> int:
        acc = getattr(self, "accelerator", None)
        if acc is not None:
            return int(acc.num_processes)
        return int(getattr(self, "num_processes", 1))

    def compute_loss(self, model: nn.Module, inputs: dict, return_outputs: bool = False,
                     num_items_in_batch: Optional[Union[torch.Tensor, int]] = None
                     ) -> Union[torch.Tensor, Tuple[torch.Tensor, Any]]:
        if getattr(self, "label_smoother", None) is not None or getattr(self, "compute_loss_func", None) is not None:
            # not the reference's configuration; leave those paths to the stock implementation
            return super().compute_loss(model, inputs, return_outputs=return_outputs,
                                        num_items_in_batch=num_items_in_batch)
        self._b200sft_prepare(model)
        kwargs = {}
        if num_items_in_batch is not None and getattr(self, "model_accepts_loss_kwargs", True):
            kwargs["num_items_in_batch"] = num_items_in_batch
        inputs = dict(inputs)
        plan = inputs.pop("b200sft_plan", None)  # labels built on the device ahead of the forward
        exchange = getattr(model, "_b200sft_tied_exchange", None)
        if exchange is not None:  # tied-embedding gradient split: padded positions carry no embedding gradient
            exchange.set_attention_mask(inputs.get("attention_mask"))
        with fused_head_armed(model, plan):
            outputs = model(**{**inputs, **kwargs})
        if isinstance(outputs, dict) and "loss" not in outputs:
            raise ValueError(
                "The model did not return a loss from the inputs, only the following keys: "
                f"{','.join(outputs.keys())}. For reference, the inputs it received are {','.join(inputs.keys())}."
            )
        loss = outputs["loss"] if isinstance(outputs, dict) else outputs[0]
        stats = _unwrap(model).lm_head.last_stats
        if self.b200sft_token_stats and stats is not None and hasattr(self, "_metrics"):
            # [TRL, not on disk] SFTTrainer keeps per-mode lists in self._metrics and averages them at log time
            mode = "train" if model.training else "eval"
            nv = stats["n_valid"].clamp(min=1).float()
            self._metrics[mode]["mean_token_accuracy"].append((stats["n_correct"].float() / nv).item())
            self._metrics[mode]["entropy"].append((stats["entropy_sum"] / nv).item())
        args = getattr(self, "args", None)
        average = getattr(args, "average_tokens_across_devices", True)
        if average and num_items_in_batch is not None and getattr(self, "model_accepts_loss_kwargs", True):
            loss = loss * self._b200sft_num_processes()
        if return_outputs:
            return loss, _without_logits(outputs)
        return loss


def _without_logits(outputs: Any) -> Any:
    """The fused path never forms logits; what sits under the ``logits`` key is the [B,T,H] hidden states the armed
    head passed through.  Hand back outputs WITHOUT that key, so a caller that expects logits
    (``Trainer.prediction_step`` builds them from ``outputs.items()``, SP/transformers/trainer.py:2953-2975;
    ``compute_metrics``; ``preprocess_logits_for_metrics``) fails visibly instead of scoring hidden states.
    Setting the attribute to None is not enough: ``ModelOutput.__setattr__`` leaves the dict entry in place."""
    if isinstance(outputs, dict):  # ModelOutput is an OrderedDict subclass
        kept = {k: v for k, v in outputs.items() if k != "logits"}
        try:
            return type(outputs)(**kept)
        except TypeError:  # an output class whose constructor insists on logits: a plain dict carries the rest
            return kept
    return outputs
