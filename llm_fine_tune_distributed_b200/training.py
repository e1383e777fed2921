"""SFT driver with the reference's surface (training.py of thesteve0/llm-fine-tune-distributed),
running the data-parallel hot path on the B200 kernels.

Kept from the reference: the environment-variable configuration (training.py:19-23,56-60,240), the
freeze policy (:114-141), the parquet dataset + 90/10 split + chat formatting (:159-205), the
``SFTConfig`` field values (:258-287), ``trainer.train()`` (:300) and the JSON artefacts written by
rank 0 (:309-335).

``trl`` / ``accelerate`` are not installable in the build image, so :class:`MiniSFTTrainer`
restates the slice of ``transformers.Trainer._inner_training_loop`` the reference exercises
(SP/transformers/trainer.py:1697-1777): gradient-accumulation windows, ``num_items_in_batch`` over
the window and across ranks, ``DDP.no_sync`` on all but the last micro-batch, clip, AdamW, linear
schedule.  When ``trl`` is importable, :func:`build_trainer` returns a real ``SFTTrainer`` subclass
with the same two seams plugged in instead.
"""
from __future__ import annotations

import contextlib
import json
import math
import os
import re
import time
import zlib
from dataclasses import dataclass, field
from typing import Any, Callable, Dict, Iterator, List, Optional, Sequence

import torch
import torch.distributed as dist
import torch.nn as nn
from torch.nn.parallel import DistributedDataParallel as DDP

from . import ops
from .loss import FusedCELossMixin

MODEL_NAME = "HuggingFaceTB/SmolLM3-3B"  # training.py:54

# training.py:176-186 -- data, not code: the system prompt every sample is formatted with, in full (it is part of
# every tokenised row, so a shortened prompt would change the row lengths and hence the padding mask)
SYSTEM_PROMPT = """You are a wilderness survival and practical skills expert. Your mission is to provide comprehensive, detailed guidance on essential survival and practical skills. Give thorough, step-by-step instructions with explanations of why each step matters.

Your expertise covers:
- Wilderness Survival Basics: Rule of 3s (3 minutes without air, 3 hours without shelter in harsh conditions, 3 days without water, 3 weeks without food), emergency signaling techniques, essential knots, identifying poisonous plants and safe alternatives
- Basic First Aid: Treatment for cuts, burns, sprains, shock, and emergency care procedures
- Simple Car Maintenance: Checking fluids (oil, coolant, brake, transmission), tire inspection and pressure, lights and electrical systems
- Basic Cooking Techniques: Food safety, preparation methods, cooking over open fires, food preservation
- Common Measurement Conversions: Imperial to metric, cooking measurements, distance and weight conversions
- Essential Knots: Bowline, clove hitch, trucker's hitch, figure-eight, sheet bend, and their practical applications

Always provide detailed explanations, safety warnings when relevant, and multiple approaches when possible. Your responses should be comprehensive enough to help someone learn and apply these skills safely and effectively. Aim for thorough, educational responses rather than brief answers."""


# --------------------------------------------------------------------------------------------
# configuration (environment variables, as the reference reads them)
# --------------------------------------------------------------------------------------------
def setup_distributed() -> tuple:
    """training.py:16-42: read WORLD_SIZE / RANK / LOCAL_RANK / MASTER_* and re-export them."""
    world_size = int(os.getenv("WORLD_SIZE", "1"))
    rank = int(os.getenv("RANK", "0"))
    local_rank = int(os.getenv("LOCAL_RANK", "0"))
    if world_size > 1:
        os.environ["MASTER_ADDR"] = os.getenv("MASTER_ADDR", "127.0.0.1")
        os.environ["MASTER_PORT"] = os.getenv("MASTER_PORT", "23456")
        os.environ["WORLD_SIZE"], os.environ["RANK"], os.environ["LOCAL_RANK"] = map(str, (world_size, rank, local_rank))
    return world_size, rank, local_rank


@dataclass
class SFTSettings:
    """The reference's ``SFTConfig(...)`` literal (training.py:258-287), field for field."""

    output_dir: str = "/tmp/models/checkpoints"
    per_device_train_batch_size: int = 8
    per_device_eval_batch_size: int = 8
    gradient_accumulation_steps: int = 4
    learning_rate: float = 5e-5
    max_grad_norm: float = 1.0
    num_train_epochs: float = 4
    max_steps: int = -1
    logging_steps: int = 2
    logging_first_step: bool = True
    bf16: bool = True
    eval_steps: int = 10
    gradient_checkpointing: bool = True
    dataloader_drop_last: bool = True
    max_seq_length: int = 1024
    packing: bool = False
    ddp_backend: Optional[str] = None
    ddp_find_unused_parameters: bool = False
    ddp_bucket_cap_mb: int = 50
    local_rank: int = 0
    weight_decay: float = 0.0
    average_tokens_across_devices: bool = True
    seed: int = 42

    @classmethod
    def from_env(cls, world_size: int, local_rank: int) -> "SFTSettings":
        lr = float(os.getenv("LEARNING_RATE", "5e-5"))
        out = os.getenv("OUTPUT_DIR", "/tmp/models")
        return cls(
            output_dir=f"{out}/checkpoints",
            per_device_train_batch_size=int(os.getenv("BATCH_SIZE", "8")),
            per_device_eval_batch_size=int(os.getenv("BATCH_SIZE", "8")),
            learning_rate=lr * world_size if world_size > 1 else lr,  # training.py:263
            num_train_epochs=float(os.getenv("EPOCHS", "4")),
            max_steps=int(os.getenv("MAX_STEPS", "-1")),
            eval_steps=int(os.getenv("EVAL_STEPS", "10")),
            gradient_accumulation_steps=int(os.getenv("GRAD_ACCUM", "4")),
            ddp_backend="nccl" if world_size > 1 else None,
            local_rank=local_rank,
            packing=os.getenv("PACKING", "0") == "1",  # reference: False (training.py:283); 1 = padding-free batches
        )


def apply_freeze_policy(model: nn.Module) -> tuple:
    """training.py:114-145: freeze everything, then unfreeze the last two decoder layers and the
    output head (which is tied to the input embedding on SmolLM3).  Returns (trainable, total)."""
    for p in model.parameters():
        p.requires_grad = False
    layers = None
    if hasattr(model, "model") and hasattr(model.model, "layers"):
        layers = model.model.layers
    elif hasattr(model, "transformer") and hasattr(model.transformer, "h"):
        layers = model.transformer.h
    elif hasattr(model, "gpt_neox") and hasattr(model.gpt_neox, "layers"):
        layers = model.gpt_neox.layers
    if layers is None:
        for p in model.parameters():
            p.requires_grad = True
    else:
        for p in layers[-2:].parameters():
            p.requires_grad = True
    head = getattr(model, "embed_out", None) or getattr(model, "lm_head", None)
    if head is not None:
        for p in head.parameters():
            p.requires_grad = True
    trainable = sum(p.numel() for p in model.parameters() if p.requires_grad)
    total = sum(p.numel() for p in model.parameters())
    return trainable, total


# --------------------------------------------------------------------------------------------
# data: parquet rows -> chat text -> token ids -> right-padded batches with -100 labels
# --------------------------------------------------------------------------------------------
def load_qa_rows(path: str) -> List[Dict[str, str]]:
    import pandas as pd

    df = pd.read_parquet(path)
    return [{"full-question": q, "answer": a} for q, a in zip(df["full-question"], df["answer"])]


def train_test_split(rows: Sequence[Any], test_size: float = 0.1, seed: int = 42) -> tuple:
    """training.py:164 ``train_test_split(test_size=0.1, seed=42)`` (same sizes; the permutation is
    numpy's rather than datasets', which is not installed here)."""
    import numpy as np

    idx = np.random.default_rng(seed).permutation(len(rows))
    n_test = int(math.ceil(len(rows) * test_size))
    test = [rows[i] for i in idx[:n_test]]
    train = [rows[i] for i in idx[n_test:]]
    return train, test


def format_prompt(example: Dict[str, str]) -> Dict[str, Any]:
    """training.py:189-199."""
    return {"messages": [
        {"role": "system", "content": SYSTEM_PROMPT},
        {"role": "user", "content": example["full-question"]},
        {"role": "assistant", "content": example["answer"]},
    ]}


class ByteChatTokenizer:
    """Deterministic stand-in used when the SmolLM3 tokenizer cannot be fetched (no network): ChatML-style framing;
    the text is cut into pieces of at most four word characters (with their leading space) or one punctuation /
    whitespace character, and each piece's UTF-8 bytes are hashed (CRC-32) into the vocabulary.  That gives about one
    token per 3.5 characters, the scale of a BPE vocabulary, so that the reference's full system prompt
    (training.py:176-186, ~1,500 characters) takes ~400 of the 1,024 positions as it does with the real tokenizer
    instead of all of them.  Only shapes and masking matter for the hot path; real runs pass a HF tokenizer."""

    pad_token_id = 0
    eos_token_id = 1
    _PIECES = re.compile(r" ?\w{1,4}| ?[^\w\s]|\s", re.UNICODE)

    def __init__(self, vocab_size: int = 128256):
        self.vocab_size = vocab_size

    def encode(self, text: str) -> List[int]:
        return [2 + zlib.crc32(p.encode("utf-8")) % (self.vocab_size - 2) for p in self._PIECES.findall(text)]

    def apply_chat_template(self, messages: List[Dict[str, str]], return_assistant_mask: bool = False):
        """Token ids of the templated conversation; with ``return_assistant_mask`` also a 0/1 list
        marking the assistant's tokens (what TRL's ``assistant_only_loss`` trains on)."""
        ids: List[int] = []
        mask: List[int] = []
        for m in messages:
            head = self.encode(f"<|im_start|>{m['role']}\n")
            body = self.encode(f"{m['content']}<|im_end|>\n")
            ids.extend(head + body)
            mask.extend([0] * len(head) + [1 if m["role"] == "assistant" else 0] * len(body))
        ids.append(self.eos_token_id)
        mask.append(1)
        return (ids, mask) if return_assistant_mask else ids


def collate(samples: List[List[int]], pad_id: int, max_len: int, pad_to: Optional[int] = None,
            masks: Optional[List[List[int]]] = None, device_labels: bool = False) -> Dict[str, torch.Tensor]:
    """[TRL collator semantics, SURVEY.md a3] truncate to ``max_len``, right-pad ``input_ids`` with
    the pad id, ``attention_mask`` with 0 and ``labels`` with -100.  Padding is the only mask the
    reference produces for plain ``messages`` data.  ``device_labels``: leave ``labels`` out and hand
    over the raw masks (``completion_mask`` when ``masks`` is given); ``ops.label_frontend`` then
    builds the same labels on the GPU."""
    samples = [s[:max_len] for s in samples]
    width = pad_to or max(len(s) for s in samples)
    B = len(samples)
    input_ids = torch.full((B, width), pad_id, dtype=torch.int64)
    attn = torch.zeros((B, width), dtype=torch.int64)
    if device_labels:
        comp = torch.zeros((B, width), dtype=torch.int64) if masks is not None else None
        for i, s in enumerate(samples):
            input_ids[i, : len(s)] = torch.tensor(s, dtype=torch.int64)
            attn[i, : len(s)] = 1
            if comp is not None:
                comp[i, : len(s)] = torch.tensor(masks[i][: len(s)], dtype=torch.int64).ne(0)
        out = {"input_ids": input_ids, "attention_mask": attn}
        if comp is not None:
            out["completion_mask"] = comp
        return out
    labels = torch.full((B, width), -100, dtype=torch.int64)
    for i, s in enumerate(samples):
        t = torch.tensor(s, dtype=torch.int64)
        input_ids[i, : len(s)] = t
        attn[i, : len(s)] = 1
        labels[i, : len(s)] = t
        if masks is not None:  # completion-only loss: score the assistant's tokens only
            keep = torch.tensor(masks[i][: len(s)], dtype=torch.bool)
            labels[i, : len(s)] = torch.where(keep, t, torch.full_like(t, -100))
    return {"input_ids": input_ids, "attention_mask": attn, "labels": labels}


def collate_flattened(samples: List[List[int]], max_len: int, masks: Optional[List[List[int]]] = None,
                      separator_id: int = -100, device_labels: bool = False) -> Dict[str, Any]:
    """Padding-free batch, as ``transformers.DataCollatorWithFlattening(return_flash_attn_kwargs=True)``
    builds it (SP/transformers/data/data_collator.py:1364-1470): the samples are concatenated into one
    ``[1, total]`` row, ``position_ids`` restart at every sample, the first label of every sample is
    ``separator_id`` (:1426-1428) so no sample is scored against its predecessor, and the FlashAttention
    varlen arguments (``cu_seq_lens_*`` int32, ``max_length_*`` int) ride along.  ``masks`` (0/1 per
    token) additionally turn non-assistant tokens into -100 (completion-only loss).  ``device_labels``:
    no ``labels``; ``position_ids`` (and ``completion_mask``) are what ``ops.label_frontend`` makes them from."""
    samples = [s[:max_len] for s in samples]
    ids: List[int] = []
    labels: List[int] = []
    comp: List[int] = []
    pos: List[int] = []
    cu = [0]
    for i, smp in enumerate(samples):
        lab = list(smp)
        if masks is not None:
            lab = [t if m else -100 for t, m in zip(smp, masks[i][:max_len])]
            comp += [1 if m else 0 for m in masks[i][: len(smp)]]
        ids += smp
        labels += [separator_id] + lab[1:]
        pos += list(range(len(smp)))
        cu.append(cu[-1] + len(smp))
    longest = max(len(smp) for smp in samples)
    out = {"input_ids": torch.tensor([ids], dtype=torch.int64), "labels": torch.tensor([labels], dtype=torch.int64),
           "position_ids": torch.tensor([pos], dtype=torch.int64),
           "cu_seq_lens_q": torch.tensor(cu, dtype=torch.int32), "cu_seq_lens_k": torch.tensor(cu, dtype=torch.int32),
           "max_length_q": longest, "max_length_k": longest}
    if device_labels:
        del out["labels"]
        if masks is not None:
            out["completion_mask"] = torch.tensor([comp], dtype=torch.int64)
    return out


class ShardedBatches:
    """Per-rank micro-batch iterator: shuffles with the epoch seed, shards by rank, drops the
    ragged tail (``dataloader_drop_last=True``, training.py:281)."""

    def __init__(self, tokens: List[List[int]], batch_size: int, rank: int, world: int, pad_id: int, max_len: int, seed: int,
                 masks: Optional[List[List[int]]] = None, padding_free: bool = False, device_labels: bool = False):
        self.tokens, self.bs, self.rank, self.world = tokens, batch_size, rank, world
        self.pad_id, self.max_len, self.seed = pad_id, max_len, seed
        self.masks, self.padding_free, self.device_labels = masks, padding_free, device_labels

    def __len__(self) -> int:
        return len(self.tokens) // (self.bs * self.world)

    def epoch(self, epoch: int) -> Iterator[Dict[str, torch.Tensor]]:
        g = torch.Generator().manual_seed(self.seed + epoch)
        perm = torch.randperm(len(self.tokens), generator=g).tolist()
        per_step = self.bs * self.world
        for b in range(len(self)):
            chunk = perm[b * per_step: (b + 1) * per_step][self.rank * self.bs: (self.rank + 1) * self.bs]
            toks = [self.tokens[i] for i in chunk]
            msk = [self.masks[i] for i in chunk] if self.masks is not None else None
            if self.padding_free:
                yield collate_flattened(toks, self.max_len, msk, device_labels=self.device_labels)
            else:
                yield collate(toks, self.pad_id, self.max_len, masks=msk, device_labels=self.device_labels)


# --------------------------------------------------------------------------------------------
# trainer
# --------------------------------------------------------------------------------------------
class TrainingHistoryCallback:
    """training.py:215-221."""

    def __init__(self):
        self.history: List[Dict[str, float]] = []

    def on_log(self, logs: Dict[str, float]) -> None:
        if logs:
            self.history.append(dict(logs))


class PerplexityCallback:
    """training.py:224-234."""

    def on_log(self, logs: Dict[str, float]) -> None:
        if logs and "loss" in logs:
            logs["perplexity"] = math.exp(min(logs["loss"], 80.0))
        if logs and "eval_loss" in logs:
            logs["eval_perplexity"] = math.exp(min(logs["eval_loss"], 80.0))


@dataclass
class TrainResult:
    training_loss: float
    global_step: int
    tokens_seen: int
    seconds: float
    metrics: Dict[str, float] = field(default_factory=dict)


class MiniSFTTrainer(FusedCELossMixin):
    """The loop ``SFTTrainer.train()`` runs for the reference's configuration, with both seams
    plugged in: :meth:`compute_loss` (inherited) and the DDP comm hook.

    ``loss_impl(model, inputs, num_items_in_batch)`` replaces the fused loss when given -- the
    CPU/gloo plumbing tests inject the oracle there; product runs leave it None and need a B200.
    """

    def __init__(self, model: nn.Module, args: SFTSettings, train_batches: ShardedBatches,
                 eval_batches: Optional[ShardedBatches] = None, callbacks: Sequence[Any] = (),
                 device: Optional[torch.device] = None, loss_impl: Optional[Callable] = None,
                 comm_hook_state: Any = None):
        self.args = args
        self.callbacks = list(callbacks)
        self.device = device or torch.device("cuda", args.local_rank)
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.num_processes = self.world
        self.model_accepts_loss_kwargs = True  # SmolLM3 forward takes **kwargs (trainer.py:1925-1927)
        self.loss_impl = loss_impl
        env = os.getenv("FUSE_FINAL_NORM")  # final RMSNorm inside the head's gather: unset = when it applies, 0 / 1 = off / required
        self.b200sft_fuse_final_norm = None if env is None else env == "1"
        self.model = model.to(self.device)
        if args.gradient_checkpointing and hasattr(self.model, "gradient_checkpointing_enable"):
            self.model.gradient_checkpointing_enable()
            if hasattr(self.model, "config"):
                self.model.config.use_cache = False
        self.model_wrapped: nn.Module = self.model
        if self.world > 1:
            kw = {"device_ids": [self.device.index]} if self.device.type == "cuda" else {}
            self.model_wrapped = DDP(self.model, bucket_cap_mb=args.ddp_bucket_cap_mb,
                                     find_unused_parameters=args.ddp_find_unused_parameters, **kw)
            if comm_hook_state is not None:
                from .comm_hook import b200_allreduce_hook

                self.model_wrapped.register_comm_hook(comm_hook_state, b200_allreduce_hook)
                self.model_wrapped._b200sft_hook = True
            if os.getenv("TIED_EXCHANGE", "0") == "1" and self.device.type == "cuda":
                # reduce the tied head/embedding bucket at the start of backward (tied_embedding.py)
                self.b200sft_tied_exchange_tokens = args.per_device_train_batch_size * args.max_seq_length
        self.train_batches, self.eval_batches = train_batches, eval_batches
        params = [p for p in self.model.parameters() if p.requires_grad]
        self.fused_tail = self.device.type == "cuda" and all(p.dtype == torch.bfloat16 for p in params)
        if self.fused_tail:  # grad-norm clip + AdamW in this library's multi-tensor kernels
            from .optim import B200AdamW

            self.optimizer = B200AdamW(params, lr=args.learning_rate, weight_decay=args.weight_decay)
        else:                # CPU plumbing runs (gloo) and non-bf16 models: torch's own step tail
            self.optimizer = torch.optim.AdamW(params, lr=args.learning_rate, weight_decay=args.weight_decay)
        steps_per_epoch = max(len(train_batches) // args.gradient_accumulation_steps, 1)
        self.max_steps = args.max_steps if args.max_steps > 0 else int(math.ceil(args.num_train_epochs * steps_per_epoch))
        self.lr_scheduler = torch.optim.lr_scheduler.LambdaLR(
            self.optimizer, lambda s: max(0.0, 1.0 - s / max(self.max_steps, 1)))  # linear decay, Trainer default
        self.global_step = 0
        self.best_eval_loss: Optional[float] = None
        self.best_step: Optional[int] = None

    # -- pieces of _inner_training_loop ------------------------------------------------------
    def get_batch_samples(self, it: Iterator, n: int) -> tuple:
        """trainer.py:2092-2159: pull the window's micro-batches, count non-ignored labels over the
        whole window (UNSHIFTED labels, :2136) and sum the count across ranks (:2141-2143)."""
        batches = []
        for _ in range(n):
            try:
                batches.append(next(it))
            except StopIteration:
                break
        if not batches:
            return batches, None
        if any("labels" not in b for b in batches):  # labels are built on the device; the count comes with them
            batches = [self._on_device(b) for b in batches]
            count = sum(b["b200sft_plan"].counts[0] if "b200sft_plan" in b else b["labels"].ne(-100).sum()
                        for b in batches)
        else:
            count = sum(b["labels"].ne(-100).sum() for b in batches).to(self.device)
        if self.args.average_tokens_across_devices and self.world > 1:
            dist.all_reduce(count, op=dist.ReduceOp.SUM)
        return batches, count

    def _on_device(self, inputs: Dict[str, Any]) -> Dict[str, Any]:
        """H2D; when the collator left ``labels`` out (``device_labels``) build them here with the label
        front-end -- one launch gives the labels, the window's count and the plan the fused loss reuses."""
        inputs = {k: (v.to(self.device, non_blocking=True) if torch.is_tensor(v) else v) for k, v in inputs.items()}
        if "labels" not in inputs:
            plan = ops.label_frontend(inputs["input_ids"], None, inputs.get("attention_mask"),
                                      inputs.pop("completion_mask", None), inputs.get("position_ids"),
                                      compact=bool(getattr(self, "b200sft_skip_ignored_rows", True)), return_labels=True)
            inputs["labels"] = plan.labels
            if self.loss_impl is None:
                inputs["b200sft_plan"] = plan
        return inputs

    def training_step(self, inputs: Dict[str, torch.Tensor], num_items_in_batch: Optional[torch.Tensor]) -> torch.Tensor:
        """trainer.py:1867-1936 for this configuration: H2D, compute_loss, backward.  No division
        by the accumulation steps: the loss is already normalised by the window's token count."""
        self.model_wrapped.train()
        inputs = self._on_device(inputs)
        if self.loss_impl is not None:
            loss = self.loss_impl(self.model_wrapped, inputs, num_items_in_batch)
            if num_items_in_batch is not None and self.args.average_tokens_across_devices:
                loss = loss * self.num_processes
        else:
            loss = self.compute_loss(self.model_wrapped, inputs, num_items_in_batch=num_items_in_batch)
        loss.backward()
        return loss.detach()

    def _log(self, logs: Dict[str, float]) -> None:
        for cb in self.callbacks:
            if hasattr(cb, "on_log"):
                cb.on_log(logs)

    @torch.no_grad()
    def evaluate(self) -> Dict[str, float]:
        """``Trainer.evaluate`` for this configuration (eval_strategy="steps", eval_steps=10,
        metric_for_best_model="eval_loss": training.py:270-275): the loss of every validation batch as
        ``Trainer.prediction_step`` computes it (SP/transformers/trainer.py:2951-2957) -- same fused path, no_grad,
        ``num_items_in_batch`` = the batch's UNSHIFTED non-ignored label count summed over ranks
        (``_get_num_items_in_batch([inputs])``), handed to ``compute_loss``, which multiplies by the number of
        processes -- averaged over batches and ranks."""
        if self.eval_batches is None or len(self.eval_batches) == 0:
            return {}
        was_training = self.model_wrapped.training
        self.model.eval()
        t0 = time.time()
        total = torch.zeros((), device=self.device)
        n = 0
        for inputs in self.eval_batches.epoch(0):
            inputs = self._on_device(inputs)
            plan = inputs.get("b200sft_plan")
            count = (plan.counts[0] if plan is not None else inputs["labels"].ne(-100).sum()).to(self.device).clone()
            if self.args.average_tokens_across_devices and self.world > 1:
                dist.all_reduce(count, op=dist.ReduceOp.SUM)
            if self.loss_impl is not None:
                loss = self.loss_impl(self.model, inputs, count).float()
                if self.args.average_tokens_across_devices:
                    loss = loss * self.num_processes
                total += loss
            else:  # unwrapped model, as prediction_step uses it
                total += self.compute_loss(self.model, inputs, num_items_in_batch=count).float()
            n += 1
        if self.world > 1:
            dist.all_reduce(total, op=dist.ReduceOp.SUM)
            total /= self.world
        self.model.train(was_training)
        dt = time.time() - t0
        samples = n * self.eval_batches.bs * self.world
        logs = {"eval_loss": float(total) / max(n, 1), "eval_runtime": round(dt, 4),
                "eval_samples_per_second": round(samples / dt, 3) if dt > 0 else 0.0,
                "eval_steps_per_second": round(n / dt, 3) if dt > 0 else 0.0,
                "epoch": round(self.global_step / max(len(self.train_batches) // self.args.gradient_accumulation_steps, 1), 4)}
        if self.best_eval_loss is None or logs["eval_loss"] < self.best_eval_loss:   # greater_is_better=False
            self.best_eval_loss, self.best_step = logs["eval_loss"], self.global_step
        self._log(logs)
        return logs

    def train(self) -> TrainResult:
        args = self.args
        ga = args.gradient_accumulation_steps
        t0 = time.time()
        tr_loss = torch.zeros((), device=self.device)
        logged_loss, tokens_seen, epoch = 0.0, 0, 0
        last_logged_step = 0
        while self.global_step < self.max_steps:
            it = self.train_batches.epoch(epoch)
            while self.global_step < self.max_steps:
                batches, num_items = self.get_batch_samples(it, ga)
                if len(batches) < ga:  # epoch exhausted (drop the ragged window)
                    break
                for i, inputs in enumerate(batches):
                    sync = (i == len(batches) - 1)
                    ctx = contextlib.nullcontext() if (sync or self.world == 1) else self.model_wrapped.no_sync()
                    with ctx:  # trainer.py:1725-1733 -> DDP.no_sync
                        tr_loss += self.training_step(inputs, num_items)
                    tokens_seen += int(inputs["attention_mask"].sum()) if "attention_mask" in inputs else inputs["input_ids"].numel()
                if self.fused_tail:
                    grad_norm = self.optimizer.clip_and_step(args.max_grad_norm)
                else:
                    grad_norm = torch.nn.utils.clip_grad_norm_(
                        [p for p in self.model.parameters() if p.requires_grad], args.max_grad_norm)
                    self.optimizer.step()
                self.lr_scheduler.step()
                self.optimizer.zero_grad(set_to_none=True)
                self.global_step += 1
                if (self.global_step % args.logging_steps == 0) or (args.logging_first_step and self.global_step == 1):
                    total = tr_loss.clone()
                    if self.world > 1:  # nested_gather(tr_loss).mean(), trainer.py:2060
                        dist.all_reduce(total, op=dist.ReduceOp.SUM)
                        total /= self.world
                    total = float(total)
                    logs = {"loss": round((total - logged_loss) / (self.global_step - last_logged_step), 4),
                            "grad_norm": float(grad_norm), "learning_rate": self.lr_scheduler.get_last_lr()[0],
                            "epoch": round(self.global_step / max(len(self.train_batches) // ga, 1), 4)}
                    logged_loss, last_logged_step = total, self.global_step
                    self._log(logs)
                if args.eval_steps > 0 and self.global_step % args.eval_steps == 0:
                    self.evaluate()
            epoch += 1
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        dt = time.time() - t0
        total_tokens = tokens_seen * self.world
        return TrainResult(training_loss=float(tr_loss) / max(self.global_step, 1), global_step=self.global_step,
                           tokens_seen=total_tokens, seconds=dt,
                           metrics={"train_tokens_per_second": total_tokens / dt if dt > 0 else 0.0})

    def save_model(self, path: str) -> None:
        os.makedirs(path, exist_ok=True)
        if hasattr(self.model, "save_pretrained"):
            self.model.save_pretrained(path)
        else:
            torch.save(self.model.state_dict(), os.path.join(path, "model.pt"))


def build_trainer(model: nn.Module, settings: SFTSettings, train_rows, val_rows, tokenizer, rank: int, world: int,
                  callbacks: Sequence[Any] = (), **mini_kwargs) -> Any:
    """A ``trl.SFTTrainer`` subclass with both seams when trl is importable (the reference's exact
    host), otherwise :class:`MiniSFTTrainer`."""
    try:  # pragma: no cover - trl is not installable in the build image
        from trl import SFTConfig, SFTTrainer

        class FusedSFTTrainer(FusedCELossMixin, SFTTrainer):
            pass

        from datasets import Dataset

        fields = SFTConfig.__dataclass_fields__
        kw = {k: v for k, v in settings.__dict__.items() if k in fields}
        if "max_seq_length" not in fields and "max_length" in fields:  # TRL >= 0.20 renamed the field
            kw["max_length"] = settings.max_seq_length
        cfg = SFTConfig(**kw)
        # [TRL, not on disk] untested in the build image (trl / accelerate are not installable there): INTEGRATION.md 1
        return FusedSFTTrainer(model=model, args=cfg,
                               train_dataset=Dataset.from_list([format_prompt(r) for r in train_rows]),
                               eval_dataset=Dataset.from_list([format_prompt(r) for r in val_rows]),
                               processing_class=tokenizer if hasattr(tokenizer, "save_pretrained") else None,
                               callbacks=list(callbacks))
    except ImportError:
        pass
    completion_only = os.getenv("COMPLETION_ONLY", "0") == "1" and isinstance(tokenizer, ByteChatTokenizer)

    def mk(rows, bs):
        msgs = [format_prompt(r)["messages"] for r in rows]
        if completion_only:
            pairs = [tokenizer.apply_chat_template(m, return_assistant_mask=True) for m in msgs]
            toks, masks = [p[0] for p in pairs], [p[1] for p in pairs]
        else:
            toks, masks = [tokenizer.apply_chat_template(m) for m in msgs], None
        return ShardedBatches(toks, bs, rank, world, tokenizer.pad_token_id, settings.max_seq_length, settings.seed,
                              masks=masks, padding_free=settings.packing,
                              device_labels=os.getenv("DEVICE_LABELS", "0") == "1")  # labels built by ops.label_frontend

    return MiniSFTTrainer(model, settings, mk(train_rows, settings.per_device_train_batch_size),
                          mk(val_rows, settings.per_device_eval_batch_size), callbacks=callbacks, **mini_kwargs)


def make_model(kind: str, device: torch.device, attn: Optional[str] = None) -> nn.Module:
    """Random-init SmolLM3 in bf16: ``3b`` is the architecture of HuggingFaceTB/SmolLM3-3B
    (configuration_smollm3.py defaults), ``tiny`` is BASELINE.json config 1 (2 layers, hidden 256,
    full 128,256 vocabulary)."""
    from transformers import SmolLM3Config, SmolLM3ForCausalLM

    if kind == "tiny":
        cfg = SmolLM3Config(num_hidden_layers=2, hidden_size=256, intermediate_size=688, num_attention_heads=4,
                            num_key_value_heads=1, vocab_size=128256, max_position_embeddings=4096,
                            tie_word_embeddings=True)
    else:
        cfg = SmolLM3Config(tie_word_embeddings=True)
    if attn:
        cfg._attn_implementation = attn
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        with torch.device(device):
            model = SmolLM3ForCausalLM(cfg)
    finally:
        torch.set_default_dtype(prev)
    return model


def main() -> None:  # mirrors training.py:49-349
    world_size, rank, local_rank = setup_distributed()
    if not torch.cuda.is_available():
        raise RuntimeError("CUDA not available")  # training.py:81-83
    torch.cuda.set_device(local_rank)  # the reference's bare "cuda" lands every rank on GPU 0
    device = torch.device("cuda", local_rank)
    if world_size > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=device)
    output_dir = os.getenv("OUTPUT_DIR", "/tmp/models")
    dataset_path = os.getenv("DATASET_PATH", "data/qa_dataset.parquet")
    settings = SFTSettings.from_env(world_size, local_rank)
    if rank == 0:
        os.makedirs(f"{output_dir}/best_model", exist_ok=True)

    tokenizer = None
    model = None
    if os.getenv("RANDOM_INIT", "0") != "1":
        try:
            from transformers import AutoModelForCausalLM, AutoTokenizer

            tokenizer = AutoTokenizer.from_pretrained(MODEL_NAME, cache_dir="/tmp/model_cache")
            tokenizer.pad_token = tokenizer.eos_token
            tokenizer.padding_side = "right"
            model = AutoModelForCausalLM.from_pretrained(MODEL_NAME, torch_dtype=torch.bfloat16,
                                                         cache_dir="/tmp/model_cache",
                                                         attn_implementation="flash_attention_2")
        except Exception as e:  # offline: fall through to random init of the same architecture
            print(f"could not load {MODEL_NAME} ({type(e).__name__}); using random-init weights")
    if model is None:
        attn = os.getenv("ATTN_IMPL")
        if attn is None:  # the reference loads with attn_implementation="flash_attention_2" (training.py:101)
            try:
                import flash_attn  # noqa: F401

                attn = "flash_attention_2"
            except Exception:
                attn = None
        torch.manual_seed(settings.seed)  # random init is the only seeded draw before Trainer's own set_seed(args.seed)
        model = make_model(os.getenv("MODEL_CONFIG", "3b"), device, attn)
    if tokenizer is None:
        tokenizer = ByteChatTokenizer(model.config.vocab_size)
    model = model.to(device)
    trainable, total = apply_freeze_policy(model)
    print(f"Trainable parameters: {trainable:,} / {total:,} ({100 * trainable / total:.2f}%)")

    rows = load_qa_rows(dataset_path)
    train_rows, val_rows = train_test_split(rows, 0.1, 42)
    history, ppl = TrainingHistoryCallback(), PerplexityCallback()
    trainer = build_trainer(model, settings, train_rows, val_rows, tokenizer, rank, world_size,
                            callbacks=[ppl, history], device=device)
    result = trainer.train()

    if rank == 0:
        trainer.save_model(f"{output_dir}/best_model")
        with open(f"{output_dir}/training_history.json", "w") as f:
            json.dump(history.history, f, indent=2)
        summary = {
            "model_name": MODEL_NAME, "dataset_path": dataset_path, "epochs": settings.num_train_epochs,
            "batch_size": settings.per_device_train_batch_size, "learning_rate": settings.learning_rate,
            "trainable_params": trainable, "total_params": total, "training_samples": len(train_rows),
            "validation_samples": len(val_rows), "final_train_loss": getattr(result, "training_loss", None),
            "world_size": world_size, "distributed_training": world_size > 1,
            "train_tokens_per_second": getattr(result, "metrics", {}).get("train_tokens_per_second"),
        }
        exchange = getattr(getattr(trainer, "model_wrapped", None), "_b200sft_tied_exchange", None)
        if exchange is not None:  # beyond the reference's keys: optimizer steps whose tied gradient went through the split
            summary["tied_exchange_steps"] = exchange.exchanges
        with open(f"{output_dir}/training_summary.json", "w") as f:
            json.dump(summary, f, indent=2)
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
