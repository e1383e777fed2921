#!/usr/bin/env python
"""bench.py -- headline benchmark: SFT trained tokens/sec, SmolLM3-3B, on N B200s of one box.

    python bench.py --gpus N --steps K --warmup W            (N=1; N>1 under torchrun)
    python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1]/[2]): SmolLM3-3B architecture, random-init bf16, sequence 1024,
per-device batch 8, gradient accumulation 2, the reference's freeze policy (last two decoder layers
+ tied lm_head/embedding trainable, training.py:114-141), gradient checkpointing, FlashAttention-2,
fused AdamW, grad clip 1.0.  A "step" is one optimizer step = 2 micro-batches of [8,1024] per GPU.
The LM head + cross-entropy (forward and backward) run in this repository's sm_100a kernels through
the compute_loss seam; with N>1 the model is wrapped in DDP (bucket_cap_mb=50) with this
repository's comm hook.  Synthetic token ids; weights and tokens are random (no network).

One JSON line is printed by rank 0; see the task contract for the keys.  ``--impl reference`` times
the reference's own path (stock transformers lm_head + ForCausalLMLoss + autograd, torch CPU
kernels) on the host cores on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "SFT trained tokens/sec SmolLM3-3B"
UNIT = "tokens/s"
H, V = 2048, 128256


def read_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"burst": d["bf16_tflops"], "sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "hbm": d["hbm_gbs"], "kind": "measured (MEASURED_PEAKS.json)"}
    return {"burst": 1590.0, "sustained": 1400.0, "hbm": 6650.0, "kind": "fallback (B200_PROFILING.md)"}


# ------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi, during the timed region)
# ------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int, interval_ms: int = 200):
        self.gpu_index, self.proc, self.lines, self.interval_ms = gpu_index, None, [], interval_ms

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", str(self.interval_ms), "-i", str(self.gpu_index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------
# model / data
# ------------------------------------------------------------------------------------------
def pick_attn() -> str:
    try:
        import flash_attn  # noqa: F401

        return "flash_attention_2"
    except Exception:
        return "sdpa"


def build_model(device: torch.device, attn: str, layers: int | None = None):
    from transformers import SmolLM3Config, SmolLM3ForCausalLM

    cfg = SmolLM3Config(tie_word_embeddings=True)
    if layers is not None:
        cfg.num_hidden_layers = layers
        for k in ("layer_types", "no_rope_layers"):
            if getattr(cfg, k, None) is not None:
                setattr(cfg, k, getattr(cfg, k)[:layers])
    cfg._attn_implementation = attn
    cfg.use_cache = False
    torch.manual_seed(0)
    prev = torch.get_default_dtype()
    torch.set_default_dtype(torch.bfloat16)
    try:
        with torch.device(device):
            model = SmolLM3ForCausalLM(cfg)
    finally:
        torch.set_default_dtype(prev)
    return model


def _sample_lengths(g: torch.Generator, n: int, lo: int, hi: int):
    """Variable sample lengths ~ LogNormal(median 420 tokens, sigma 0.6), clipped to [lo, hi]: a
    stand-in for the tokenised-length distribution of data/qa_dataset.parquet (no tokenizer offline)."""
    z = torch.randn(n, generator=g) * 0.6 + 6.04
    return z.exp().clamp(lo, hi).long().tolist()


def synthetic_window(step: int, ga: int, B: int, T: int, pin: bool, workload: str = "full"):
    """GA micro-batches as dicts of host tensors (pinned when ``pin``).

    full   -- full-length rows: ids ~ U[0,V), labels = ids (every position trains).
    completion -- full-length rows, the first 60 % of every row is prompt (labels -100): completion-only SFT.
    padded -- the reference collator's shape: variable-length rows right-padded to T, attention_mask 0
              and labels -100 on the padding (the only mask the reference produces).
    packed -- BASELINE config 5: variable-length samples packed into rows of T tokens, flattened to
              [1, B*T] with restarting position_ids, -100 at the first token of every sample
              (SP/transformers/data/data_collator.py:1426-1428) and FA2 varlen cu_seq_lens.
    """
    g = torch.Generator().manual_seed(1234 + step)
    out = []
    for _ in range(ga):
        ids = torch.randint(0, V, (B, T), generator=g, dtype=torch.int64)
        mb = {"input_ids": ids, "labels": ids.clone()}
        if workload == "completion":
            mb["labels"][:, : int(0.6 * T)] = -100
        if workload == "padded":
            lens = _sample_lengths(g, B, 48, T)
            mask = (torch.arange(T)[None, :] < torch.tensor(lens)[:, None]).long()
            mb["attention_mask"] = mask
            mb["labels"] = torch.where(mask.bool(), ids, torch.full_like(ids, -100))
        elif workload == "packed":
            pos = torch.empty(B, T, dtype=torch.int64)
            labels = ids.clone()
            cu = [0]
            for r in range(B):
                filled = 0
                while filled < T:
                    n = min(_sample_lengths(g, 1, 48, 1024)[0], T - filled)
                    pos[r, filled:filled + n] = torch.arange(n)
                    labels[r, filled] = -100
                    filled += n
                    cu.append(r * T + filled)
            mb = {"input_ids": ids.reshape(1, B * T), "labels": labels.reshape(1, B * T),
                  "position_ids": pos.reshape(1, B * T),
                  "cu_seq_lens_q": torch.tensor(cu, dtype=torch.int32), "cu_seq_lens_k": torch.tensor(cu, dtype=torch.int32),
                  "max_length_q": int(max(b - a for a, b in zip(cu[:-1], cu[1:]))),
                  "max_length_k": int(max(b - a for a, b in zip(cu[:-1], cu[1:])))}
        if pin:
            mb = {k: (v.pin_memory() if torch.is_tensor(v) else v) for k, v in mb.items()}
        out.append(mb)
    return out


def window_tokens(window) -> tuple:
    """(trained tokens, valid-label tokens) of a window: attention_mask.sum() (all tokens when absent)."""
    trained = sum(int(mb["attention_mask"].sum()) if "attention_mask" in mb else mb["input_ids"].numel() for mb in window)
    valid = sum(int(mb["labels"].ne(-100).sum()) for mb in window)
    return trained, valid


# ------------------------------------------------------------------------------------------
# this repository's arm
# ------------------------------------------------------------------------------------------
def run_b200(args) -> dict:
    """This repository's arm (``--impl b200``) and, with ``--impl stock``, the same step on the stock GPU path the
    reference would take on this box: transformers' lm_head + ForCausalLMLoss (logits materialised), torch's
    clip_grad_norm_ + fused AdamW, DDP's built-in allreduce -- none of this library's kernels."""
    stock = args.impl == "stock"
    from llm_fine_tune_distributed_b200.comm_hook import (B200HookState, B200NvlinkState, b200_allreduce_hook,
                                                          b200_nvlink_allreduce_hook)
    from llm_fine_tune_distributed_b200.loss import FusedCELossMixin, install_fused_head
    from llm_fine_tune_distributed_b200.training import apply_freeze_policy

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from llm_fine_tune_distributed_b200 import _cabi, build as _build

    if not _cabi.LIB_PATH.exists():  # normally prebuilt by __graft_entry__.build(); one rank compiles if it is not
        if local_rank == 0:
            _build.build()
        if world > 1:
            dist.barrier()
    from llm_fine_tune_distributed_b200 import ops

    B, T, GA = args.batch, args.seq, args.ga
    attn = pick_attn() if args.attn == "auto" else args.attn
    model = build_model(dev, attn, args.layers)
    model.gradient_checkpointing_enable()
    trainable, total = apply_freeze_policy(model)
    if not stock:
        install_fused_head(model, fuse_final_norm=(args.final_norm == "fused"))  # True: must apply, else raise
    wrapped = model
    hook_state = None
    if world > 1:
        wrapped = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local_rank], bucket_cap_mb=50,
                                                            find_unused_parameters=False)
    if world > 1 and stock:
        args.comm = "ddp-default"
    elif world > 1:
        comm = os.environ.get("B200SFT_COMM", args.comm)
        hook_state = None
        if comm in ("auto", "nvlink"):
            # symmetric-memory arena + rendezvous is a collective; every rank must end up on the same hook
            try:
                hook_state = B200NvlinkState.for_model(model, None)
                ok = torch.ones((), device=dev)
            except Exception as e:  # no symmetric memory on this box / torch build
                if comm == "nvlink":
                    raise
                print(f"[rank {rank}] NVLink hook unavailable ({type(e).__name__}: {e}); using the NCCL hook", file=sys.stderr)
                ok = torch.zeros((), device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if float(ok) < 1:
                hook_state = None
        if hook_state is not None:
            args.comm = "nvlink"
            wrapped.register_comm_hook(hook_state, b200_nvlink_allreduce_hook)
        else:
            args.comm = "nccl"
            hook_state = B200HookState(None)
            wrapped.register_comm_hook(hook_state, b200_allreduce_hook)
        wrapped._b200sft_hook = True
        if args.tied_exchange == "on" or (args.tied_exchange == "auto" and args.comm == "nvlink"):
            from llm_fine_tune_distributed_b200.tied_embedding import TiedEmbeddingExchange

            wrapped._b200sft_tied_exchange = TiedEmbeddingExchange(wrapped, max_tokens=B * T)
            args.tied_exchange = "on"
        else:
            args.tied_exchange = "off"
    params = [p for p in model.parameters() if p.requires_grad]
    if stock:
        args.tail = "torch"
    if args.tail == "b200":
        from llm_fine_tune_distributed_b200.optim import B200AdamW

        opt = B200AdamW(params, lr=5e-5 * world, weight_decay=0.0)
    else:
        opt = torch.optim.AdamW(params, lr=5e-5 * world, weight_decay=0.0, fused=True)

    class Seam(FusedCELossMixin):
        num_processes = world
        model_accepts_loss_kwargs = True
        args = None

    seam = Seam()

    bwd_done = [] if (world > 1 and isinstance(hook_state, B200NvlinkState)) else None

    def stock_loss(mb, n_items) -> torch.Tensor:
        """Trainer.compute_loss's default path (SP/transformers/trainer.py:1973-2019): the model's own loss."""
        return wrapped(**mb, num_items_in_batch=n_items).loss * world

    def opt_step(window_dev, use_stock: bool = stock) -> torch.Tensor:
        """One optimizer step on device-resident micro-batches; returns the summed loss."""
        n_items = sum(mb["labels"].ne(-100).sum() for mb in window_dev)  # Trainer._get_num_items_in_batch
        if world > 1:
            dist.all_reduce(n_items)
        total_loss = torch.zeros((), device=dev)
        for i, mb in enumerate(window_dev):
            last = i == len(window_dev) - 1
            ctx = contextlib.nullcontext() if (last or world == 1) else wrapped.no_sync()
            with ctx:
                if use_stock:
                    loss = stock_loss(dict(mb), n_items)
                else:
                    loss = seam.compute_loss(wrapped, dict(mb), num_items_in_batch=n_items)
                loss.backward()
            if last and bwd_done is not None:   # backward() returns after DDP has waited for every bucket
                ev = torch.cuda.Event(enable_timing=True)
                ev.record()
                bwd_done.append(ev)
            total_loss += loss.detach()
        if use_stock and not stock:         # the in-line stock_gpu block of this library's arm: torch's own tail
            torch.nn.utils.clip_grad_norm_(params, 1.0)
            stock_opt[0].step()
            stock_opt[0].zero_grad(set_to_none=True)
        elif args.tail == "b200":
            opt.clip_and_step(1.0)          # grad-norm + clip + AdamW in this library's multi-tensor kernels
        else:
            torch.nn.utils.clip_grad_norm_(params, 1.0)
            opt.step()
        opt.zero_grad(set_to_none=True)
        return total_loss

    stock_opt = [None]

    to_dev = lambda mb: {k: (v.to(dev, non_blocking=True) if torch.is_tensor(v) else v) for k, v in mb.items()}
    token_counts = {}

    def timed(steps: int, e2e: bool, first_step: int, use_stock: bool = stock):
        # every rank draws its own data (seed offset by rank), as a sharded sampler would hand it out
        windows = [synthetic_window((first_step + s) * 64 + rank, GA, B, T, pin=e2e, workload=args.workload)
                   for s in range(steps)]
        counts = [window_tokens(w) for w in windows]
        token_counts[first_step] = (sum(c[0] for c in counts), sum(c[1] for c in counts))
        if not e2e:
            windows = [[to_dev(mb) for mb in w] for w in windows]
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        host_losses = torch.zeros(steps, dtype=torch.float32).pin_memory() if e2e else None
        t0 = time.perf_counter()
        ev0.record()
        last_loss = None
        for i, w in enumerate(windows):
            if e2e:
                w = [to_dev(mb) for mb in w]                        # H2D of the step's inputs (pinned)
            loss = opt_step(w, use_stock)
            if e2e:
                # D2H of the step's result: every step's summed loss goes to pinned host memory as soon as it exists.
                # The copy is asynchronous, as a trainer's logging is (Trainer reads tr_loss every logging_steps only,
                # training.py:266): a blocking .item() per step would drain the launch queue at every step boundary.
                # All of them are on the host, and read, before the timed region ends.
                host_losses[i].copy_(loss, non_blocking=True)
        ev1.record()
        torch.cuda.synchronize()
        if e2e:
            last_loss = float(host_losses[-1])
            assert bool(torch.isfinite(host_losses).all()), "a step's loss did not arrive on the host"
        if world > 1:
            dist.barrier()
        wall = time.perf_counter() - t0
        ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms) / steps, wall, last_loss

    # warm-up (also lets DDP rebuild its buckets after the first iteration)
    timed(max(args.warmup, 3), e2e=False, first_step=0)

    sampler = ClockSampler(local_rank)
    if rank == 0 and not args.no_clocks:
        sampler.start()
    ops.STATS.reset(with_events=True)
    torch.cuda.reset_peak_memory_stats(dev)
    if bwd_done is not None:
        hook_state.timeline = []
        hook_state.phases = []
        bwd_done.clear()
        if getattr(wrapped, "_b200sft_tied_exchange", None) is not None:
            wrapped._b200sft_tied_exchange.events = []
    ms_step, wall_dev, _ = timed(args.steps, e2e=False, first_step=100)
    comm_times = None
    if bwd_done is not None and hook_state.timeline:
        # per step: buckets in arrival order; the last one is handed over when backward compute ends, so the time
        # from that hand-over to the end of backward() is communication the step could not hide
        tl, per = hook_state.timeline, len(hook_state.timeline) // max(len(bwd_done), 1)
        tied = getattr(wrapped, "_b200sft_tied_exchange", None)
        if tied is not None and tied.events and len(tied.events) == len(bwd_done):
            # the tied bucket left early: backward compute ends when the embedding-output gradient is captured
            exposed = [tied.events[i][0].elapsed_time(bwd_done[i]) for i in range(len(bwd_done))]
        else:
            exposed = [tl[(i + 1) * per - 1][0].elapsed_time(bwd_done[i]) for i in range(len(bwd_done))]
        busy = [sum(a.elapsed_time(b) for a, b, _ in tl[i * per:(i + 1) * per]) for i in range(len(bwd_done))]
        big = max(tl, key=lambda t: t[2])
        ph = hook_state.phases
        big_ph = max(ph, key=lambda r: r[6]) if ph else None   # the reduce kernel of the largest bucket alone
        names = ("queue_ms", "copy_in_ms", "wait_for_peers_ms", "kernel_ms", "barrier_after_ms")

        def phase_ms(rows):
            return {nm: round(sum(r[i].elapsed_time(r[i + 1]) for r in rows) / max(len(bwd_done), 1), 3)
                    for i, nm in enumerate(names)}

        last_rows = [ph[(i + 1) * per - 1] for i in range(len(bwd_done))] if len(ph) == len(tl) else []
        breakdown = None
        if last_rows:
            # the same phases on every rank, max and min over ranks: a large max-min of wait_for_peers is rank skew
            # (the early ranks wait for the late ones), not transfer
            mine = phase_ms(last_rows)
            vec = torch.tensor([mine[nm] for nm in names], device=dev, dtype=torch.float64)
            hi, lo = vec.clone(), vec.clone()
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            after = [last_rows[i][5].elapsed_time(bwd_done[i]) for i in range(len(bwd_done))]
            breakdown = {"last_bucket_rank0": mine, "last_bucket_max_over_ranks": dict(zip(names, [round(v, 3) for v in hi.tolist()])),
                         "last_bucket_min_over_ranks": dict(zip(names, [round(v, 3) for v in lo.tolist()])),
                         # with the tied exchange the last bucket is done mid-backward: see comm.tied_rows instead
                         "copy_back_and_finalize_ms": None if tied is not None else round(sum(after) / len(after), 3),
                         "all_buckets_rank0": phase_ms(ph),
                         "note": "per optimizer step; phases of the hook on the side stream: queue = bucket ready -> side stream "
                                 "free; copy_in = staging into the symmetric arena; wait_for_peers = first barrier; kernel = "
                                 "reduce+scale+broadcast; barrier_after = second barrier; copy_back_and_finalize = last bucket "
                                 "done -> backward() returns (DDP copies 838 MB of results into .grad)"}
        hook_state.phases = []
        ex_mine = torch.tensor([sum(exposed) / len(exposed)], device=dev, dtype=torch.float64)
        ex_hi, ex_lo = ex_mine.clone(), ex_mine.clone()
        dist.all_reduce(ex_hi, op=dist.ReduceOp.MAX)
        dist.all_reduce(ex_lo, op=dist.ReduceOp.MIN)
        comm_times = {"exposed_ms_per_step": round(sum(exposed) / len(exposed), 3),
                      # the latest rank pays only its own copies and kernels; the others also wait for it (rank skew)
                      "exposed_ms_per_step_over_ranks": {"min": round(float(ex_lo), 3), "max": round(float(ex_hi), 3)},
                      "hook_busy_ms_per_step": round(sum(busy) / len(busy), 3),
                      "largest_bucket": {"MiB": round(big[2] / 2**20, 1), "ms": round(big[0].elapsed_time(big[1]), 3),
                                         "bus_GBps": round(2 * (world - 1) / world * big[2] / big[0].elapsed_time(big[1]) / 1e6, 1),
                                         "note": "ms = hand-over to done, i.e. including the wait for the other ranks at the barrier"}}
        if big_ph is not None:
            k_ms = big_ph[3].elapsed_time(big_ph[4])
            comm_times["largest_bucket"].update({"kernel_ms": round(k_ms, 3),
                                                 "kernel_bus_GBps": round(2 * (world - 1) / world * big_ph[6] / k_ms / 1e6, 1)})
        if breakdown:
            comm_times["breakdown"] = breakdown
        if tied is not None and tied.events:
            ev = tied.events
            seg = lambda i: round(sum(e[i].elapsed_time(e[i + 1]) for e in ev) / len(ev), 3)
            mine = {"publish_ms": seg(0), "ddp_finalize_ms": seg(1), "wait_for_peers_ms": seg(2), "row_kernels_ms": seg(3),
                    "barrier_after_ms": seg(4)}
            vec = torch.tensor(list(mine.values()), device=dev, dtype=torch.float64)
            hi, lo = vec.clone(), vec.clone()
            dist.all_reduce(hi, op=dist.ReduceOp.MAX)
            dist.all_reduce(lo, op=dist.ReduceOp.MIN)
            comm_times["tied_rows"] = {
                "rank0": mine, "max_over_ranks": dict(zip(mine, [round(v, 3) for v in hi.tolist()])),
                "min_over_ranks": dict(zip(mine, [round(v, 3) for v in lo.tolist()])),
                "note": "per optimizer step, after backward compute: publish = copy of (ids, dY) into symmetric memory; "
                        "ddp_finalize = DDP waits for every bucket and copies the results into .grad; wait_for_peers = first "
                        "barrier (rank skew: the early ranks wait here for the late ones); row_kernels = the six kernels of "
                        "b200sft_embed_grad_rows; barrier_after = second barrier"}
            comm_times["exposed_definition"] = "embedding-output gradient captured (end of backward compute) -> backward() returns"
            tied.events = None
        hook_state.timeline = None
    launches = ops.STATS.kernel_launches
    fwd_ms, bwd_ms = ops.STATS.mean_ms("flce_fwd"), ops.STATS.mean_ms("flce_bwd")
    ops.STATS.reset(with_events=False)
    if args.no_e2e:
        ms_e2e, wall_e2e, last_loss = float("nan"), float("nan"), None
    else:
        ms_e2e, wall_e2e, last_loss = timed(args.steps, e2e=True, first_step=200)
    clocks = sampler.stop() if rank == 0 else None
    peak_mem_gb = torch.cuda.max_memory_allocated(dev) / 1e9
    stock_gpu = None
    if world == 1 and not stock and not args.no_stock_gpu:
        # the same step on the stock GPU path (what training.py would run on this box without this library):
        # same model object, weights, data and GA; transformers' lm_head + ForCausalLMLoss, torch's clip + fused AdamW
        stock_opt[0] = torch.optim.AdamW(params, lr=5e-5 * world, weight_decay=0.0, fused=True)
        torch.cuda.reset_peak_memory_stats(dev)
        timed(1, e2e=False, first_step=300, use_stock=True)
        ms_stock, _, _ = timed(3, e2e=False, first_step=301, use_stock=True)
        stock_gpu = {"ms_per_step": round(ms_stock, 3),
                     "value": round(token_counts[301][0] / 3 / (ms_stock / 1e3), 1), "unit": UNIT, "steps": 3, "warmup": 1,
                     "max_memory_allocated_GB": round(torch.cuda.max_memory_allocated(dev) / 1e9, 2),
                     "path": "stock transformers lm_head + ForCausalLMLoss (logits materialised), torch clip_grad_norm_ + "
                             "fused AdamW; same model, data, GA; peak memory includes this arm's 1.7 GB of AdamW state"}
        stock_opt[0] = None

    def all_ranks(x: float) -> float:
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t)
        return float(t)

    trained_dev, valid_dev = (all_ranks(v) for v in token_counts[100])
    per_rank = torch.tensor([token_counts[100][1] / args.steps], device=dev, dtype=torch.float64)  # valid-label tokens / step
    rank_hi, rank_lo = per_rank.clone(), per_rank.clone()
    if world > 1:
        dist.all_reduce(rank_hi, op=dist.ReduceOp.MAX)
        dist.all_reduce(rank_lo, op=dist.ReduceOp.MIN)
    if args.kernel_table and rank == 0:
        from torch.profiler import ProfilerActivity, profile

        win = [to_dev(mb) for mb in synthetic_window(999 * 64 + rank, GA, B, T, pin=False, workload=args.workload)]
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
            opt_step(win)
            torch.cuda.synchronize()
        rows = [(e.key, e.count, e.device_time_total) for e in prof.key_averages() if e.device_time_total > 0
                and e.device_type == torch.autograd.DeviceType.CUDA]
        rows.sort(key=lambda r: -r[2])
        tot = sum(r[2] for r in rows) or 1.0
        ours = sum(r[2] for r in rows if "b200::" in r[0])
        with open(args.kernel_table, "w") as f:
            f.write("# Kernel device-time table of one bench.py optimizer step (torch.profiler / CUPTI)\n\n"
                    f"workload {args.workload}, B={B} T={T} GA={GA}, {model.config.num_hidden_layers} layers; "
                    f"total kernel time {tot / 1e3:.2f} ms; this library's kernels {ours / 1e3:.2f} ms "
                    f"({100 * ours / tot:.1f} %)\n\n| kernel | launches | total ms | share |\n|---|---:|---:|---:|\n")
            for k, n, t in rows[:45]:
                f.write(f"| `{k[:110]}` | {n} | {t / 1e3:.3f} | {100 * t / tot:.1f}% |\n")
    tokens_per_step = trained_dev / args.steps
    value = tokens_per_step / (ms_step / 1e3)
    if args.no_e2e:
        e2e_value = float("nan")
    else:
        e2e_value = all_ranks(token_counts[200][0]) / args.steps / (ms_e2e / 1e3)
    N = B * T
    peaks = read_peaks()
    fwd_flops = 2.0 * N * H * V
    achieved = fwd_flops / (fwd_ms / 1e3) / 1e12 if not stock else float("nan")
    traffic = tensor_pct = None
    tp = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tp):
        prof = json.load(open(tp))
        traffic = prof.get("flce_fwd_dram_bytes_per_launch")
        tensor_pct = prof.get("flce_kernels_tensor_pipe_active_pct")
    probe = synthetic_window(0, GA, B, T, pin=False, workload=args.workload)
    h2d_bytes = sum(v.numel() * v.element_size() for mb in probe for v in mb.values() if torch.is_tensor(v))
    res = {
        "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(ms_step, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {
            "workload": f"SmolLM3-3B random-init bf16 SFT optimizer step: seq {T}, per-device batch {B}, grad-accum {GA}, "
                        "freeze policy last-2-layers+tied lm_head, gradient checkpointing, "
                        + ("stock lm_head + ForCausalLMLoss (logits materialised), " if stock else "fused LM-head+CE (sm_100a), ")
                        + f"clip 1.0 + AdamW ({args.tail} kernels)"
                        + (f", DDP bucket 50 MiB + {'built-in allreduce' if stock else f'b200 comm hook ({args.comm})'}" if world > 1 else ""),
            "global_batch": B * GA * world, "seq_len": T, "parallelism": f"dp{world}", "attn": attn,
            "rows": {"full": "full-length rows, every position trains",
                     "completion": "full-length rows, first 60 % of each row is prompt (labels -100, completion-only loss)",
                     "padded": "variable-length rows right-padded to seq_len (reference collator), -100 on padding",
                     "packed": "variable-length samples packed to seq_len, flattened [1, B*T], restarting position_ids, "
                               "-100 at sample starts, FA2 varlen"}[args.workload],
            "tokens_per_step": round(tokens_per_step, 1), "valid_label_tokens_per_step": round(valid_dev / args.steps, 1),
            # load balance of the loss work across ranks (valid-label tokens per optimizer step on one rank)
            "valid_tokens_per_rank": {"min": round(float(rank_lo), 1), "max": round(float(rank_hi), 1),
                                      "imbalance": round(float(rank_hi) / max(float(rank_lo), 1.0) - 1.0, 5)},
            "skip_ignored_rows": not stock, "final_norm": "stock" if stock else args.final_norm,
            "layers": model.config.num_hidden_layers, "trainable_params": trainable, "total_params": total,
            "l2": "working set (6.2 GB of weights + activations per pass) far exceeds the 126 MB L2; no explicit flush",
            "attention_mask": "None (full-length rows; identical to an all-ones mask)",
        },
        "clocks": clocks,
        "e2e": {"value": round(e2e_value, 1), "unit": UNIT, "ms_per_step": round(ms_e2e, 3),
                "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": 4,
                "note": "every step: pinned int64 inputs copied H2D, the step's summed loss copied D2H into pinned memory "
                        "(asynchronously); all losses are on the host and checked before the timed region ends",
                "last_loss": last_loss},
        "gpu_launches": launches,
        "roofline": {
            "kernel": "gemm_tcgen05_2cta_kernel<FwdSoftmaxEpi> (+ target-logit memset, row merge and loss reduce: the whole "
                      f"b200sft_flce_fwd call), N={N} x H={H} x V={V}",
            "bound": "tensor", "achieved": round(achieved, 1), "peak": peaks["sustained"], "unit": "TFLOP/s",
            "frac": round(achieved / peaks["sustained"], 4), "traffic": traffic,
            "peak_kind": "bf16 sustained, " + peaks["kind"], "algorithmic_flops_per_launch": fwd_flops,
            "avg_launch_ms": round(fwd_ms, 4),
            "frac_of_burst_peak": round(achieved / peaks["burst"], 4),
            "ncu_tensor_pipe_active_pct": tensor_pct,  # from the committed ncu --set full capture, not measured live
            "note": "timed inside the training step, so the denominator is the sustained cuBLAS figure; the step is not "
                    "tensor-bound end to end, the SMs hold burst clocks (see clocks.sm_mhz) and the kernel runs above it",
        },
        "flce_bwd": {"avg_call_ms": round(bwd_ms, 4), "algorithmic_flops_per_call": 4.0 * N * H * V,
                     "achieved_tflops": round(4.0 * N * H * V / (bwd_ms / 1e3) / 1e12, 1),
                     "frac_of_sustained_peak": round(4.0 * N * H * V / (bwd_ms / 1e3) / 1e12 / peaks["sustained"], 4)},
        "wall_s": {"device_timed": round(wall_dev, 3), "e2e_timed": round(wall_e2e, 3)},
        "max_memory_allocated_GB": round(peak_mem_gb, 2),
    }
    if stock_gpu is not None:
        res["stock_gpu"] = stock_gpu
    if stock:  # none of this library's kernels ran: no roofline of ours to report
        res["impl"] = "stock"
        for k in ("roofline", "flce_bwd"):
            res.pop(k)
        res["gpu_launches"] = 0
    if hook_state is not None:
        n_steps = args.steps * (1 if args.no_e2e else 2) + max(args.warmup, 3)
        res["comm"] = {"hook": args.comm, "path": getattr(hook_state, "path_used", "nccl"),
                       "tied_exchange": args.tied_exchange,
                       "buckets_per_step": round(hook_state.buckets_seen / max(n_steps, 1), 2),
                       "bytes_reduced_per_step": round(hook_state.bytes_reduced / max(n_steps, 1))}
        if comm_times:
            res["comm"].update(comm_times)
    if world > 1:
        dist.destroy_process_group()   # the other ranks leave; rank 0 alone times the CPU baseline on the free cores
    if rank == 0 and not args.no_cpu_baseline and not stock:
        del opt, wrapped, seam
        torch.cuda.empty_cache()
        res["cpu_baseline"] = cpu_reference_step(sample_tokens=args.cpu_tokens, steps=1, warmup=1)
    return res if rank == 0 else {}


# ------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the reference's own arithmetic on the host cores
# ------------------------------------------------------------------------------------------
def host_threads() -> int:
    """Cores this process may run on.  torchrun exports OMP_NUM_THREADS=1 to every rank, which would leave the CPU
    arm on one core at N > 1; the arm runs on rank 0 alone, so it takes every core of the affinity mask."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:  # not Linux
        return max(1, os.cpu_count() or 1)


def cpu_reference_step(sample_tokens: int, steps: int, warmup: int = 1, budget_s: float = 240.0) -> dict:
    """Optimizer steps of the reference path on CPU: stock SmolLM3ForCausalLM forward with labels (lm_head ->
    ForCausalLMLoss inside transformers), autograd backward, clip, AdamW (training.py:300's step) -- each on a
    bounded sample of B=1 x ``sample_tokens`` tokens of the same 3B architecture, on all host threads.
    ``warmup`` untimed steps of the same size (after one 16-token pass that spins up the thread pool), then up to
    ``steps`` timed ones; stops early once ``budget_s`` seconds of timed work are spent (reported as steps_run)."""
    from llm_fine_tune_distributed_b200.training import apply_freeze_policy

    threads = host_threads()
    torch.set_num_threads(threads)
    t_build = time.perf_counter()
    if torch.cuda.is_available():  # initialise on the GPU (fast), then move: initialisation is not timed
        model = build_model(torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))), "sdpa").cpu()
        torch.cuda.empty_cache()
    else:
        model = build_model(torch.device("cpu"), "sdpa")
    model.gradient_checkpointing_enable()
    apply_freeze_policy(model)
    model.train()
    params = [p for p in model.parameters() if p.requires_grad]
    opt = torch.optim.AdamW(params, lr=5e-5)
    t_build = time.perf_counter() - t_build
    times = []
    spent = 0.0
    for s in range(1 + warmup + steps):
        T = 16 if s == 0 else sample_tokens
        ids = torch.randint(0, V, (1, T), generator=torch.Generator().manual_seed(99 + s))
        n_items = ids.ne(-100).sum()
        t0 = time.perf_counter()
        out = model(input_ids=ids, labels=ids, num_items_in_batch=n_items)
        out.loss.backward()
        torch.nn.utils.clip_grad_norm_(params, 1.0)
        opt.step()
        opt.zero_grad(set_to_none=True)
        dt = time.perf_counter() - t0
        if s > warmup:
            times.append(dt)
        if s > 0:
            spent += dt
        if times and spent > budget_s:
            break
    sec = sum(times) / len(times)
    return {"value": round(sample_tokens / sec, 2), "unit": UNIT, "cores": threads, "kind": "port",
            "os_cpu_count": os.cpu_count(), "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS"),
            "torch_num_threads": torch.get_num_threads(), "seconds_per_sample_step": round(sec, 3),
            "steps_run": len(times), "warmup_run": warmup,
            "sample": f"{len(times)} optimizer step(s) on B=1 x T={sample_tokens} tokens of the same SmolLM3-3B workload "
                      "(stock transformers lm_head + ForCausalLMLoss + autograd + AdamW, sdpa attention, torch CPU "
                      f"kernels, {threads} threads); model build {t_build:.1f}s not timed"}


def run_reference(args) -> dict:
    """``--impl reference``: under torchrun rank 0 alone runs and prints; the other ranks exit 0 without work."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return {}
    cb = cpu_reference_step(sample_tokens=args.cpu_tokens, steps=max(1, args.steps), warmup=max(1, args.warmup))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    return {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": world,
        "steps": cb["steps_run"], "steps_requested": args.steps, "warmup": cb["warmup_run"],
        "warmup_requested": args.warmup, "ms_per_step": round(cb["seconds_per_sample_step"] * 1e3, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "SmolLM3-3B random-init bf16 SFT optimizer step, reference path on host CPU cores "
                               f"(bounded sample B=1 x T={args.cpu_tokens} per step; rank 0 only, {cb['cores']} threads)",
                   "global_batch": 1, "seq_len": args.cpu_tokens, "parallelism": "cpu"},
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "stock"],
                    help="b200 = this repository; reference = the reference step on the host CPU cores; stock = the same "
                         "step on the GPU through stock transformers / torch / DDP (what training.py runs without this library)")
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--seq", type=int, default=1024)
    ap.add_argument("--ga", type=int, default=2)
    ap.add_argument("--layers", type=int, default=None, help="debug only: fewer decoder layers (invalid as a result)")
    ap.add_argument("--attn", default="auto")
    ap.add_argument("--tail", default="b200", choices=["b200", "torch"],
                    help="step tail: this library's fused grad-norm clip + AdamW kernels, or torch's clip_grad_norm_ + fused AdamW")
    ap.add_argument("--final-norm", default="fused", choices=["stock", "fused"],
                    help="fused = the decoder's final RMSNorm runs inside the head's row gather (scored rows only)")
    ap.add_argument("--comm", default="auto", choices=["auto", "nccl", "nvlink"],
                    help="N>1: gradient hook.  nvlink = this library's fused NVLink reduce+scale kernel over symmetric "
                         "memory; nccl = scale kernel + NCCL allreduce; auto (default) = nvlink when symmetric memory "
                         "can be set up on every rank, else nccl.  Env B200SFT_COMM overrides.")
    ap.add_argument("--tied-exchange", default="auto", choices=["auto", "on", "off"],
                    help="N>1: split the tied lm_head/embed_tokens gradient so its 521 MiB bucket is reduced at the start of "
                         "backward and only the embedding's rows are exchanged at the end (auto = on with the NVLink hook)")
    ap.add_argument("--workload", default="full", choices=["full", "completion", "padded", "packed"],
                    help="full = headline (configs 1-3); packed with --batch 2 --seq 4096 = BASELINE config 5")
    ap.add_argument("--cpu-tokens", type=int, default=1024,
                    help="CPU baseline / reference arm: tokens in the bounded sample (one row of the workload by default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-stock-gpu", action="store_true", help="skip the in-line stock-GPU-path block (N=1 only)")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs only: skip the host-buffer pass")
    ap.add_argument("--no-clocks", action="store_true", help="profiling runs only: do not spawn the nvidia-smi sampler")
    ap.add_argument("--kernel-table", default=None, metavar="PATH",
                    help="after the timed passes, trace ONE more step with torch.profiler (CUPTI) and write the "
                         "per-kernel device-time table (markdown) to PATH")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus != world and world == 1 and args.gpus > 1:
        print(f"bench.py: --gpus {args.gpus} needs torchrun (WORLD_SIZE={world}); see the module docstring", file=sys.stderr)
        sys.exit(2)
    res = run_reference(args) if args.impl == "reference" else run_b200(args)
    if res:
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
