"""LM-head + masked-CE microbenchmark (BASELINE.json config 4): N tokens x hidden 2048 x vocab
128256, masked fraction m, fused CUDA path vs torch ``F.linear -> .float() -> F.cross_entropy``.

Prints one JSON line per (N, masked) point: ms and TFLOP/s (algorithmic 2NHV fwd / 6NHV fwd+bwd),
fraction of the measured bf16 peak, peak memory of both paths.  CUDA events, L2 flushed between
iterations by streaming a 512 MiB buffer.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_fine_tune_distributed_b200 import ops  # noqa: E402


def peaks():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["bf16_tflops"], d["hbm_gbs"], "measured"
    return 1590.0, 6650.0, "fallback"


def sustained_peak():
    """cuBLAS bf16 throughput under sustained load (the denominator for cases that run long enough to sit at the power
    cap, i.e. every case of a 20-iteration sweep); falls back to the burst figure."""
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops_sustained", d["bf16_tflops"])
    return 1400.0


def time_ms(fn, iters, warmup, flush):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    total = 0.0
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        total += s.elapsed_time(e)
    return total / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, nargs="+", default=[8192])
    ap.add_argument("--masked", type=float, nargs="+", default=[0.0])
    ap.add_argument("--hidden", type=int, default=2048)
    ap.add_argument("--vocab", type=int, default=128256)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--chunk-cols", type=int, default=0)
    ap.add_argument("--no-torch", action="store_true")
    ap.add_argument("--cpu-tokens", type=int, default=0,
                    help="also time the reference head + CE forward/backward on the host cores at this many tokens "
                         "(the same torch path as the GPU comparison: F.linear -> .float() -> F.cross_entropy + autograd, bf16)")
    ap.add_argument("--kernel-times", action="store_true", help="also print per-kernel device times (torch.profiler)")
    ap.add_argument("--skip-ignored", action="store_true", help="compact away ignore_index rows first (valid-row compaction)")
    ap.add_argument("--modes", nargs="+", default=None, choices=["dense", "compact"],
                    help="run these modes for every case (overrides --skip-ignored); the torch comparison runs once per case")
    ap.add_argument("--patterns", nargs="+", default=["scatter"], choices=["scatter", "prefix"],
                    help="where the masked labels sit: scattered at random, or the first m*1024 positions of every "
                         "1024-token sequence (completion-only SFT)")
    ap.add_argument("--clocks", action="store_true", help="sample nvidia-smi clocks / throttle reasons during each case")
    args = ap.parse_args()
    modes = args.modes or (["compact"] if args.skip_ignored else ["dense"])
    dev = torch.device("cuda:0")
    peak_tf, _, peak_kind = peaks()
    H, V = args.hidden, args.vocab
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    gw = torch.Generator(device="cpu").manual_seed(1)
    w = (torch.randn(V, H, generator=gw) * 0.02).to(torch.bfloat16).to(dev)
    if args.cpu_tokens > 0:
        import time

        def cpu_head_ce(xh, wh, lh, n_items):  # the torch comparison path below, on the host cores
            xh, wh = xh.clone().requires_grad_(True), wh.clone().requires_grad_(True)
            logits = torch.nn.functional.linear(xh, wh).float()
            (torch.nn.functional.cross_entropy(logits, lh, ignore_index=-100, reduction="sum") / n_items).backward()

        n = args.cpu_tokens
        xc = torch.randn(n, H, generator=torch.Generator().manual_seed(0)).to(torch.bfloat16)
        wc = w.cpu()
        lc = torch.randint(0, V, (n,), generator=torch.Generator().manual_seed(2))
        cpu_head_ce(xc[:64], wc, lc[:64], 64.0)      # warm the thread pool
        t0 = time.perf_counter()
        cpu_head_ce(xc, wc, lc, float(n))
        dt = time.perf_counter() - t0
        print(json.dumps({"cpu_reference_head_ce": {"tokens": n, "seconds_fwd_bwd": round(dt, 3),
                                                    "tflops_6NHV": round(6.0 * n * H * V / dt / 1e12, 3),
                                                    "threads": torch.get_num_threads(), "os_cpu_count": os.cpu_count()}}), flush=True)
    if args.clocks:
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        from bench import ClockSampler
    for N in args.tokens:
        gx = torch.Generator(device="cpu").manual_seed(0)
        x = torch.randn(N, H, generator=gx).to(torch.bfloat16).to(dev)
        for pattern in args.patterns:
          for m in args.masked:
            gl = torch.Generator(device="cpu").manual_seed(2)
            labels = torch.randint(0, V, (N,), generator=gl)
            if pattern == "scatter":
                labels[torch.rand(N, generator=gl) < m] = -100
            else:
                T = 1024 if N % 1024 == 0 else N
                labels.view(-1, T)[:, : int(round(m * T))] = -100
            labels = labels.to(dev)
            lab32 = labels.to(torch.int32)
            n_items = torch.tensor(max(int((labels != -100).sum()), 1), dtype=torch.int64, device=dev)
            g = torch.ones((), device=dev)
            def sampled(fn_):
                """Run fn_ under its own nvidia-smi sampler: the part is power-capped, so what a case measures depends on
                what ran just before it; every mode carries the clocks it actually got."""
                if not args.clocks:
                    return fn_(), None
                sm = ClockSampler(0, interval_ms=50)
                sm.start()
                val = fn_()
                return val, sm.stop()

            recs = []
            for mode in modes:
                skip = mode == "compact"
                torch.cuda.reset_peak_memory_stats()
                base_mem = torch.cuda.memory_allocated()
                out = {}
                if skip:
                    def fwd():
                        xc, lc, _ri, pos, nl = ops.compact_rows(x, lab32)
                        out["c"] = (xc, lc, pos, nl)
                        out["f"] = ops.flce_forward(xc, w, lc, denom=n_items, n_rows=nl)

                    def bwd():
                        xc, lc, pos, nl = out["c"]
                        dxc, dw = ops.flce_backward(xc, w, lc, out["f"][1], g, denom=n_items, chunk_cols=args.chunk_cols,
                                                    n_rows=nl)
                        out["b"] = (ops.scatter_rows(dxc, pos), dw)
                else:
                    def fwd():
                        out["f"] = ops.flce_forward(x, w, lab32, denom=n_items)

                    def bwd():
                        out["b"] = ops.flce_backward(x, w, lab32, out["f"][1], g, denom=n_items, chunk_cols=args.chunk_cols)

                (t_f, t_b), mode_clocks = sampled(lambda: (time_ms(fwd, args.iters, args.warmup, flush),
                                                           time_ms(bwd, args.iters, args.warmup, flush)))
                fused_mem = torch.cuda.max_memory_allocated() - base_mem
                if args.kernel_times:
                    from torch.profiler import ProfilerActivity, profile

                    with profile(activities=[ProfilerActivity.CUDA]) as prof:
                        for _ in range(3):
                            fwd()
                            bwd()
                        torch.cuda.synchronize()
                    for e in sorted(prof.key_averages(), key=lambda e: -e.device_time_total):
                        if e.device_time_total > 0:
                            print(f"   {e.device_time_total / e.count / 1e3:8.3f} ms x{e.count:3d}  {e.key[:100]}")
                fl_f, fl_fb = 2.0 * N * H * V, 6.0 * N * H * V
                recs.append({
                    "N": N, "H": H, "V": V, "masked": m, "pattern": pattern, "valid_rows": int((labels != -100).sum()),
                    "chunk_cols": args.chunk_cols, "skip_ignored_rows": skip,
                    "fused_fwd_ms": round(t_f, 4), "fused_bwd_ms": round(t_b, 4),
                    "fused_fwd_tflops": round(fl_f / t_f / 1e9, 1),
                    "fused_fwd_bwd_tflops": round(fl_fb / (t_f + t_b) / 1e9, 1),
                    "fused_fwd_frac_of_peak": round(fl_f / t_f / 1e9 / peak_tf, 4),
                    "fused_fwd_bwd_frac_of_peak": round(fl_fb / (t_f + t_b) / 1e9 / peak_tf, 4),
                    "peak_tflops": peak_tf, "peak_kind": peak_kind,
                    "sustained_peak_tflops": sustained_peak(),
                    "fused_fwd_frac_of_sustained_peak": round(fl_f / t_f / 1e9 / sustained_peak(), 4),
                    "fused_fwd_bwd_frac_of_sustained_peak": round(fl_fb / (t_f + t_b) / 1e9 / sustained_peak(), 4),
                    "fused_peak_mem_gb": round(fused_mem / 2**30, 3),
                    "fused_loss": float(out["f"][0]),
                })
                if mode_clocks is not None:
                    recs[-1]["clocks"] = mode_clocks
                del out
            torch_rec = {}
            if not args.no_torch:
                try:
                    torch.cuda.reset_peak_memory_stats()
                    xr = x.clone().requires_grad_(True)
                    wr = w.clone().requires_grad_(True)
                    base_mem = torch.cuda.memory_allocated()
                    st = {}

                    def tfwd():
                        st.pop("loss", None)  # free the previous graph before building the next
                        logits = torch.nn.functional.linear(xr, wr).float()
                        st["loss"] = torch.nn.functional.cross_entropy(logits, labels, ignore_index=-100,
                                                                       reduction="sum") / n_items

                    def tfb():
                        tfwd()
                        xr.grad = None
                        wr.grad = None
                        st["loss"].backward()

                    (t_tf, t_tfb), torch_clocks = sampled(lambda: (time_ms(tfwd, args.iters, args.warmup, flush),
                                                                   time_ms(tfb, args.iters, args.warmup, flush)))
                    torch_rec = {"torch_fwd_ms": round(t_tf, 4), "torch_fwd_bwd_ms": round(t_tfb, 4), "torch_clocks": torch_clocks,
                                 "torch_peak_mem_gb": round((torch.cuda.max_memory_allocated() - base_mem) / 2**30, 3),
                                 "torch_loss": float(st["loss"])}
                    del st, xr, wr
                except torch.cuda.OutOfMemoryError:
                    torch_rec = {"torch": "OOM"}
                torch.cuda.empty_cache()
            for rec in recs:
                rec.update(torch_rec)
                if "torch_fwd_ms" in rec:
                    rec["speedup_fwd"] = round(rec["torch_fwd_ms"] / rec["fused_fwd_ms"], 3)
                    rec["speedup_fwd_bwd"] = round(rec["torch_fwd_bwd_ms"] / (rec["fused_fwd_ms"] + rec["fused_bwd_ms"]), 3)
                print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
