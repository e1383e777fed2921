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
    args = ap.parse_args()
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
    for N in args.tokens:
        gx = torch.Generator(device="cpu").manual_seed(0)
        x = torch.randn(N, H, generator=gx).to(torch.bfloat16).to(dev)
        for m in args.masked:
            gl = torch.Generator(device="cpu").manual_seed(2)
            labels = torch.randint(0, V, (N,), generator=gl)
            labels[torch.rand(N, generator=gl) < m] = -100
            labels = labels.to(dev)
            lab32 = labels.to(torch.int32)
            n_items = torch.tensor(max(int((labels != -100).sum()), 1), dtype=torch.int64, device=dev)
            g = torch.ones((), device=dev)

            torch.cuda.reset_peak_memory_stats()
            base_mem = torch.cuda.memory_allocated()
            out = {}

            if args.skip_ignored:
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

            t_f = time_ms(fwd, args.iters, args.warmup, flush)
            t_b = time_ms(bwd, args.iters, args.warmup, flush)
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
            rec = {
                "N": N, "H": H, "V": V, "masked": m, "chunk_cols": args.chunk_cols, "skip_ignored_rows": args.skip_ignored,
                "fused_fwd_ms": round(t_f, 4), "fused_bwd_ms": round(t_b, 4),
                "fused_fwd_tflops": round(fl_f / t_f / 1e9, 1),
                "fused_fwd_bwd_tflops": round(fl_fb / (t_f + t_b) / 1e9, 1),
                "fused_fwd_frac_of_peak": round(fl_f / t_f / 1e9 / peak_tf, 4),
                "fused_fwd_bwd_frac_of_peak": round(fl_fb / (t_f + t_b) / 1e9 / peak_tf, 4),
                "peak_tflops": peak_tf, "peak_kind": peak_kind,
                "fused_peak_mem_gb": round(fused_mem / 2**30, 3),
                "fused_loss": float(out["f"][0]),
            }
            del out
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

                    t_tf = time_ms(tfwd, args.iters, args.warmup, flush)
                    t_tfb = time_ms(tfb, args.iters, args.warmup, flush)
                    rec.update({
                        "torch_fwd_ms": round(t_tf, 4), "torch_fwd_bwd_ms": round(t_tfb, 4),
                        "torch_peak_mem_gb": round((torch.cuda.max_memory_allocated() - base_mem) / 2**30, 3),
                        "torch_loss": float(st["loss"]),
                        "speedup_fwd": round(t_tf / t_f, 3), "speedup_fwd_bwd": round(t_tfb / (t_f + t_b), 3),
                    })
                    del st, xr, wr
                except torch.cuda.OutOfMemoryError:
                    rec["torch"] = "OOM"
                torch.cuda.empty_cache()
            print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
