"""Final RMSNorm inside the head's row gather (b200sft_rmsnorm_gather / _scatter_bwd) against what it
replaces: the stock six-pass RMSNorm over all rows (modeling_smollm3.py:271-276, torch autograd) followed
by gather_rows, and scatter_rows followed by the norm's autograd backward.  CUDA-event timing.

    python benchmarks/rmsnorm_bench.py >> profiles/r01_rmsnorm_microbench.jsonl
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_fine_tune_distributed_b200 import ops  # noqa: E402

HBM_GBPS = 6554.6  # MEASURED_PEAKS.json


def stock_norm(h, g, eps):
    x = h.to(torch.float32)
    x = x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps)
    return g * x.to(h.dtype)


def timed(fn, iters=50, warmup=10):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3  # us


def main():
    dev = torch.device("cuda:0")
    N, H, eps = 8192, 2048, 1e-6
    g = (1 + 0.1 * torch.randn(H, device=dev)).to(torch.bfloat16)
    for masked in (0.0, 0.5, 0.9):
        h = torch.randn(N, H, device=dev).to(torch.bfloat16)
        labels = torch.randint(0, 1000, (N,), device=dev, dtype=torch.int32)
        labels[torch.rand(N, device=dev) < masked] = -100
        plan = ops.compaction_list(labels)
        live = int(plan.n_live)
        dyc = (torch.randn(N, H, device=dev) * 0.01).to(torch.bfloat16)
        _, rstd = ops.rmsnorm_gather(h, g, eps, plan)

        fused_fwd = timed(lambda: ops.rmsnorm_gather(h, g, eps, plan))
        fused_bwd = timed(lambda: ops.rmsnorm_scatter_bwd(dyc, h, g, rstd, plan, True, False))
        fused_bwd_dg = timed(lambda: ops.rmsnorm_scatter_bwd(dyc, h, g, rstd, plan, True, True))
        stock_fwd = timed(lambda: ops.gather_rows(stock_norm(h, g, eps), plan))

        hh = h.clone().requires_grad_(True)

        def stock_fb():
            hh.grad = None
            y = stock_norm(hh, g, eps)
            y.backward(ops.scatter_rows(dyc, plan.pos))

        stock_fwd_bwd = timed(stock_fb)
        # algorithmic bytes: read live rows + write live rows (fwd); read dy + h live rows, write all dh rows (bwd)
        fwd_bytes = 2 * live * H * 2
        bwd_bytes = 2 * live * H * 2 + N * H * 2
        print(json.dumps({"bench": "rmsnorm_gather", "N": N, "H": H, "masked": masked, "live_rows": live,
                          "fused_fwd_us": round(fused_fwd, 1), "fused_fwd_GBps": round(fwd_bytes / fused_fwd / 1e3, 0),
                          "fused_bwd_us": round(fused_bwd, 1), "fused_bwd_GBps": round(bwd_bytes / fused_bwd / 1e3, 0),
                          "fused_bwd_with_dgamma_us": round(fused_bwd_dg, 1),
                          "stock_norm_plus_gather_fwd_us": round(stock_fwd, 1),
                          "stock_fwd_plus_scatter_plus_autograd_bwd_us": round(stock_fwd_bwd, 1),
                          "fused_fwd_plus_bwd_us": round(fused_fwd + fused_bwd, 1),
                          "hbm_peak_GBps": HBM_GBPS}), flush=True)


if __name__ == "__main__":
    main()
