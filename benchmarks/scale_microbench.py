"""Gradient-bucket scale / cast kernels vs the HBM roofline (MEASURED_PEAKS.json hbm_gbs).

Bucket sizes are DDP's for the reference freeze policy (~43/86/63/86/521 MiB bf16)
plus the whole 837.8 MB trainable set.  Buckets far exceed the 126 MB L2 except the smallest, for
which a 512 MiB flush buffer is streamed between iterations.  One JSON line per case, with torch's
own ``div_`` / ``.to(bf16).div_`` on the same buffers beside it.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_fine_tune_distributed_b200 import ops  # noqa: E402


def timeit(fn, flush, iters=10, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot / iters


def main():
    dev = torch.device("cuda:0")
    peak = 6650.0
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p))["hbm_gbs"]
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    for mib in (43, 86, 521, 799):
        n = (mib << 20) // 2
        buf = torch.randn(n, device=dev, dtype=torch.float32).to(torch.bfloat16)
        t = timeit(lambda: ops.scale_(buf, 0.125), flush)
        tt = timeit(lambda: buf.div_(8), flush)
        print(json.dumps({"kernel": "scale_bf16_inplace", "elements": n, "MiB": mib, "ms": round(t, 4),
                          "GBps": round(4 * n / t / 1e6, 1), "frac_of_hbm_peak": round(4 * n / t / 1e6 / peak, 4),
                          "torch_div_ms": round(tt, 4), "hbm_peak_GBps": peak}), flush=True)
        del buf
    n = 200 << 20
    src = torch.randn(n, device=dev)
    dst = torch.empty(n, device=dev, dtype=torch.bfloat16)
    t = timeit(lambda: ops.scale_cast_f32_to_bf16(src, 0.125, out=dst), flush)
    tt = timeit(lambda: src.to(torch.bfloat16).div_(8), flush)
    print(json.dumps({"kernel": "scale_cast_f32_to_bf16", "elements": n, "ms": round(t, 4), "GBps": round(6 * n / t / 1e6, 1),
                      "frac_of_hbm_peak": round(6 * n / t / 1e6 / peak, 4), "torch_ms": round(tt, 4)}), flush=True)
    t = timeit(lambda: ops.cast_bf16_to_f32(dst, src), flush)
    tt = timeit(lambda: src.copy_(dst), flush)
    print(json.dumps({"kernel": "cast_bf16_to_f32", "elements": n, "ms": round(t, 4), "GBps": round(6 * n / t / 1e6, 1),
                      "frac_of_hbm_peak": round(6 * n / t / 1e6 / peak, 4), "torch_ms": round(tt, 4)}), flush=True)


if __name__ == "__main__":
    main()
