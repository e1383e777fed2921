"""Label front-end: one launch (b200sft_label_frontend) against the two-step path it replaces inside
the fused loss (b200sft_label_prep = memset + kernel, then compact_labels inside b200sft_compact_rows).
CUDA-event timing over many back-to-back calls (launch-latency regime: N*8 B = 64 KiB at 8 x 1024).

    python benchmarks/label_frontend_bench.py >> profiles/r01_label_frontend_microbench.jsonl
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_fine_tune_distributed_b200 import _cabi, ops  # noqa: E402


def timed(fn, iters=500, warmup=50):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3  # microseconds


def main():
    dev = torch.device("cuda:0")
    lib = _cabi.load()
    for B, T, p in [(8, 1024, 0.5), (1, 8192, 0.5), (8, 8192, 0.9)]:
        N = B * T
        labels = torch.randint(0, 128256, (B, T), device=dev)
        labels[torch.rand(B, T, device=dev) < p] = -100
        shifted = torch.empty(N, dtype=torch.int32, device=dev)
        counts = torch.empty(2, dtype=torch.int64, device=dev)
        ri, lc, pos = (torch.empty(N, dtype=torch.int32, device=dev) for _ in range(3))
        n_live = torch.empty((), dtype=torch.int32, device=dev)
        x = torch.empty(N, 64, dtype=torch.bfloat16, device=dev)   # narrow rows: the gather is not what is compared
        xc = torch.empty_like(x)
        st = torch.cuda.current_stream().cuda_stream
        P = lambda t: t.data_ptr()

        def one_launch():
            lib.b200sft_label_frontend(None, P(labels), None, None, None, B, T, -100, None, P(shifted), P(counts),
                                       P(ri), P(lc), P(pos), P(n_live), st)
            lib.b200sft_gather_rows(P(x), P(ri), P(n_live), N, 64, P(xc), st)

        def two_step():
            lib.b200sft_label_prep(P(labels), B, T, -100, P(shifted), P(counts), st)
            lib.b200sft_compact_rows(P(x), P(shifted), N, 64, -100, P(xc), P(lc), P(ri), P(pos), P(n_live), st)

        print(json.dumps({"bench": "label_frontend", "B": B, "T": T, "masked": p,
                          "frontend_plus_gather_us": round(timed(one_launch), 2),
                          "label_prep_plus_compact_rows_us": round(timed(two_step), 2),
                          "through_ops_frontend_us": round(timed(lambda: ops.label_frontend(labels=labels)), 2),
                          "through_ops_label_prep_us": round(timed(lambda: ops.label_prep(labels)), 2)}), flush=True)


if __name__ == "__main__":
    main()
