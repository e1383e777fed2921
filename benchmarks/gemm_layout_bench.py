"""Throughput of the tcgen05 GEMM engine per operand layout (K-major vs MN-major) at the shapes
the backward uses, next to torch.matmul (cuBLAS) on the same operands."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_fine_tune_distributed_b200 import ops  # noqa: E402


def t_ms(fn, iters=5, warmup=2):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    dev = torch.device("cuda:0")
    shapes = [("dW-like", 64256, 2048, 8192), ("dX-like", 8192, 2048, 64256), ("logit-like", 8192, 64256, 2048)]
    for name, M, N, K in shapes:
        for a_mn in (False, True):
            for b_mn in (False, True):
                a = torch.randn((K, M) if a_mn else (M, K), device=dev, dtype=torch.bfloat16)
                b = torch.randn((K, N) if b_mn else (N, K), device=dev, dtype=torch.bfloat16)
                ms = t_ms(lambda: ops.gemm_bf16(a, b, M, N, K, a_mn, b_mn))
                A = a.t() if a_mn else a
                B = b if b_mn else b.t()
                ms_t = t_ms(lambda: torch.matmul(A, B))
                fl = 2.0 * M * N * K
                print(json.dumps({"shape": name, "M": M, "N": N, "K": K, "a_mn": a_mn, "b_mn": b_mn, "ms": round(ms, 3),
                                  "tflops": round(fl / ms / 1e9, 1), "cublas_ms": round(ms_t, 3),
                                  "cublas_tflops": round(fl / ms_t / 1e9, 1)}), flush=True)
                del a, b


if __name__ == "__main__":
    main()
