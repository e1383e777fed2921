"""Step tail (global grad-norm clip + AdamW) on the reference freeze policy's trainable tensors
(418.9 M bf16 parameters: last two SmolLM3-3B decoder layers + the tied 128256 x 2048 head):
this library's two multi-tensor kernels vs torch.nn.utils.clip_grad_norm_ + torch.optim.AdamW(fused).
Algorithmic bytes: 2 B (norm pass) + 8 B read + 6 B written per parameter = 16 B."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_fine_tune_distributed_b200.optim import B200AdamW  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    peak = 6650.0
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p))["hbm_gbs"]
    Hd, I, V = 2048, 11008, 128256
    layer = [(2048, Hd), (512, Hd), (512, Hd), (Hd, 2048), (I, Hd), (I, Hd), (Hd, I), (Hd,), (Hd,)]
    shapes = [(V, Hd)] + layer + layer
    mk = lambda: [torch.nn.Parameter(torch.randn(*s, device=dev, dtype=torch.bfloat16) * 0.02) for s in shapes]
    ours, ref = mk(), mk()
    n = sum(p.numel() for p in ours)
    for ps in (ours, ref):
        for q in ps:
            q.grad = torch.randn_like(q)
    opt = B200AdamW(ours, lr=5e-5, weight_decay=0.0)
    opt_ref = torch.optim.AdamW(ref, lr=5e-5, weight_decay=0.0, fused=True)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)

    def t_ms(fn, iters=8, warmup=3):
        for _ in range(warmup):
            fn()
        tot = 0.0
        for _ in range(iters):
            flush.zero_()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            tot += a.elapsed_time(b)
        return tot / iters

    def torch_tail():
        torch.nn.utils.clip_grad_norm_(ref, 1.0)
        opt_ref.step()

    ms = t_ms(lambda: opt.clip_and_step(1.0))
    ms_t = t_ms(torch_tail)
    print(json.dumps({"params": n, "tensors": len(shapes), "b200_ms": round(ms, 4), "torch_ms": round(ms_t, 4),
                      "b200_GBps": round(16 * n / ms / 1e6, 1), "frac_of_hbm_peak": round(16 * n / ms / 1e6 / peak, 4),
                      "speedup": round(ms_t / ms, 3), "hbm_peak_GBps": peak}))


if __name__ == "__main__":
    main()
