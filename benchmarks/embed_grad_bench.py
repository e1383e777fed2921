"""b200sft_embed_grad_rows on one GPU with `world` emulated ranks (raw pointers per rank): time per call and the bytes it
has to move (world * tokens dY rows read + touched grad rows read and written), against the dense alternative --
torch's embedding backward + a 525 MB add, which is what keeps the tied bucket until the end of backward."""
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_fine_tune_distributed_b200 import _cabi  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    lib = _cabi.load()
    V, H, cap = 128256, 2048, 8192
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)
    for world in (1, 2, 4, 8):
        for dist_name in ("uniform", "zipf"):
            g = torch.Generator().manual_seed(world)
            if dist_name == "uniform":
                ids = [torch.randint(0, V, (cap,), generator=g, dtype=torch.int32).to(dev) for _ in range(world)]
            else:  # text-like: a few very frequent tokens
                w = 1.0 / torch.arange(1, V + 1, dtype=torch.float64)
                ids = [torch.multinomial(w, cap, replacement=True, generator=g).to(torch.int32).to(dev) for _ in range(world)]
            dys = [(torch.randn(cap, H, generator=g) * 0.05).to(torch.bfloat16).to(dev) for _ in range(world)]
            ns = [torch.tensor(cap, dtype=torch.int32, device=dev) for _ in range(world)]
            grad = torch.zeros(V, H, dtype=torch.bfloat16, device=dev)
            ws_bytes = lib.b200sft_embed_grad_workspace_bytes(world, cap, V, H)
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
            tab = lambda ts: (ctypes.c_uint64 * world)(*[t.data_ptr() for t in ts])
            a, b, c = tab(ids), tab(dys), tab(ns)
            stream = torch.cuda.current_stream(dev).cuda_stream

            def call():
                _cabi.check(lib.b200sft_embed_grad_rows(a, b, c, world, cap, V, H, 1.0 / world, grad.data_ptr(), ws.data_ptr(),
                                                        ws_bytes, stream), "b200sft_embed_grad_rows")

            for _ in range(3):
                call()
            torch.cuda.synchronize()
            tot = 0.0
            for _ in range(10):
                flush.zero_()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                call()
                e.record()
                torch.cuda.synchronize()
                tot += s.elapsed_time(e)
            if "--kernel-times" in sys.argv:
                from torch.profiler import ProfilerActivity, profile

                with profile(activities=[ProfilerActivity.CUDA]) as prof:
                    for _ in range(5):
                        call()
                    torch.cuda.synchronize()
                for ev in sorted(prof.key_averages(), key=lambda e: -e.device_time_total):
                    if ev.device_time_total > 0:
                        print(f"   {ev.device_time_total / ev.count:8.1f} us x{ev.count:3d}  {ev.key[:90]}")
            touched = int(torch.unique(torch.cat(ids)).numel())
            moved = world * cap * H * 2 + 2 * touched * H * 2
            # the dense alternative on one rank: embedding backward ([V,H] zeros + index_add) and the add into the gradient
            wt = torch.zeros(V, H, dtype=torch.bfloat16, device=dev, requires_grad=True)
            t2 = 0.0
            for it in range(5):
                flush.zero_()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                wt.grad = None
                torch.nn.functional.embedding(ids[0].long(), wt).backward(dys[0])
                grad.add_(wt.grad)
                e.record()
                torch.cuda.synchronize()
                if it:
                    t2 += s.elapsed_time(e)
            print(json.dumps({"world": world, "tokens_per_rank": cap, "V": V, "H": H, "ids": dist_name, "touched_rows": touched,
                              "rows_kernel_ms": round(tot / 10, 4), "bytes_moved": moved, "GBps": round(moved / (tot / 10) / 1e6, 1),
                              "workspace_MiB": round(ws_bytes / 2**20, 1),
                              "torch_dense_embedding_bwd_plus_add_ms_one_rank": round(t2 / 4, 4)}), flush=True)


if __name__ == "__main__":
    main()
