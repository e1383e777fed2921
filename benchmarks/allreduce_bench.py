"""Gradient-bucket averaging over NVLink: this library's fused reduce+scale kernel (multimem / p2p)
vs scale kernel + NCCL allreduce, per DDP bucket size of the reference freeze policy.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 benchmarks/allreduce_bench.py

Times are CUDA events on the launch stream, max over ranks, L2 flushed between iterations; "bus
GB/s" uses the allreduce convention 2(n-1)/n * bytes / time.  The nvlink rows include the copy into
symmetric memory and both barriers (everything the hook does)."""
import ctypes
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from llm_fine_tune_distributed_b200 import _cabi, ops  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    lib = _cabi.load()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()

    def timed(fn, iters=8, warmup=3):
        for _ in range(warmup):
            fn()
        tot = 0.0
        for _ in range(iters):
            flush.zero_()
            dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            t = torch.tensor([a.elapsed_time(b)], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            tot += float(t)
        return tot / iters

    for mib in (43, 86, 521):
        n = (mib << 20) // 2
        n = n // (8 * world) * (8 * world)
        bucket = torch.randn(n, device=dev).to(torch.bfloat16)
        sbuf = symm_mem.empty(n, dtype=torch.bfloat16, device=dev)
        hdl = symm_mem.rendezvous(sbuf, dist.group.WORLD)
        ptrs = (ctypes.c_uint64 * world)(*[int(p) for p in hdl.buffer_ptrs])

        def nccl():
            ops.scale_(bucket, 1.0 / world)
            dist.all_reduce(bucket)

        def nvl(multimem):
            sbuf.copy_(bucket)
            hdl.barrier(channel=0)
            if multimem:
                _cabi.check(lib.b200sft_allreduce_scale_bf16_multimem(hdl.multicast_ptr, n, rank, world, 1.0 / world,
                                                                      stream.cuda_stream), "multimem")
            else:
                _cabi.check(lib.b200sft_allreduce_scale_bf16_p2p(ptrs, n, rank, world, 1.0 / world, stream.cuda_stream), "p2p")
            hdl.barrier(channel=0)

        res = {"MiB": mib, "world": world, "nccl_ms": round(timed(nccl), 4)}
        if hdl.multicast_ptr:
            res["multimem_ms"] = round(timed(lambda: nvl(True)), 4)
        res["p2p_ms"] = round(timed(lambda: nvl(False)), 4)
        bus = lambda ms: round(2 * (world - 1) / world * n * 2 / ms / 1e6, 1)
        for k in ("nccl", "multimem", "p2p"):
            if k + "_ms" in res:
                res[k + "_bus_GBps"] = bus(res[k + "_ms"])
        if rank == 0:
            print(json.dumps(res), flush=True)
        del sbuf, hdl, bucket
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
