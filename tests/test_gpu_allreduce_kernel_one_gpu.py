"""Seam #2 on ONE GPU: the peer-pointer reduce+scale kernel (b200sft_allreduce_scale_bf16_p2p) takes raw device
addresses, so `world` "ranks" can be emulated with `world` buffers on one device -- launch it once per emulated
rank, each of which reduces its 1/world slice of every buffer and writes the result back into all of them.

Reference arithmetic: default_hooks.allreduce_hook (SP/torch/distributed/algorithms/ddp_comm_hooks/default_hooks.py:18-32,
configured by training.py:249-256) averages a bucket as div_(world) + all_reduce(SUM).  This kernel keeps the
sum in fp32 and rounds once: result = bf16(fp32_sum(bufs) * fp32(1/world)), checked bit for bit here, and
against the reference's own order (round after the divide, bf16 sum) to bf16 round-off.
"""
import ctypes

import pytest
import torch

from oracle import sft_oracle as oracle

pytestmark = pytest.mark.gpu


def _f32_recip(world: int) -> float:
    return float(torch.tensor(1.0, dtype=torch.float32) / torch.tensor(float(world), dtype=torch.float32))


def _run_p2p(lib, bufs, n, world, scale, dev):
    ptrs = (ctypes.c_uint64 * world)(*[b.data_ptr() for b in bufs])
    stream = torch.cuda.current_stream(dev).cuda_stream
    for rank in range(world):  # each emulated rank owns slice `rank`; slices are disjoint, so order does not matter
        rc = lib.b200sft_allreduce_scale_bf16_p2p(ptrs, n, rank, world, scale, stream)
        assert rc == 0, lib.b200sft_last_error().decode()
    torch.cuda.synchronize(dev)


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("vecs_per_rank", [1, 37, 4099, 1 << 16])
def test_p2p_kernel_every_rank_gets_the_rounded_fp32_average(cuda_device, world, vecs_per_rank):
    from llm_fine_tune_distributed_b200 import _cabi

    lib = _cabi.load()
    n = 8 * world * vecs_per_rank                      # the kernel's granularity: 16-byte vectors, one slice per rank
    g = torch.Generator().manual_seed(world * 1000 + vecs_per_rank)
    with torch.cuda.device(cuda_device):
        # one allocation carved into `world` buffers, as a symmetric-memory arena would be
        arena = (torch.randn(world, n, generator=g) * 3).to(torch.bfloat16).to(cuda_device)
        inputs = arena.clone()
        scale = _f32_recip(world)
        _run_p2p(lib, [arena[r] for r in range(world)], n, world, scale, cuda_device)
    acc = torch.zeros(n, dtype=torch.float32, device=cuda_device)
    for r in range(world):                             # the kernel adds the peers in rank order
        acc += inputs[r].float()
    expect = (acc * scale).to(torch.bfloat16)
    for r in range(world):
        assert torch.equal(arena[r], expect), f"rank {r}'s copy differs from bf16(fp32 sum * 1/world)"
    # the reference's order of operations (oracle.allreduce_hook_bucket: divide and round per rank, sum in bf16)
    # differs from the single fp32 rounding by bf16 round-off only
    total = torch.zeros(n, dtype=torch.bfloat16, device=cuda_device)
    divided = [inputs[r].clone().div_(world) for r in range(world)]
    for d in divided:
        total += d
    ref = oracle.allreduce_hook_bucket(inputs[0].clone(), world, lambda t: t.copy_(total))
    # (world divides + world - 1 bf16 additions + this kernel's one rounding, each within 2^-9 of a term <= max |x_r|)
    bound = inputs.float().abs().amax(dim=0) * ((world + 1) * 2.0 ** -9)
    assert bool(((ref.float() - expect.float()).abs() <= bound).all())


def test_p2p_kernel_power_of_two_world_matches_reference_bits_when_exact(cuda_device):
    """With inputs that are exactly representable after the divide and whose sums need no rounding (small integers),
    every order of operations agrees: the kernel must reproduce the oracle's allreduce_hook bit for bit."""
    from llm_fine_tune_distributed_b200 import _cabi

    lib = _cabi.load()
    world, n = 4, 8 * 4 * 513
    g = torch.Generator().manual_seed(3)
    with torch.cuda.device(cuda_device):
        arena = torch.randint(-32, 33, (world, n), generator=g).to(torch.bfloat16).to(cuda_device)
        inputs = arena.clone()
        _run_p2p(lib, [arena[r] for r in range(world)], n, world, 0.25, cuda_device)
    total = torch.zeros(n, dtype=torch.bfloat16, device=cuda_device)
    for r in range(world):
        total += inputs[r].clone().div_(world)
    ref = oracle.allreduce_hook_bucket(inputs[0].clone(), world, lambda t: t.copy_(total))
    for r in range(world):
        assert torch.equal(arena[r], ref)


def test_p2p_kernel_error_paths(cuda_device):
    from llm_fine_tune_distributed_b200 import _cabi

    lib = _cabi.load()
    buf = torch.zeros(64, dtype=torch.bfloat16, device=cuda_device)
    stream = torch.cuda.current_stream(cuda_device).cuda_stream
    two = (ctypes.c_uint64 * 2)(buf.data_ptr(), buf.data_ptr())
    assert lib.b200sft_allreduce_scale_bf16_p2p(two, 24, 0, 2, 0.5, stream) != 0        # not a multiple of 8 * world
    assert "multiple of 8*world" in lib.b200sft_last_error().decode()
    assert lib.b200sft_allreduce_scale_bf16_p2p(two, 32, 2, 2, 0.5, stream) != 0        # rank out of range
    assert lib.b200sft_allreduce_scale_bf16_p2p(None, 32, 0, 2, 0.5, stream) != 0       # no pointer table
    bad = (ctypes.c_uint64 * 2)(buf.data_ptr(), buf.data_ptr() + 2)                     # misaligned peer
    assert lib.b200sft_allreduce_scale_bf16_p2p(bad, 32, 0, 2, 0.5, stream) != 0
    many = (ctypes.c_uint64 * 17)(*([buf.data_ptr()] * 17))
    assert lib.b200sft_allreduce_scale_bf16_p2p(many, 8 * 17, 0, 17, 1.0, stream) != 0  # more peers than the kernel holds
    assert lib.b200sft_allreduce_scale_bf16_p2p(two, 0, 0, 2, 0.5, stream) == 0         # empty bucket: nothing to do
    assert lib.b200sft_allreduce_scale_bf16_multimem(None, 32, 0, 2, 0.5, stream) != 0  # no multicast address
    torch.cuda.synchronize(cuda_device)


@pytest.mark.parametrize("world", [2, 3, 5, 6, 7, 8])
def test_scale_kernel_has_the_bits_of_div_for_every_world(cuda_device, world):
    """default_hooks.py:25 `tensor.div_(world_size)` on a CUDA bf16 bucket vs ops.scale_(buf, 1/world): bit-equal for
    every world size of one box, not only powers of two (ATen's CUDA div by a host scalar multiplies by the fp32
    reciprocal).  Same for the compress hook's `buffer.to(bf16).div_(world)` (default_hooks.py:74)."""
    from llm_fine_tune_distributed_b200 import ops

    g = torch.Generator().manual_seed(world)
    buf = (torch.randn((1 << 20) + 3, generator=g) * 5).to(torch.bfloat16).to(cuda_device)
    ref = buf.clone().div_(world)
    ops.scale_(buf, 1.0 / world)
    assert torch.equal(buf, ref)
    src = (torch.randn((1 << 20) + 3, generator=g) * 5).to(cuda_device)
    ref_c = src.to(torch.bfloat16).div_(world)
    assert torch.equal(ops.scale_cast_f32_to_bf16(src, 1.0 / world), ref_c)
