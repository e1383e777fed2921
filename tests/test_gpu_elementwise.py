"""Integer / byte kernels vs the oracle: bit-exact for labels and counts; the scale / cast kernels
must equal torch's own bf16 arithmetic bit for bit (same single rounding)."""
import pytest
import torch

from oracle import sft_oracle as oracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,T,masked", [(1, 1, 0.0), (3, 17, 0.5), (8, 1024, 0.6), (2, 4096, 0.95), (5, 33, 1.0)])
def test_label_prep_bit_exact(cuda_device, B, T, masked):
    from llm_fine_tune_distributed_b200 import ops

    g = torch.Generator().manual_seed(B * 1000 + T)
    labels = torch.randint(0, 128256, (B, T), generator=g)
    labels[torch.rand(B, T, generator=g) < masked] = -100
    shifted, counts = ops.label_prep(labels.to(cuda_device))
    ref = oracle.shift_labels(labels)
    assert torch.equal(shifted.cpu().long().view(B, T), ref)
    assert int(counts[0]) == oracle.num_items_in_batch([labels])
    assert int(counts[1]) == int((ref != -100).sum())


def test_label_prep_empty(cuda_device):
    from llm_fine_tune_distributed_b200 import ops

    shifted, counts = ops.label_prep(torch.empty(0, 16, dtype=torch.int64, device=cuda_device))
    assert shifted.numel() == 0 and counts.tolist() == [0, 0]


@pytest.mark.parametrize("n,offset", [(1, 0), (7, 0), (8, 0), (1000, 1), (12345, 3), (1 << 20, 0), ((1 << 22) + 5, 5)])
def test_scale_bf16_inplace(cuda_device, n, offset):
    from llm_fine_tune_distributed_b200 import ops

    g = torch.Generator().manual_seed(n)
    base = torch.randn(n + offset, generator=g).to(torch.bfloat16).to(cuda_device)
    view = base[offset:]                      # unaligned bucket views must work too
    ref = (view.clone()).div_(8)              # default_hooks.py:25 tensor.div_(world_size), world = 8
    ops.scale_(view, 1.0 / 8)
    assert torch.equal(view, ref)
    ref3 = view.clone().float().mul(1.0 / 3).to(torch.bfloat16)
    ops.scale_(view, 1.0 / 3)
    assert torch.equal(view, ref3)


@pytest.mark.parametrize("n", [1, 9, 4096, (1 << 20) + 3])
def test_compress_decompress(cuda_device, n):
    from llm_fine_tune_distributed_b200 import ops

    g = torch.Generator().manual_seed(n)
    src = torch.randn(n, generator=g).to(cuda_device)
    ref = src.to(torch.bfloat16).div_(4)      # default_hooks.py:74 buffer.to(bf16).div_(world)
    got = ops.scale_cast_f32_to_bf16(src, 0.25)
    # torch rounds to bf16 THEN divides; a power-of-two scale commutes with rounding, so bit-exact
    assert torch.equal(got, ref)
    back = torch.empty_like(src)
    ops.cast_bf16_to_f32(got, back)
    assert torch.equal(back, ref.float())
    f = src.clone()
    ops.scale_(f, 0.5)
    assert torch.equal(f, src * 0.5)
