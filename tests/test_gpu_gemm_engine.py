"""tcgen05 GEMM engine vs torch.matmul (fp32 reference), one test per operand-layout path.

C = A . B^T with fp32 accumulation and a single bf16 rounding at the end, so the result must agree
with an fp32 matmul of the same bf16 inputs to bf16 resolution (rel 2^-8) -- written below.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a, b, a_mn, b_mn):
    A = a.float().t() if a_mn else a.float()
    B = b.float().t() if b_mn else b.float()
    return A @ B.t()


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 512, 256), (384, 768, 1024), (200, 328, 136),
                                   (1024, 2048, 4096)])
def test_gemm_layouts(cuda_device, a_mn, b_mn, M, N, K):
    from llm_fine_tune_distributed_b200 import ops

    g = torch.Generator(device="cpu").manual_seed(M * 7 + N * 3 + K)
    a = torch.randn((K, M) if a_mn else (M, K), generator=g).to(torch.bfloat16).to(cuda_device)
    b = torch.randn((K, N) if b_mn else (N, K), generator=g).to(torch.bfloat16).to(cuda_device)
    c = ops.gemm_bf16(a, b, M, N, K, a_mn, b_mn).float()
    ref = _ref(a, b, a_mn, b_mn)
    torch.cuda.synchronize()
    # bf16 output rounding: |err| <= 2^-8 |ref| + small absolute term from fp32 accumulation order
    tol = 2.0 ** -8 * ref.abs() + 1e-3 * (K ** 0.5)
    bad = (c - ref).abs() > tol
    assert not bad.any(), f"{int(bad.sum())} / {bad.numel()} elements off; max abs err {(c - ref).abs().max().item()}"


def test_gemm_many_tiles_persistent(cuda_device):
    """More tiles than SMs, so every CTA loops and both TMEM accumulator stages recycle."""
    from llm_fine_tune_distributed_b200 import ops

    M, N, K = 128 * 20, 256 * 20, 192
    g = torch.Generator(device="cpu").manual_seed(5)
    a = torch.randn(M, K, generator=g).to(torch.bfloat16).to(cuda_device)
    b = torch.randn(N, K, generator=g).to(torch.bfloat16).to(cuda_device)
    c = ops.gemm_bf16(a, b, M, N, K).float()
    ref = a.float() @ b.float().t()
    assert torch.allclose(c, ref, rtol=2.0 ** -7, atol=0.05)
