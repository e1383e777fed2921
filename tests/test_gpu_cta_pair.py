"""cta_group::2 engine variant (two-CTA clusters, 256x256 tiles) vs the single-CTA engine: the
contraction order per output element is the same, so results must be bit-identical -- for every
operand layout of the plain GEMM and for the fused forward / backward (dense and compacted)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture
def pair_mode():
    from llm_fine_tune_distributed_b200 import _cabi

    def use(n):
        _cabi.set_cta_group(n)

    yield use
    _cabi.set_cta_group(2)  # library default


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (200, 328, 136), (1024, 2048, 1024), (128 * 41, 256 * 9, 320)])
def test_pair_gemm_equals_single_cta(cuda_device, pair_mode, a_mn, b_mn, M, N, K):
    from llm_fine_tune_distributed_b200 import ops

    g = torch.Generator().manual_seed(M + N + K)
    a = torch.randn((K, M) if a_mn else (M, K), generator=g).to(torch.bfloat16).to(cuda_device)
    b = torch.randn((K, N) if b_mn else (N, K), generator=g).to(torch.bfloat16).to(cuda_device)
    pair_mode(1)
    ref = ops.gemm_bf16(a, b, M, N, K, a_mn, b_mn).clone()
    pair_mode(2)
    got = ops.gemm_bf16(a, b, M, N, K, a_mn, b_mn)
    torch.cuda.synchronize()
    assert torch.equal(got, ref)
    full = (a.float().t() if a_mn else a.float()) @ (b.float() if b_mn else b.float().t())
    assert torch.allclose(got.float(), full, rtol=2.0 ** -7, atol=1e-3 * K ** 0.5)


@pytest.mark.parametrize("N,H,V,masked,skip", [(384, 128, 1536, 0.3, False), (1000, 72, 1000, 0.6, True),
                                               (2048, 256, 8192, 0.0, False)])
def test_pair_flce_equals_single_cta(cuda_device, pair_mode, N, H, V, masked, skip):
    from llm_fine_tune_distributed_b200 import ops

    g = torch.Generator().manual_seed(N + V)
    x = torch.randn(N, H, generator=g).to(torch.bfloat16).to(cuda_device)
    w = (torch.randn(V, H, generator=g) * 0.05).to(torch.bfloat16).to(cuda_device)
    labels = torch.randint(0, V, (N,), generator=g, dtype=torch.int32)
    labels[torch.rand(N, generator=g) < masked] = -100
    labels = labels.to(cuda_device)
    denom = torch.tensor(max(int((labels != -100).sum()), 1), dtype=torch.int64, device=cuda_device)
    outs = []
    for grp in (1, 2):
        pair_mode(grp)
        xg, wg = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
        loss = ops.fused_linear_cross_entropy(xg, wg, labels, denom, skip_ignored_rows=skip, chunk_cols=512)
        loss.backward()
        torch.cuda.synchronize()
        outs.append((loss.detach().clone(), xg.grad.clone(), wg.grad.clone()))
    assert torch.equal(outs[0][0], outs[1][0])
    assert torch.equal(outs[0][1], outs[1][1])
    assert torch.equal(outs[0][2], outs[1][2])
