"""Final RMSNorm fused into the head's row gather (b200sft_rmsnorm_gather / _scatter_bwd) against the
oracle's restatement of SmolLM3RMSNorm (modeling_smollm3.py:271-276) and its autograd backward."""
import numpy as np
import pytest
import torch

from oracle import sft_oracle as oracle

pytestmark = pytest.mark.gpu
EPS = 1e-6
cos = lambda a, b: torch.nn.functional.cosine_similarity(a.reshape(-1).float(), b.reshape(-1).float(), dim=0).item()


def _inputs(dev, N, H, p_ignore, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    h = (torch.randn(N, H, generator=g) * 3.0 + 0.3).to(torch.bfloat16).to(dev)
    gamma = (1.0 + 0.2 * torch.randn(H, generator=g)).to(torch.bfloat16).to(dev)
    labels = torch.randint(0, 1000, (N,), generator=g, dtype=torch.int32)
    labels[torch.rand(N, generator=g) < p_ignore] = -100
    return h, gamma, labels.to(dev)


@pytest.mark.parametrize("N,H", [(1, 8), (37, 64), (300, 2048), (1000, 264), (8192, 2048)])
@pytest.mark.parametrize("p_ignore", [None, 0.0, 0.6, 1.0])
def test_rmsnorm_gather_matches_oracle(cuda_device, N, H, p_ignore):
    from llm_fine_tune_distributed_b200 import ops

    h, gamma, labels = _inputs(cuda_device, N, H, p_ignore or 0.0, seed=N + H)
    plan = ops.compaction_list(labels) if p_ignore is not None else None
    xc, rstd = ops.rmsnorm_gather(h, gamma, EPS, plan)
    rows = plan.row_index[: int(plan.n_live)].long() if plan is not None else torch.arange(N, device=cuda_device)
    n = rows.numel()
    want = oracle.rms_norm(h[rows], gamma, EPS)
    want_rstd = torch.rsqrt(h[rows].float().pow(2).mean(-1) + EPS)
    torch.testing.assert_close(rstd[:n], want_rstd, rtol=2e-6, atol=0)
    got = xc[:n]
    # fp32 summation order can move 1/rms by an ulp, which (rarely) flips the bf16 rounding of the normalised
    # value and then of its product with the weight: two roundings, one bf16 ulp (<= 2^-7 relative) each
    torch.testing.assert_close(got.float(), want.float(), rtol=2 ** -6, atol=1e-30)
    assert n == 0 or (got == want).float().mean().item() > 0.995
    if plan is not None:  # zero rows up to the next 128-row boundary
        pad = min((n + 127) // 128 * 128, N)
        assert torch.count_nonzero(xc[n:pad]) == 0


@pytest.mark.parametrize("N,H", [(37, 64), (300, 2048), (1000, 264), (4096, 2048)])
@pytest.mark.parametrize("p_ignore", [None, 0.5, 1.0])
def test_rmsnorm_scatter_bwd_matches_autograd(cuda_device, N, H, p_ignore):
    from llm_fine_tune_distributed_b200 import ops

    h, gamma, labels = _inputs(cuda_device, N, H, p_ignore or 0.0, seed=7 * N + H)
    plan = ops.compaction_list(labels) if p_ignore is not None else None
    n = int(plan.n_live) if plan is not None else N
    rows = plan.row_index[:n].long() if plan is not None else torch.arange(N, device=cuda_device)
    dyc = torch.zeros(N, H, dtype=torch.bfloat16, device=cuda_device)
    dyc[:n] = (torch.randn(n, H, device=cuda_device) * 0.01).to(torch.bfloat16)
    _, rstd = ops.rmsnorm_gather(h, gamma, EPS, plan)
    dh, dgamma = ops.rmsnorm_scatter_bwd(dyc, h, gamma, rstd, plan, need_dh=True, need_dgamma=True)
    hh = h.clone().requires_grad_(True)
    gg = gamma.clone().requires_grad_(True)
    if n:
        oracle.rms_norm(hh[rows], gg, EPS).backward(dyc[:n])
    want_dh = hh.grad if hh.grad is not None else torch.zeros_like(h)
    want_dg = gg.grad if gg.grad is not None else torch.zeros_like(gamma)
    dead = torch.ones(N, dtype=torch.bool, device=cuda_device)
    dead[rows] = False
    assert torch.count_nonzero(dh[dead]) == 0          # ignored rows: exactly zero
    if n:
        assert cos(dh, want_dh) > 0.99999 and cos(dgamma, want_dg) > 0.9999
        torch.testing.assert_close(dh.float(), want_dh.float(), rtol=2 ** -6, atol=float(want_dh.abs().max()) * 2 ** -8)
        torch.testing.assert_close(dgamma.float(), want_dg.float(), rtol=2 ** -5, atol=float(want_dg.abs().max()) * 2 ** -6)
    else:
        assert torch.count_nonzero(dh) == 0 and torch.count_nonzero(dgamma) == 0
    # twice in a row: same bits (fixed summation order)
    dh2, dg2 = ops.rmsnorm_scatter_bwd(dyc, h, gamma, rstd, plan, need_dh=True, need_dgamma=True)
    assert torch.equal(dh, dh2) and torch.equal(dgamma, dg2)


@pytest.mark.parametrize("skip", [False, True])
def test_fused_norm_head_loss_and_grads(cuda_device, skip):
    """norm -> head -> CE as ONE differentiable op against the oracle chain rms_norm -> lm_head_ce."""
    from llm_fine_tune_distributed_b200 import ops

    N, H, V = 512, 256, 3000
    h, gamma, labels = _inputs(cuda_device, N, H, 0.4, seed=3)
    labels = torch.where(labels >= 0, labels % V, labels)
    w = (torch.randn(V, H, device=cuda_device) * 0.05).to(torch.bfloat16)
    n_items = int((labels != -100).sum())
    hp, wp, gp = (t.clone().requires_grad_(True) for t in (h, w, gamma))
    loss = ops.fused_linear_cross_entropy(hp, wp, labels, n_items, skip_ignored_rows=skip, norm_weight=gp, norm_eps=EPS)
    loss.backward()
    ho, wo, go = (t.clone().requires_grad_(True) for t in (h, w, gamma))
    ref = oracle.lm_head_ce(oracle.rms_norm(ho, go, EPS), wo, labels.long(), float(n_items))
    ref.backward()
    assert abs(loss.item() - ref.item()) <= 1e-3 * abs(ref.item())  # north_star tolerance
    assert cos(hp.grad, ho.grad) >= 0.999 and cos(wp.grad, wo.grad) >= 0.999 and cos(gp.grad, go.grad) >= 0.999
    assert torch.count_nonzero(hp.grad[labels == -100]) == 0
    # frozen norm weight (the reference's freeze policy): no dgamma work, same dX
    hp2 = h.clone().requires_grad_(True)
    ops.fused_linear_cross_entropy(hp2, w, labels, n_items, skip_ignored_rows=skip, norm_weight=gamma, norm_eps=EPS).backward()
    assert torch.equal(hp2.grad, hp.grad)


def test_model_with_fused_final_norm_matches_stock_path(cuda_device):
    """Tiny SmolLM3 through the compute_loss seam: final norm inside the head's gather vs the model's own."""
    import copy

    from transformers import SmolLM3Config, SmolLM3ForCausalLM

    from llm_fine_tune_distributed_b200 import FusedCELossMixin, install_fused_head
    from llm_fine_tune_distributed_b200.loss import FusedFinalNorm

    cfg = SmolLM3Config(num_hidden_layers=2, hidden_size=128, intermediate_size=256, num_attention_heads=4,
                        num_key_value_heads=2, vocab_size=1024, max_position_embeddings=256, tie_word_embeddings=True,
                        pad_token_id=0, bos_token_id=1, eos_token_id=2)
    cfg._attn_implementation = "sdpa"
    torch.manual_seed(0)
    base = SmolLM3ForCausalLM(cfg).to(torch.bfloat16).to(cuda_device)
    with torch.no_grad():
        base.model.norm.weight.add_(0.1 * torch.randn_like(base.model.norm.weight))
    keys = set(base.state_dict().keys())
    ids = torch.randint(3, 1024, (3, 64), device=cuda_device)
    labels = ids.clone()
    labels[:, :20] = -100
    n = torch.tensor(int((labels != -100).sum()), device=cuda_device)

    class Seam(FusedCELossMixin):
        num_processes = 1
        model_accepts_loss_kwargs = True
        args = None

    out = {}
    for fused in (False, True):
        m = copy.deepcopy(base)
        install_fused_head(m, fuse_final_norm=fused)
        assert isinstance(m.model.norm, FusedFinalNorm) == fused and set(m.state_dict().keys()) == keys
        seam = Seam()
        seam.b200sft_fuse_final_norm = fused
        loss = seam.compute_loss(m, {"input_ids": ids, "labels": labels}, num_items_in_batch=n)
        loss.backward()
        assert not m.model.norm.armed if fused else True
        out[fused] = (loss.item(), m.model.norm.weight.grad.clone(), m.model.layers[-1].mlp.down_proj.weight.grad.clone(),
                      m.lm_head.weight.grad.clone())
        if fused:  # unarmed, the stand-in computes the stock norm: plain forward gives the stock logits
            with torch.no_grad():
                torch.testing.assert_close(m(ids).logits, base(ids).logits, rtol=0, atol=0)
    assert abs(out[True][0] - out[False][0]) <= 1e-3 * abs(out[False][0])
    for a, b in zip(out[True][1:], out[False][1:]):
        assert cos(a, b) >= 0.999
