"""b200sft_label_frontend (one launch: collator masks + shift + counts + compaction list) against
the numpy oracle -- bit-exact integer work -- and its use by the fused loss."""
import numpy as np
import pytest
import torch

from oracle import sft_oracle as oracle

pytestmark = pytest.mark.gpu


def _case(rng, B, T, masks, p_ignore=0.3):
    ids = rng.integers(0, 50000, size=(B, T)).astype(np.int64)
    labels = att = comp = posid = None
    if "labels" in masks:
        labels = ids.copy()
        labels[rng.random((B, T)) < p_ignore] = -100
    if "att" in masks:
        lens = rng.integers(0, T + 1, size=B)
        att = (np.arange(T)[None, :] < lens[:, None]).astype(np.int64)
    if "comp" in masks:
        comp = (rng.random((B, T)) > 0.5).astype(np.int64)
    if "pos" in masks:  # packed rows: position ids restart at every sequence start
        posid = np.zeros((B, T), dtype=np.int64)
        for b in range(B):
            t = 0
            while t < T:
                n = int(rng.integers(1, max(2, T // 3)))
                posid[b, t:t + n] = np.arange(min(n, T - t))
                t += n
    return ids, labels, att, comp, posid


def _run(ops, dev, ids, labels, att, comp, posid, **kw):
    t = lambda a: None if a is None else torch.from_numpy(a).to(dev)
    return ops.label_frontend(t(ids), t(labels), t(att), t(comp), t(posid), return_labels=True, **kw)


def _check(plan, want, compact=True):
    np.testing.assert_array_equal(plan.labels.cpu().numpy(), want["labels"])
    np.testing.assert_array_equal(plan.shifted.cpu().numpy(), want["shifted"])
    assert tuple(plan.counts.tolist()) == want["counts"]
    if not compact:
        assert plan.row_index is None and plan.n_live is None
        return
    n = int(plan.n_live)
    assert n == want["n_live"]
    np.testing.assert_array_equal(plan.row_index.cpu().numpy()[:n], want["row_index"])
    np.testing.assert_array_equal(plan.labels_c.cpu().numpy(), want["labels_c"])
    np.testing.assert_array_equal(plan.pos.cpu().numpy(), want["pos"])


# up to 16383 tokens: one block; from 16384: one 8-CTA cluster exchanging totals through distributed shared memory
@pytest.mark.parametrize("B,T", [(1, 1), (1, 31), (3, 32), (2, 1023), (8, 1024), (5, 777), (1, 65536), (64, 1),
                                 (1, 16383), (4, 4096), (3, 5462), (5, 7001), (8, 8192), (16385, 1)])
@pytest.mark.parametrize("masks", [(), ("labels",), ("att",), ("labels", "comp"), ("pos",), ("labels", "att", "comp", "pos")])
def test_label_frontend_matches_oracle(cuda_device, B, T, masks):
    from llm_fine_tune_distributed_b200 import ops

    rng = np.random.default_rng(B * 7919 + T + len(masks))
    case = _case(rng, B, T, masks)
    _check(_run(ops, cuda_device, *case), oracle.label_frontend_numpy(*case))


@pytest.mark.parametrize("p_ignore", [0.0, 1.0])
def test_label_frontend_all_or_nothing(cuda_device, p_ignore):
    from llm_fine_tune_distributed_b200 import ops

    case = _case(np.random.default_rng(5), 4, 257, ("labels",), p_ignore)
    plan = _run(ops, cuda_device, *case)
    _check(plan, oracle.label_frontend_numpy(*case))
    assert int(plan.n_live) == (0 if p_ignore == 1.0 else 4 * 256)


def test_label_frontend_no_compaction_empty_and_errors(cuda_device):
    from llm_fine_tune_distributed_b200 import ops

    case = _case(np.random.default_rng(6), 3, 100, ("labels", "att"))
    _check(_run(ops, cuda_device, *case, compact=False), oracle.label_frontend_numpy(*case), compact=False)
    empty = ops.label_frontend(labels=torch.empty(0, 16, dtype=torch.int64, device=cuda_device))
    assert empty.counts.tolist() == [0, 0] and int(empty.n_live) == 0 and empty.shifted.numel() == 0
    with pytest.raises(ValueError):
        ops.label_frontend(labels=torch.zeros(2, 4, dtype=torch.int32, device=cuda_device))
    with pytest.raises(ValueError):
        ops.label_frontend(labels=torch.zeros(2, 4, dtype=torch.int64, device=cuda_device),
                           attention_mask=torch.ones(2, 5, dtype=torch.int64, device=cuda_device))
    with pytest.raises(ValueError):
        ops.label_frontend()


def test_label_frontend_agrees_with_label_prep_and_compact_rows(cuda_device):
    """Same answers as the two-step path it replaces inside the fused loss, twice in a row."""
    from llm_fine_tune_distributed_b200 import ops

    rng = np.random.default_rng(8)
    ids, labels, *_ = _case(rng, 8, 1024, ("labels",), 0.6)
    lab = torch.from_numpy(labels).to(cuda_device)
    x = torch.randn(8 * 1024, 64, device=cuda_device).to(torch.bfloat16)
    for _ in range(2):
        plan = ops.label_frontend(labels=lab)
        shifted, counts = ops.label_prep(lab)
        xc, lc, ri, pos, n_live = ops.compact_rows(x, shifted)
        n = int(n_live)
        assert torch.equal(plan.shifted, shifted) and torch.equal(plan.counts, counts) and int(plan.n_live) == n
        assert torch.equal(plan.row_index[:n], ri[:n]) and torch.equal(plan.labels_c, lc) and torch.equal(plan.pos, pos)
        pad = min((n + 127) // 128 * 128, x.shape[0])  # rows past the zero-filled boundary are unspecified
        assert torch.equal(ops.gather_rows(x, plan)[:pad], xc[:pad])


def test_fused_loss_through_frontend_matches_oracle(cuda_device):
    """fused_causal_lm_loss (labels -> front-end -> gather -> head) against the reference math."""
    from llm_fine_tune_distributed_b200.loss import fused_causal_lm_loss

    torch.manual_seed(0)
    B, T, H, V = 2, 96, 128, 1000
    hs = (torch.randn(B, T, H, device=cuda_device) * 0.5).to(torch.bfloat16).requires_grad_(True)
    w = (torch.randn(V, H, device=cuda_device) * 0.05).to(torch.bfloat16).requires_grad_(True)
    labels = torch.randint(0, V, (B, T), device=cuda_device)
    labels[torch.rand(B, T, device=cuda_device) < 0.4] = -100
    n_items = int((labels != -100).sum())
    loss = fused_causal_lm_loss(hs, w, labels, num_items_in_batch=n_items)
    loss.backward()
    ref, dh, dw = oracle.lm_head_ce_with_grads(hs.detach().reshape(B * T, H), w.detach(),
                                               oracle.shift_labels(labels).reshape(-1), n_items)
    assert abs(loss.item() - ref.item()) <= 1e-3 * abs(ref.item())  # north_star: loss within 1e-3 relative
    cos = torch.nn.functional.cosine_similarity
    assert cos(hs.grad.reshape(-1).float(), dh.reshape(-1).float(), dim=0) >= 0.999
    assert cos(w.grad.reshape(-1).float(), dw.reshape(-1).float(), dim=0) >= 0.999
