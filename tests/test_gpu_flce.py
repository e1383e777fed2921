"""Fused LM-head + cross-entropy (CUDA, through the C ABI) vs the oracle.

Tolerances are BASELINE.json's: loss within 1e-3 relative; gradients cosine >= 0.999 with the
max-relative error (against the float64 oracle, normalised by the largest gradient magnitude)
printed and bounded; lse within 1e-3 absolute of the fp32 oracle.  Label handling is bit-exact.
"""
import numpy as np
import pytest
import torch

from oracle import sft_oracle as oracle

pytestmark = pytest.mark.gpu


def _case(N, H, V, masked, seed, device):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(N, H, generator=g).to(torch.bfloat16)
    w = (torch.randn(V, H, generator=g) * 0.02 * (2048 / H) ** 0.5).to(torch.bfloat16)
    labels = torch.randint(0, V, (N,), generator=g, dtype=torch.int64)
    drop = torch.rand(N, generator=g) < masked
    labels[drop] = -100
    return x.to(device), w.to(device), labels.to(device)


def _cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm() + 1e-300))


CASES = [
    # N, H, V, masked fraction
    (128, 64, 256, 0.0),
    (256, 128, 1024, 0.25),
    (200, 72, 1000, 0.5),      # ragged rows, ragged vocabulary tile, K tail
    (1024, 256, 4096, 0.9),
    (512, 2048, 8192, 0.3),    # SmolLM3 hidden size
]


@pytest.mark.parametrize("N,H,V,masked", CASES)
@pytest.mark.parametrize("mode", ["num_items", "mean"])
def test_flce_forward_backward_vs_oracle(cuda_device, N, H, V, masked, mode):
    from llm_fine_tune_distributed_b200 import ops

    x, w, labels = _case(N, H, V, masked, seed=N + V, device=cuda_device)
    n_valid = int((labels != -100).sum())
    num_items = None if mode == "mean" else max(n_valid, 1) + 3  # denominator is an input, not the local count
    if mode == "mean" and n_valid == 0:
        pytest.skip("mean over zero rows is NaN by definition; covered separately")

    # oracle: reference arithmetic (bf16 linear -> fp32 CE) on the same device, and float64 on the host
    ref_loss, ref_dx, ref_dw = oracle.lm_head_ce_with_grads(x, w, labels, num_items, grad_out=1.7)
    loss64, lse64, dx64, dw64 = oracle.lm_head_ce_float64(x.float().cpu().numpy(), w.float().cpu().numpy(),
                                                          labels.cpu().numpy(), num_items)

    xg = x.clone().requires_grad_(True)
    wg = w.clone().requires_grad_(True)
    denom = None if num_items is None else torch.tensor(num_items, dtype=torch.int64, device=cuda_device)
    loss = ops.fused_linear_cross_entropy(xg, wg, labels.to(torch.int32), denom)
    (loss * 1.7).backward()
    torch.cuda.synchronize()

    assert abs(loss.item() - ref_loss.item()) <= 1e-3 * abs(ref_loss.item()), (loss.item(), ref_loss.item())
    assert abs(loss.item() - loss64) <= 1e-3 * abs(loss64)

    for name, got, ref_bf16, ref64 in (("dx", xg.grad, ref_dx, dx64 * 1.7), ("dw", wg.grad, ref_dw, dw64 * 1.7)):
        ref64_t = torch.from_numpy(ref64)
        cos_ref = _cos(got.cpu(), ref_bf16.cpu())
        cos64 = _cos(got.cpu(), ref64_t)
        scale = float(ref64_t.abs().max()) + 1e-30
        max_rel = float((got.cpu().double() - ref64_t).abs().max()) / scale
        max_rel_ref = float((ref_bf16.cpu().double() - ref64_t).abs().max()) / scale
        print(f"{name}: cos(vs ref)={cos_ref:.6f} cos(vs f64)={cos64:.6f} max_rel_err={max_rel:.3e} "
              f"(reference's own: {max_rel_ref:.3e})")
        assert cos_ref >= 0.999 and cos64 >= 0.999
        assert max_rel <= max(2.0 * max_rel_ref, 2e-2)

    # ignored rows contribute exactly zero to dX
    ign = labels == -100
    if ign.any():
        assert torch.count_nonzero(xg.grad[ign]) == 0


def test_flce_lse_and_round_flag(cuda_device):
    from llm_fine_tune_distributed_b200 import ops

    N, H, V = 384, 256, 2048
    x, w, labels = _case(N, H, V, 0.2, seed=11, device=cuda_device)
    logits_bf16 = torch.nn.functional.linear(x, w).float()
    ref_lse = torch.logsumexp(logits_bf16, dim=1)
    loss, lse, n_valid = ops.flce_forward(x, w, labels.to(torch.int32), reduction_mean=True,
                                          flags=ops.FLAG_ROUND_LOGITS_BF16)
    assert int(n_valid) == int((labels != -100).sum())
    assert torch.allclose(lse, ref_lse, atol=1e-3, rtol=0)
    ref = oracle.lm_head_ce(x, w, labels, None)
    # with the reference's bf16 logit rounding emulated the loss agrees far inside the 1e-3 bar
    assert abs(loss.item() - ref.item()) <= 2e-5 * abs(ref.item())


def test_flce_all_rows_ignored(cuda_device):
    """All-masked batch: sum-reduction gives 0 loss and zero grads; mean gives NaN (torch semantics)."""
    from llm_fine_tune_distributed_b200 import ops

    N, H, V = 128, 64, 512
    x, w, _ = _case(N, H, V, 0.0, seed=3, device=cuda_device)
    labels = torch.full((N,), -100, dtype=torch.int32, device=cuda_device)
    denom = torch.tensor(7, dtype=torch.int64, device=cuda_device)
    xg, wg = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    loss = ops.fused_linear_cross_entropy(xg, wg, labels, denom)
    loss.backward()
    assert loss.item() == 0.0
    assert torch.count_nonzero(xg.grad) == 0 and torch.count_nonzero(wg.grad) == 0
    loss_mean, _, _ = ops.flce_forward(x, w, labels, reduction_mean=True)
    assert torch.isnan(loss_mean)


def test_flce_slabbed_backward_matches_single_slab(cuda_device):
    """Vocabulary-slabbed dlogits scratch (fp32 dX accumulation across slabs) vs one slab."""
    from llm_fine_tune_distributed_b200 import ops

    N, H, V = 640, 128, 1536
    x, w, labels = _case(N, H, V, 0.4, seed=21, device=cuda_device)
    lab = labels.to(torch.int32)
    loss, lse, n_valid = ops.flce_forward(x, w, lab, reduction_mean=True)
    g = torch.ones((), device=cuda_device)
    dx1, dw1 = ops.flce_backward(x, w, lab, lse, g, denom=n_valid, chunk_cols=2048)   # one slab (V = 1536)
    dx2, dw2 = ops.flce_backward(x, w, lab, lse, g, denom=n_valid, chunk_cols=512)    # three slabs
    dx3, dw3 = ops.flce_backward(x, w, lab, lse, g, denom=n_valid, chunk_cols=1024)   # two balanced slabs of 768
    for dx, dw in ((dx2, dw2), (dx3, dw3)):
        assert torch.equal(dw1, dw)            # every dW slab is produced by one GEMM either way
        assert _cos(dx1, dx) > 0.99999         # dX: fp32 accumulation across slabs, one bf16 rounding
        assert torch.allclose(dx1.float(), dx.float(), rtol=2.0 ** -7, atol=1e-6)


def test_flce_rejects_cpu_tensors():
    from llm_fine_tune_distributed_b200 import ops
    from llm_fine_tune_distributed_b200._cabi import B200SftError

    with pytest.raises(B200SftError):
        ops.flce_forward(torch.zeros(8, 8, dtype=torch.bfloat16), torch.zeros(8, 8, dtype=torch.bfloat16),
                         torch.zeros(8, dtype=torch.int32))


@pytest.mark.parametrize("N,H,V,masked", [(512, 128, 2048, 0.0), (640, 256, 4096, 0.5), (1000, 72, 1000, 0.9),
                                          (384, 64, 512, 0.997)])
def test_flce_skip_ignored_rows_matches_dense_path_and_oracle(cuda_device, N, H, V, masked):
    """Valid-row compaction (the live count never leaves the device) must not change the result."""
    from llm_fine_tune_distributed_b200 import ops

    x, w, labels = _case(N, H, V, masked, seed=7 * N + V, device=cuda_device)
    lab = labels.to(torch.int32)
    n_valid = int((labels != -100).sum())
    denom = torch.tensor(max(n_valid, 1), dtype=torch.int64, device=cuda_device)
    outs = []
    for skip in (False, True):
        xg, wg = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
        loss = ops.fused_linear_cross_entropy(xg, wg, lab, denom, skip_ignored_rows=skip)
        loss.backward()
        outs.append((loss.detach(), xg.grad, wg.grad))
    torch.cuda.synchronize()
    (l0, dx0, dw0), (l1, dx1, dw1) = outs
    assert abs(l0.item() - l1.item()) <= 1e-5 * max(abs(l0.item()), 1e-6)
    # each dX row depends on its own row only; the dense and the compacted GEMM may cut K into a different
    # number of split-K pieces (they have different tile counts), so the fp32 sums can differ in the last bit
    assert torch.allclose(dx0.float(), dx1.float(), rtol=2.0 ** -7, atol=1e-7)
    assert n_valid == 0 or _cos(dx0, dx1) > 0.99999
    assert torch.count_nonzero(dx1[labels == -100]) == 0
    assert torch.allclose(dw0.float(), dw1.float(), rtol=2.0 ** -7, atol=1e-6)   # same sum, other order over rows
    ref_loss, ref_dx, ref_dw = oracle.lm_head_ce_with_grads(x, w, labels, float(denom))
    assert abs(l1.item() - ref_loss.item()) <= 1e-3 * abs(ref_loss.item()) + 1e-7
    if n_valid > 0:
        assert _cos(dx1, ref_dx) >= 0.999 and _cos(dw1, ref_dw) >= 0.999


def test_flce_skip_ignored_rows_all_masked(cuda_device):
    from llm_fine_tune_distributed_b200 import ops

    N, H, V = 256, 64, 512
    x, w, _ = _case(N, H, V, 0.0, seed=5, device=cuda_device)
    lab = torch.full((N,), -100, dtype=torch.int32, device=cuda_device)
    xg, wg = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    loss = ops.fused_linear_cross_entropy(xg, wg, lab, torch.tensor(3, dtype=torch.int64, device=cuda_device),
                                          skip_ignored_rows=True)
    loss.backward()
    assert loss.item() == 0.0
    assert torch.count_nonzero(xg.grad) == 0 and torch.count_nonzero(wg.grad) == 0


@pytest.mark.parametrize("N,masked", [(1, 0.0), (130, 0.3), (8192, 0.6), (5000, 1.0)])
def test_compact_and_scatter_rows_bit_exact(cuda_device, N, masked):
    from llm_fine_tune_distributed_b200 import ops

    H = 64
    g = torch.Generator().manual_seed(N)
    x = torch.randn(N, H, generator=g).to(torch.bfloat16).to(cuda_device)
    labels = torch.randint(0, 1000, (N,), generator=g, dtype=torch.int32)
    labels[torch.rand(N, generator=g) < masked] = -100
    xc, labels_c, row_index, pos, n_live = ops.compact_rows(x, labels.to(cuda_device))
    live, ref_pos, n = oracle.compact_rows_numpy(labels.numpy())
    assert int(n_live) == n
    assert np.array_equal(row_index.cpu().numpy()[:n], live)
    assert np.array_equal(pos.cpu().numpy(), ref_pos)
    assert np.array_equal(labels_c.cpu().numpy()[:n], labels.numpy()[live])
    assert bool((labels_c[n:] == -100).all())
    assert torch.equal(xc[:n].cpu(), x.cpu()[torch.from_numpy(live)])
    pad = min((n + 127) // 128 * 128, N)
    assert torch.count_nonzero(xc[n:pad]) == 0
    back = ops.scatter_rows(xc, pos)
    expect = x.clone()
    expect[labels.to(cuda_device) == -100] = 0
    assert torch.equal(back, expect)


@pytest.mark.parametrize("N,H,V,masked,skip", [(256, 128, 1024, 0.0, False), (512, 64, 1000, 0.5, True),
                                               (384, 256, 4096, 0.8, True)])
def test_flce_token_stats_vs_oracle(cuda_device, N, H, V, masked, skip):
    """Logits-free token accuracy / entropy vs the same metrics on materialised logits."""
    from llm_fine_tune_distributed_b200 import ops

    x, w, labels = _case(N, H, V, masked, seed=N + 3 * V, device=cuda_device)
    x = x * 4  # sharpen the distribution so that some argmaxes hit their label
    top = torch.nn.functional.linear(x, w).float().argmax(-1)
    pick = torch.rand(N, generator=torch.Generator().manual_seed(1)).to(cuda_device) < 0.3
    labels = torch.where(pick & (labels != -100), top, labels)   # make ~30 % of the valid rows "correct"
    lab = labels.to(torch.int32)
    # the reference's argmax sees bf16-rounded logits; emulate that rounding so "is the maximum" means the same
    loss, stats = ops.fused_linear_cross_entropy(x, w, lab, None, skip_ignored_rows=skip, return_stats=True,
                                                 flags=ops.FLAG_ROUND_LOGITS_BF16)
    n_correct, ent = oracle.token_stats(x, w, labels)
    n_valid = int((labels != -100).sum())
    assert int(stats["n_valid"]) == n_valid
    assert abs(int(stats["n_correct"]) - n_correct) <= max(1, n_valid // 200)   # only exact ties may differ
    assert abs(float(stats["entropy_sum"]) - ent) <= 1e-3 * abs(ent) + 1e-3
    ref = oracle.lm_head_ce(x, w, labels, None)
    assert abs(loss.item() - ref.item()) <= 1e-3 * abs(ref.item())


def test_flce_is_bitwise_deterministic(cuda_device):
    """No atomics on floating-point data anywhere in the path: two runs give identical bits."""
    from llm_fine_tune_distributed_b200 import ops

    x, w, labels = _case(1536, 256, 8192, 0.4, seed=99, device=cuda_device)
    lab = labels.to(torch.int32)
    denom = torch.tensor(int((labels != -100).sum()), dtype=torch.int64, device=cuda_device)
    runs = []
    for _ in range(2):
        for skip in (False, True):
            xg, wg = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
            loss = ops.fused_linear_cross_entropy(xg, wg, lab, denom, skip_ignored_rows=skip, chunk_cols=2048)
            loss.backward()
            runs.append((skip, loss.detach().clone(), xg.grad.clone(), wg.grad.clone()))
    torch.cuda.synchronize()
    for (s0, l0, dx0, dw0), (s1, l1, dx1, dw1) in ((runs[0], runs[2]), (runs[1], runs[3])):
        assert s0 == s1 and torch.equal(l0, l1) and torch.equal(dx0, dx1) and torch.equal(dw0, dw1)


def test_flce_random_shape_sweep(cuda_device):
    """Seeded sweep over ragged shapes (rows, hidden and vocabulary not multiples of any tile size),
    mask densities, slab widths, compaction on/off and both engine variants, against the oracle."""
    import random

    from llm_fine_tune_distributed_b200 import _cabi, ops

    rng = random.Random(1234)
    try:
        for i in range(24):
            N = rng.choice([1, 7, 127, 129, 255, 257, 500, 777])
            H = 8 * rng.randint(1, 40)
            V = 8 * rng.randint(2, 300)
            masked = rng.choice([0.0, 0.3, 0.9])
            skip = rng.random() < 0.5
            chunk = rng.choice([0, 256, 512])
            _cabi.set_cta_group(rng.choice([1, 2]))
            x, w, labels = _case(N, H, V, masked, seed=1000 + i, device=cuda_device)
            n_valid = int((labels != -100).sum())
            denom = torch.tensor(max(n_valid, 1), dtype=torch.int64, device=cuda_device)
            xg, wg = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
            loss = ops.fused_linear_cross_entropy(xg, wg, labels.to(torch.int32), denom, skip_ignored_rows=skip,
                                                  chunk_cols=chunk)
            loss.backward()
            ref_loss, ref_dx, ref_dw = oracle.lm_head_ce_with_grads(x, w, labels, float(denom))
            torch.cuda.synchronize()
            tag = f"case {i}: N={N} H={H} V={V} masked={masked} skip={skip} chunk={chunk}"
            assert abs(loss.item() - ref_loss.item()) <= 1e-3 * abs(ref_loss.item()) + 1e-7, tag
            if n_valid > 0:
                assert _cos(xg.grad, ref_dx) >= 0.999 and _cos(wg.grad, ref_dw) >= 0.999, tag
            else:
                assert torch.count_nonzero(xg.grad) == 0 and torch.count_nonzero(wg.grad) == 0, tag
            assert torch.isfinite(xg.grad.float()).all() and torch.isfinite(wg.grad.float()).all(), tag
    finally:
        _cabi.set_cta_group(2)


def test_flce_backward_skips_unneeded_gradients(cuda_device):
    """Frozen head (no dW) or frozen trunk (no dX): the corresponding GEMM is skipped and the other
    gradient is unchanged."""
    from llm_fine_tune_distributed_b200 import ops

    x, w, labels = _case(512, 128, 2048, 0.3, seed=8, device=cuda_device)
    lab = labels.to(torch.int32)
    denom = torch.tensor(int((labels != -100).sum()), dtype=torch.int64, device=cuda_device)
    xg, wg = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    ops.fused_linear_cross_entropy(xg, wg, lab, denom).backward()
    for skip in (False, True):
        x1 = x.clone().requires_grad_(True)
        ops.STATS.reset()
        ops.fused_linear_cross_entropy(x1, w, lab, denom, skip_ignored_rows=skip).backward()      # head frozen
        launches_dx_only = ops.STATS.kernel_launches
        assert torch.allclose(x1.grad.float(), xg.grad.float(), rtol=2.0 ** -7, atol=1e-7)
        w1 = w.clone().requires_grad_(True)
        ops.STATS.reset()
        ops.fused_linear_cross_entropy(x, w1, lab, denom, skip_ignored_rows=skip).backward()      # only the head trains
        assert torch.allclose(w1.grad.float(), wg.grad.float(), rtol=2.0 ** -7, atol=1e-7)
        assert ops.STATS.kernel_launches < launches_dx_only + 2


@pytest.mark.parametrize("gain", [30.0, 300.0])
def test_flce_extreme_logits_stay_finite_and_accurate(cuda_device, gain):
    """Logits in the hundreds / thousands (peaked softmax, exp underflow everywhere but the maximum):
    the online max/sum-exp must neither overflow nor lose the target term."""
    from llm_fine_tune_distributed_b200 import ops

    x, w, labels = _case(384, 128, 4096, 0.2, seed=41, device=cuda_device)
    x = (x.float() * gain).to(torch.bfloat16)
    lab = labels.to(torch.int32)
    denom = torch.tensor(int((labels != -100).sum()), dtype=torch.int64, device=cuda_device)
    xg, wg = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    loss = ops.fused_linear_cross_entropy(xg, wg, lab, denom, flags=ops.FLAG_ROUND_LOGITS_BF16)
    loss.backward()
    ref_loss, ref_dx, ref_dw = oracle.lm_head_ce_with_grads(x, w, labels, float(denom))
    assert torch.isfinite(loss) and torch.isfinite(xg.grad.float()).all() and torch.isfinite(wg.grad.float()).all()
    assert abs(loss.item() - ref_loss.item()) <= 1e-3 * abs(ref_loss.item())
    assert _cos(xg.grad, ref_dx) >= 0.999 and _cos(wg.grad, ref_dw) >= 0.999


@pytest.mark.parametrize("skip", [False, True])
def test_flce_empty_batch_matches_torch(cuda_device, skip):
    """Zero rows: torch gives loss 0 for reduction="sum" (nan for "mean"), an empty dX and a zero dW."""
    from llm_fine_tune_distributed_b200 import ops

    H, V = 64, 512
    x = torch.empty(0, H, dtype=torch.bfloat16, device=cuda_device, requires_grad=True)
    w = torch.randn(V, H, device=cuda_device).to(torch.bfloat16).requires_grad_(True)
    labels = torch.empty(0, dtype=torch.int32, device=cuda_device)
    denom = torch.tensor(5, dtype=torch.int64, device=cuda_device)
    loss = ops.fused_linear_cross_entropy(x, w, labels, denom, skip_ignored_rows=skip)
    loss.backward()
    ref = oracle.lm_head_ce(x.detach(), w.detach(), labels.long(), 5.0)
    assert loss.item() == ref.item() == 0.0
    assert x.grad.shape == (0, H) and torch.count_nonzero(w.grad) == 0
    mean = ops.fused_linear_cross_entropy(x.detach(), w.detach(), labels, None, skip_ignored_rows=skip)
    assert torch.isnan(mean) and torch.isnan(oracle.lm_head_ce(x.detach(), w.detach(), labels.long(), None))


@pytest.mark.parametrize("bad", [512, 100000, -7])
def test_flce_out_of_range_label_poisons_the_loss(cuda_device, bad):
    """torch refuses a target outside [0,V) (IndexError on CPU, device assert on CUDA).  The fused
    path cannot raise from a kernel; it must not return a finite, wrong loss either."""
    from llm_fine_tune_distributed_b200 import ops

    N, H, V = 256, 64, 512
    x = torch.randn(N, H, device=cuda_device).to(torch.bfloat16)
    w = torch.randn(V, H, device=cuda_device).to(torch.bfloat16)
    labels = torch.randint(0, V, (N,), device=cuda_device, dtype=torch.int32)
    good = ops.fused_linear_cross_entropy(x, w, labels, None)
    assert torch.isfinite(good)
    labels[17] = bad
    with pytest.raises((IndexError, RuntimeError)):
        oracle.lm_head_ce(x.cpu(), w.cpu(), labels.cpu().long(), None)
    for skip in (False, True):
        assert torch.isnan(ops.fused_linear_cross_entropy(x, w, labels, None, skip_ignored_rows=skip))
