"""Pins the oracle (oracle/sft_oracle.py) against golden outputs of the real third-party code the
reference runs (tests/golden/make_golden.py generated them with transformers 5.5.0 / torch 2.11.0).
Integer work must be bit-exact; the floating-point restatement must reproduce the bf16 path to
fp32 round-off, and the independent float64 restatement to bf16 resolution."""
import os

import numpy as np
import torch

from oracle import sft_oracle as oracle

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _bf16(a):
    return torch.from_numpy(a).to(torch.bfloat16)


def test_causal_lm_loss_sum_and_mean_match_transformers():
    d = np.load(os.path.join(G, "causal_lm_loss.npz"))
    hidden, weight = _bf16(d["hidden"]), _bf16(d["weight"])
    labels = torch.from_numpy(d["labels"])
    B, T, H = hidden.shape
    shifted = oracle.shift_labels(labels).reshape(-1)
    for tag, n in (("sum", float(d["num_items_sum"])), ("mean", None)):
        loss, dh, dw = oracle.lm_head_ce_with_grads(hidden.reshape(B * T, H), weight, shifted, n)
        assert abs(float(loss) - float(d[f"loss_{tag}"])) <= 1e-6 * abs(float(d[f"loss_{tag}"]))
        assert torch.equal(dh.float().reshape(B, T, H), torch.from_numpy(d[f"dhidden_{tag}"]))
        assert torch.equal(dw.float(), torch.from_numpy(d[f"dweight_{tag}"]))


def test_float64_restatement_agrees_with_transformers_to_bf16_resolution():
    d = np.load(os.path.join(G, "causal_lm_loss.npz"))
    B, T, H = d["hidden"].shape
    shifted = oracle.shift_labels_numpy(d["labels"]).reshape(-1)
    loss, lse, dh, dw = oracle.lm_head_ce_float64(d["hidden"].reshape(B * T, H), d["weight"], shifted,
                                                  float(d["num_items_sum"]), round_logits_bf16=True)
    assert abs(loss - float(d["loss_sum"])) <= 1e-5 * abs(float(d["loss_sum"]))
    for got, ref in ((dh.reshape(B, T, H), d["dhidden_sum"]), (dw, d["dweight_sum"])):
        scale = np.abs(ref).max()
        assert np.abs(got - ref).max() <= 2.0 ** -7 * scale  # golden grads were rounded to bf16 twice (dlogits, output)
    assert lse.shape == (B * T,)


def test_shift_and_count_bit_exact():
    d = np.load(os.path.join(G, "smollm3_tiny_step.npz"))
    window = [torch.from_numpy(d[f"labels_{i}"]) for i in range(2)]
    assert oracle.num_items_in_batch(window) == int(d["num_items"])
    for lbl in window:
        ref = oracle.shift_labels(lbl)
        assert np.array_equal(oracle.shift_labels_numpy(lbl.numpy()), ref.numpy())
        assert torch.equal(ref[:, :-1], lbl[:, 1:]) and bool((ref[:, -1] == -100).all())


def test_model_loss_and_compute_loss_match_smollm3_and_trainer():
    """hidden states captured from the real SmolLM3 model + its tied lm_head weight -> the oracle
    must reproduce SmolLM3ForCausalLM's loss and Trainer.compute_loss's (x num_processes = 2)."""
    d = np.load(os.path.join(G, "smollm3_tiny_step.npz"))
    weight = _bf16(d["weight"])
    n = int(d["num_items"])
    for i in range(2):
        hidden, labels = _bf16(d[f"hidden_{i}"]), torch.from_numpy(d[f"labels_{i}"])
        got = oracle.compute_loss_reference(hidden, weight, labels, n, num_processes=1)
        assert abs(float(got) - float(d[f"model_loss_{i}"])) <= 1e-6 * abs(float(d[f"model_loss_{i}"]))
        got2 = oracle.compute_loss_reference(hidden, weight, labels, n, num_processes=2)
        assert abs(float(got2) - float(d[f"compute_loss_{i}"])) <= 1e-6 * abs(float(d[f"compute_loss_{i}"]))
    mean = oracle.compute_loss_reference(_bf16(d["hidden_0"]), weight, torch.from_numpy(d["labels_0"]), None)
    assert abs(float(mean) - float(d["model_loss_mean_0"])) <= 1e-6 * abs(float(d["model_loss_mean_0"]))


def test_allreduce_hook_bucket_matches_default_hook():
    d = np.load(os.path.join(G, "allreduce_hook.npz"))
    world = int(d["world"])
    buckets = [_bf16(d[f"bucket_rank{r}"]) for r in range(world)]
    # emulate the collective: every rank divides first (bf16), then the sum is taken in bf16
    divided = [b.clone().div_(world) for b in buckets]
    total = divided[0].clone()
    for b in divided[1:]:
        total += b
    for r in range(world):
        got = oracle.allreduce_hook_bucket(buckets[r].clone(), world, lambda t: t.copy_(total))
        assert torch.equal(got.float(), torch.from_numpy(d[f"result_rank{r}"]))


def test_rms_norm_restatement_is_the_transformers_module():
    """oracle.rms_norm against the real SmolLM3RMSNorm (forward and autograd), bit for bit."""
    from transformers.models.smollm3.modeling_smollm3 import SmolLM3RMSNorm

    torch.manual_seed(0)
    mod = SmolLM3RMSNorm(96, eps=1e-6).to(torch.bfloat16)
    with torch.no_grad():
        mod.weight.add_(0.1 * torch.randn(96).to(torch.bfloat16))
    x = (torch.randn(5, 7, 96) * 2).to(torch.bfloat16)
    a = x.clone().requires_grad_(True)
    b = x.clone().requires_grad_(True)
    w = mod.weight.detach().clone().requires_grad_(True)
    ya, yb = mod(a), oracle.rms_norm(b, w, mod.variance_epsilon)
    assert torch.equal(ya, yb)
    g = torch.randn_like(ya)
    ya.backward(g)
    yb.backward(g)
    assert torch.equal(a.grad, b.grad) and torch.equal(mod.weight.grad, w.grad)


def test_chunked_restatement_equals_the_literal_path():
    """The row-blocked form used at sizes where [N,V] fp32 logits do not fit (N = 65,536) against the golden
    outputs of the real ForCausalLMLoss: loss to fp32 round-off, dX bit for bit when a block boundary falls
    inside the batch (rows are independent), dW to one bf16 rounding (fp32 sum across blocks, other order)."""
    d = np.load(os.path.join(G, "causal_lm_loss.npz"))
    hidden, weight = _bf16(d["hidden"]), _bf16(d["weight"])
    B, T, H = hidden.shape
    shifted = oracle.shift_labels(torch.from_numpy(d["labels"])).reshape(-1)
    n = float(d["num_items_sum"])
    ref_dw = torch.from_numpy(d["dweight_sum"])
    for rows in (B * T, 7, 1):
        loss, dh, dw = oracle.lm_head_ce_with_grads_chunked(hidden.reshape(B * T, H), weight, shifted, n, rows_per_chunk=rows)
        assert abs(float(loss) - float(d["loss_sum"])) <= 1e-6 * abs(float(d["loss_sum"]))
        assert torch.equal(dh.float().reshape(B, T, H), torch.from_numpy(d["dhidden_sum"]))
        assert float((dw.float() - ref_dw).abs().max()) <= 2.0 ** -8 * float(ref_dw.abs().max())
