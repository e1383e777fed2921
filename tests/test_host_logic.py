"""Host-side mirror of the reference interface, on CPU: the lm_head swap keeps names, tying and
unarmed behaviour; compute_loss follows Trainer.compute_loss's conventions; driver helpers follow
training.py."""
import os

import numpy as np
import pytest
import torch

import _cpu_standins
from oracle import sft_oracle as oracle

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def tiny():
    from transformers import SmolLM3Config, SmolLM3ForCausalLM

    cfg = SmolLM3Config(num_hidden_layers=2, hidden_size=64, intermediate_size=128, num_attention_heads=4,
                        num_key_value_heads=1, vocab_size=512, max_position_embeddings=128, tie_word_embeddings=True,
                        pad_token_id=0, bos_token_id=1, eos_token_id=2)
    cfg._attn_implementation = "eager"
    torch.manual_seed(0)  # same construction as tests/golden/make_golden.py -> same weights
    return SmolLM3ForCausalLM(cfg).to(torch.bfloat16).eval()


def test_fused_head_keeps_names_tying_and_unarmed_forward(tiny):
    from llm_fine_tune_distributed_b200 import FusedLMHead, install_fused_head

    ids = torch.randint(0, 512, (2, 8), generator=torch.Generator().manual_seed(1))
    before = tiny(input_ids=ids).logits
    keys = list(tiny.state_dict().keys())
    install_fused_head(tiny)
    install_fused_head(tiny)  # idempotent
    assert isinstance(tiny.lm_head, FusedLMHead)
    assert list(tiny.state_dict().keys()) == keys
    assert tiny.lm_head.weight is tiny.model.embed_tokens.weight
    assert torch.equal(tiny(input_ids=ids).logits, before)           # unarmed: an ordinary linear head
    stock = tiny(input_ids=ids, labels=ids).loss                      # unarmed + labels: stock ForCausalLMLoss
    assert torch.isfinite(stock)


def test_compute_loss_seam_reproduces_trainer_compute_loss(tiny, monkeypatch):
    """Armed path on CPU (kernels replaced by the oracle stand-ins) vs the golden values produced by
    the real Trainer.compute_loss: same loss, x num_processes, logits withheld."""
    from llm_fine_tune_distributed_b200 import FusedCELossMixin, install_fused_head

    _cpu_standins.install(monkeypatch)
    install_fused_head(tiny)
    d = np.load(os.path.join(G, "smollm3_tiny_step.npz"))

    class Seam(FusedCELossMixin):
        num_processes = 2
        model_accepts_loss_kwargs = True
        args = None

    seam = Seam()
    n = torch.tensor(int(d["num_items"]))
    for i in range(2):
        inputs = {"input_ids": torch.from_numpy(d[f"input_ids_{i}"]), "labels": torch.from_numpy(d[f"labels_{i}"])}
        loss, outputs = seam.compute_loss(tiny, inputs, return_outputs=True, num_items_in_batch=n)
        ref = float(d[f"compute_loss_{i}"])
        assert abs(float(loss) - ref) <= 1e-6 * abs(ref)
        # no logits anywhere in what a caller can read: prediction_step builds its `logits` from outputs.items()
        assert "logits" not in outputs.keys() and outputs.logits is None
        assert all(v is not None and tuple(v.shape) != tuple(inputs["input_ids"].shape) + (tiny.config.hidden_size,)
                   for v in outputs.to_tuple() if torch.is_tensor(v))
    # num_items_in_batch=None -> reduction="mean", no x num_processes (trainer.py:2013-2019)
    inputs = {"input_ids": torch.from_numpy(d["input_ids_0"]), "labels": torch.from_numpy(d["labels_0"])}
    loss = seam.compute_loss(tiny, inputs)
    assert abs(float(loss) - float(d["model_loss_mean_0"])) <= 1e-6 * abs(float(d["model_loss_mean_0"]))
    assert not tiny.lm_head.armed  # the context manager disarms the head again


def test_freeze_policy_matches_reference_counts():
    """training.py:114-145 on the real 3B config (meta tensors): 418,914,304 of 3,075,098,624
    parameters trainable (claude.md:242-243 of the reference)."""
    from transformers import SmolLM3Config, SmolLM3ForCausalLM

    from llm_fine_tune_distributed_b200.training import apply_freeze_policy

    with torch.device("meta"):
        model = SmolLM3ForCausalLM(SmolLM3Config(tie_word_embeddings=True))
    model.tie_weights()
    trainable, total = apply_freeze_policy(model)
    assert (trainable, total) == (418_914_304, 3_075_098_624)
    assert model.model.embed_tokens.weight.requires_grad  # tied to lm_head -> backward crosses all layers


def test_collate_and_settings(monkeypatch):
    from llm_fine_tune_distributed_b200.training import ByteChatTokenizer, SFTSettings, collate, format_prompt

    batch = collate([[5, 6, 7], [8, 9, 10, 11, 12, 13]], pad_id=0, max_len=5)
    assert batch["input_ids"].tolist() == [[5, 6, 7, 0, 0], [8, 9, 10, 11, 12]]
    assert batch["attention_mask"].tolist() == [[1, 1, 1, 0, 0], [1, 1, 1, 1, 1]]
    assert batch["labels"].tolist() == [[5, 6, 7, -100, -100], [8, 9, 10, 11, 12]]
    msgs = format_prompt({"full-question": "q?", "answer": "a."})["messages"]
    assert [m["role"] for m in msgs] == ["system", "user", "assistant"]
    ids, mask = ByteChatTokenizer(128256).apply_chat_template(msgs, return_assistant_mask=True)
    assert ids[-1] == 1 and max(ids) < 128256 and len(mask) == len(ids)
    # the reference's full system prompt (training.py:176-186) must leave room for the answer inside max_seq_length:
    # a sample cut off inside the prompt would have no assistant token left for completion-only loss
    assert msgs[0]["content"].startswith("You are a wilderness survival") and "Essential Knots" in msgs[0]["content"]
    assert 300 < len(ids) < 800 and sum(mask[:1024]) >= 3

    monkeypatch.setenv("BATCH_SIZE", "28")
    monkeypatch.setenv("LEARNING_RATE", "1e-4")
    monkeypatch.setenv("EPOCHS", "2")
    s = SFTSettings.from_env(world_size=2, local_rank=1)
    assert s.per_device_train_batch_size == 28 and s.learning_rate == 2e-4  # lr x world, training.py:263
    assert s.ddp_backend == "nccl" and s.ddp_bucket_cap_mb == 50 and not s.ddp_find_unused_parameters
    assert s.gradient_accumulation_steps == 4 and s.max_seq_length == 1024 and s.max_grad_norm == 1.0


def test_flattened_collator_matches_transformers():
    """collate_flattened vs the real DataCollatorWithFlattening, key for key."""
    from transformers import DataCollatorWithFlattening

    from llm_fine_tune_distributed_b200.training import collate_flattened

    samples = [[5, 6, 7, 8], [9, 10], [11, 12, 13, 14, 15, 16, 17]]
    ref = DataCollatorWithFlattening(return_flash_attn_kwargs=True)([{"input_ids": s} for s in samples], return_tensors="pt")
    got = collate_flattened(samples, max_len=100)
    assert set(got) == set(ref)
    for k in ref:
        if torch.is_tensor(ref[k]):
            assert got[k].dtype == ref[k].dtype and torch.equal(got[k], ref[k]), k
        else:
            assert got[k] == ref[k], k
    assert got["labels"].tolist() == [[-100, 6, 7, 8, -100, 10, -100, 12, 13, 14, 15, 16, 17]]


def test_completion_only_masks():
    from llm_fine_tune_distributed_b200.training import ByteChatTokenizer, collate, collate_flattened, format_prompt

    tok = ByteChatTokenizer(128256)
    msgs = format_prompt({"full-question": "q?", "answer": "an answer"})["messages"]
    ids, mask = tok.apply_chat_template(msgs, return_assistant_mask=True)
    assert ids == tok.apply_chat_template(msgs) and len(ids) == len(mask)
    n_assistant = len(tok.encode("an answer<|im_end|>\n")) + 1        # assistant body + eos
    assert sum(mask) == n_assistant and mask[-1] == 1 and mask[0] == 0
    batch = collate([ids], pad_id=0, max_len=4096, masks=[mask])
    assert int((batch["labels"] != -100).sum()) == n_assistant
    flat = collate_flattened([ids, ids], 4096, masks=[mask, mask])
    assert int((flat["labels"] != -100).sum()) == 2 * n_assistant and flat["position_ids"][0, len(ids)] == 0


def test_device_label_collators_carry_what_the_frontend_needs():
    """device_labels=True: no labels leave the host; masks + the front-end's rule reproduce exactly the labels
    the host collators write (padded and padding-free, with and without completion masks)."""
    from llm_fine_tune_distributed_b200.training import collate, collate_flattened

    rng = np.random.default_rng(0)
    samples = [rng.integers(3, 900, size=n).tolist() for n in (7, 1, 12, 5)]
    masks = [(rng.random(len(s)) > 0.4).astype(int).tolist() for s in samples]
    for m in (None, masks):
        host, dev = collate(samples, 0, 10, masks=m), collate(samples, 0, 10, masks=m, device_labels=True)
        assert "labels" not in dev and ("completion_mask" in dev) == (m is not None)
        got = oracle.collator_labels_numpy(dev["input_ids"].numpy(), None, dev["attention_mask"].numpy(),
                                           dev["completion_mask"].numpy() if m else None)
        np.testing.assert_array_equal(got, host["labels"].numpy())
        host, dev = collate_flattened(samples, 10, m), collate_flattened(samples, 10, m, device_labels=True)
        assert "labels" not in dev and set(host) - set(dev) == {"labels"}
        got = oracle.collator_labels_numpy(dev["input_ids"].numpy(), None, None,
                                           dev["completion_mask"].numpy() if m else None, dev["position_ids"].numpy())
        np.testing.assert_array_equal(got, host["labels"].numpy())


def test_compute_loss_seam_takes_a_label_plan(tiny, monkeypatch):
    """A batch that arrives with ``b200sft_plan`` (labels built by the front-end ahead of the forward) gives the
    loss of the plain call; the plan never reaches the model's forward."""
    from llm_fine_tune_distributed_b200 import FusedCELossMixin, install_fused_head, ops

    _cpu_standins.install(monkeypatch)
    install_fused_head(tiny)
    d = np.load(os.path.join(G, "smollm3_tiny_step.npz"))

    class Seam(FusedCELossMixin):
        num_processes = 1
        model_accepts_loss_kwargs = True
        args = None

    ids, labels = torch.from_numpy(d["input_ids_0"]), torch.from_numpy(d["labels_0"])
    plan = ops.label_frontend(labels=labels, return_labels=True)
    seam = Seam()
    n = torch.tensor(int(d["num_items"]))
    a = seam.compute_loss(tiny, {"input_ids": ids, "labels": labels}, num_items_in_batch=n)
    b = seam.compute_loss(tiny, {"input_ids": ids, "labels": plan.labels, "b200sft_plan": plan}, num_items_in_batch=n)
    assert float(a) == float(b) and tiny.lm_head.plan is None
    wrong = ops.label_frontend(labels=labels[:, :-1].contiguous())
    with pytest.raises(ValueError):
        seam.compute_loss(tiny, {"input_ids": ids, "labels": labels, "b200sft_plan": wrong}, num_items_in_batch=n)


def test_fused_final_norm_keeps_names_and_the_seam_result(tiny, monkeypatch):
    """model.model.norm swapped for the bypassable stand-in: same state-dict keys, same unarmed forward, and
    the armed compute_loss (kernels replaced by the oracle stand-ins) still reproduces the golden values
    of the real Trainer.compute_loss -- now with the norm applied inside the loss."""
    import copy

    from llm_fine_tune_distributed_b200 import FusedCELossMixin, install_fused_head
    from llm_fine_tune_distributed_b200.loss import FusedFinalNorm

    _cpu_standins.install(monkeypatch)
    d = np.load(os.path.join(G, "smollm3_tiny_step.npz"))
    m = copy.deepcopy(tiny)
    keys = set(m.state_dict().keys())
    ids = torch.from_numpy(d["input_ids_0"])
    with torch.no_grad():
        before = m(ids).logits
    install_fused_head(m, fuse_final_norm=True)
    assert isinstance(m.model.norm, FusedFinalNorm) and set(m.state_dict().keys()) == keys
    assert m.model.norm.weight is dict(m.named_parameters())["model.norm.weight"]
    with torch.no_grad():
        assert torch.equal(m(ids).logits, before)

    class Seam(FusedCELossMixin):
        num_processes = 2
        model_accepts_loss_kwargs = True
        args = None
        b200sft_fuse_final_norm = True

    loss = Seam().compute_loss(m, {"input_ids": ids, "labels": torch.from_numpy(d["labels_0"])},
                               num_items_in_batch=torch.tensor(int(d["num_items"])))
    ref = float(d["compute_loss_0"])
    assert abs(float(loss) - ref) <= 1e-6 * abs(ref)
    assert not m.model.norm.armed and not m.lm_head.armed


def test_patched_model_saves_and_reloads_as_stock(tiny, tmp_path):
    """The head swap must not leak into checkpoints: save_pretrained on the patched model, then a
    stock from_pretrained, gives the same logits (training.py:309-312 saves the best model this way)."""
    from transformers import SmolLM3ForCausalLM

    from llm_fine_tune_distributed_b200 import install_fused_head

    install_fused_head(tiny)
    ids = torch.randint(0, 512, (1, 6), generator=torch.Generator().manual_seed(2))
    before = tiny(input_ids=ids).logits
    tiny.save_pretrained(tmp_path / "ckpt")
    again = SmolLM3ForCausalLM.from_pretrained(tmp_path / "ckpt", dtype=torch.bfloat16, attn_implementation="eager").eval()
    assert type(again.lm_head) is torch.nn.Linear
    assert again.lm_head.weight.data_ptr() == again.model.embed_tokens.weight.data_ptr()
    sd0, sd1 = tiny.state_dict(), again.state_dict()
    assert list(sd0) == list(sd1) and all(torch.equal(sd0[k], sd1[k]) for k in sd0)
    # (logits agree to bf16 resolution only: the fixture's .to(bfloat16) also cast the rotary inv_freq buffer,
    # which a fresh from_pretrained keeps in fp32 -- stock transformers behaviour, independent of the head swap)
    assert torch.allclose(again(input_ids=ids).logits.float(), before.float(), atol=2e-2)


def test_split_sizes_match_the_reference_run():
    """2,845 QA rows -> 2,560 train / 285 validation, the sizes the reference reports for
    train_test_split(test_size=0.1, seed=42) (training.py:164; docs/AIM_WORKFLOW.md:47-54)."""
    from llm_fine_tune_distributed_b200.training import load_qa_rows, train_test_split

    path = "/root/reference/data/qa_dataset.parquet"
    rows = load_qa_rows(path) if os.path.exists(path) else [{"full-question": str(i), "answer": "a"} for i in range(2845)]
    assert len(rows) == 2845 and set(rows[0]) == {"full-question", "answer"}
    train, val = train_test_split(rows, 0.1, 42)
    assert (len(train), len(val)) == (2560, 285)
    assert train_test_split(rows, 0.1, 42)[1] == val          # deterministic in the seed
