"""The label front-end's CPU restatement (oracle/sft_oracle.py) against the real collator code in
site-packages: transformers' DataCollatorWithFlattening (packed batches) and the ForCausalLMLoss
shift.  Pins the restatement the GPU kernel is compared with in test_gpu_label_frontend.py."""
import numpy as np
import pytest
import torch

from oracle import sft_oracle as oracle


def _ragged(rng, n_seq, lo, hi, vocab=1000):
    return [rng.integers(1, vocab, size=int(rng.integers(lo, hi + 1))).tolist() for _ in range(n_seq)]


@pytest.mark.parametrize("seed,with_labels", [(0, False), (1, True), (2, True)])
def test_packed_labels_match_hf_flattening_collator(seed, with_labels):
    from transformers import DataCollatorWithFlattening

    rng = np.random.default_rng(seed)
    seqs = _ragged(rng, 9, 1, 40)
    feats = []
    for ids in seqs:
        f = {"input_ids": ids}
        if with_labels:  # completion-only samples: prompt part already -100
            cut = int(rng.integers(0, len(ids) + 1))
            f["labels"] = [-100] * cut + ids[cut:]
        feats.append(f)
    hf = DataCollatorWithFlattening(return_tensors="np")(feats)
    flat_labels = None
    if with_labels:  # the un-separated concatenation; the separator comes from position_ids == 0
        flat_labels = np.concatenate([np.asarray(f["labels"]) for f in feats])[None]
    got = oracle.label_frontend_numpy(hf["input_ids"], flat_labels, position_ids=hf["position_ids"])
    np.testing.assert_array_equal(got["labels"], hf["labels"])
    # shift + count as the loss and the trainer do them (loss_utils.py:57-60, trainer.py:2136)
    ref_shift = oracle.shift_labels(torch.from_numpy(hf["labels"])).reshape(-1).numpy()
    np.testing.assert_array_equal(got["shifted"], ref_shift)
    assert got["counts"] == (oracle.num_items_in_batch([torch.from_numpy(hf["labels"])]), int((ref_shift != -100).sum()))


def test_padded_completion_only_labels():
    """Padded batch: labels = input_ids, pads and prompt tokens -> -100 (what the TRL collator does
    with attention_mask / completion_mask), then torch's own shift."""
    rng = np.random.default_rng(3)
    B, T = 5, 33
    ids = rng.integers(1, 500, size=(B, T))
    lens = rng.integers(1, T + 1, size=B)
    att = (np.arange(T)[None, :] < lens[:, None]).astype(np.int64)
    comp = (np.arange(T)[None, :] >= rng.integers(0, T, size=B)[:, None]).astype(np.int64) * att
    got = oracle.label_frontend_numpy(ids, attention_mask=att, completion_mask=comp)
    want = torch.from_numpy(ids).clone()
    want[torch.from_numpy(att) == 0] = -100
    want[torch.from_numpy(comp) == 0] = -100
    np.testing.assert_array_equal(got["labels"], want.numpy())
    shifted = oracle.shift_labels(want).reshape(-1).numpy()
    np.testing.assert_array_equal(got["shifted"], shifted)
    live = np.flatnonzero(shifted != -100)
    np.testing.assert_array_equal(got["row_index"], live)
    np.testing.assert_array_equal(got["labels_c"][: live.size], shifted[live])
    assert (got["labels_c"][live.size:] == -100).all() and got["n_live"] == live.size
    assert (got["pos"][live] == np.arange(live.size)).all() and (np.delete(got["pos"], live) == -1).all()


def test_frontend_restatement_invariants_hypothesis():
    """Properties any label front-end output must have, on random shapes and masks (the size-independent
    checks the GPU test repeats at full size through bit-equality with this restatement)."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=60, deadline=None)
    @given(B=st.integers(1, 6), T=st.integers(1, 70), seed=st.integers(0, 2 ** 31 - 1),
           use_att=st.booleans(), use_comp=st.booleans(), use_pos=st.booleans())
    def check(B, T, seed, use_att, use_comp, use_pos):
        rng = np.random.default_rng(seed)
        ids = rng.integers(0, 1000, size=(B, T))
        att = (rng.random((B, T)) > 0.2).astype(np.int64) if use_att else None
        comp = (rng.random((B, T)) > 0.5).astype(np.int64) if use_comp else None
        posid = np.where(rng.random((B, T)) < 0.15, 0, 1 + rng.integers(0, 9, size=(B, T))) if use_pos else None
        r = oracle.label_frontend_numpy(ids, None, att, comp, posid)
        lab, sh = r["labels"], r["shifted"].reshape(B, T)
        assert (sh[:, -1] == -100).all() and (sh[:, :-1] == lab[:, 1:]).all()           # the shift
        assert r["counts"] == (int((lab != -100).sum()), int((sh != -100).sum()))
        assert r["counts"][1] == r["n_live"] <= r["counts"][0]                              # shifting only loses labels
        live = r["row_index"]
        assert (np.diff(live) > 0).all()                                                    # stable order
        assert (r["pos"][live] == np.arange(live.size)).all() and (r["pos"] >= -1).all()   # inverse maps
        assert int((r["pos"] >= 0).sum()) == r["n_live"]
        assert (r["labels_c"][: r["n_live"]] != -100).all() and (r["labels_c"][r["n_live"]:] == -100).all()
        # every mask only removes labels
        assert ((lab == ids) | (lab == -100)).all()
        if att is not None:
            assert (lab[att == 0] == -100).all()
        if comp is not None:
            assert (lab[comp == 0] == -100).all()
        if posid is not None:
            assert (lab[posid == 0] == -100).all()

    check()
