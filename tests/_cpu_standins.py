"""CPU stand-ins for the two CUDA operators, built from the oracle, so that the host-side wiring
(FusedLMHead, loss_function patch, compute_loss mixin, DDP hook, trainer loop) can be exercised
without a GPU.  Test-only: the product raises on CPU tensors."""
import torch

from oracle import sft_oracle as oracle


def label_prep(labels, ignore_index=-100):
    shifted = oracle.shift_labels(labels, ignore_index)
    counts = torch.tensor([int(labels.ne(ignore_index).sum()), int(shifted.ne(ignore_index).sum())])
    return shifted.reshape(-1).to(torch.int32), counts


def label_frontend(input_ids=None, labels=None, attention_mask=None, completion_mask=None, position_ids=None,
                   ignore_index=-100, compact=True, return_labels=False):
    from llm_fine_tune_distributed_b200.ops import LabelPlan

    np_ = lambda t: None if t is None else t.cpu().numpy()
    r = oracle.label_frontend_numpy(np_(input_ids), np_(labels), np_(attention_mask), np_(completion_mask),
                                    np_(position_ids), ignore_index)
    t = torch.from_numpy
    pos = t(r["pos"])
    row_index = torch.zeros_like(pos)
    row_index[: r["n_live"]] = t(r["row_index"])
    return LabelPlan(t(r["shifted"]), torch.tensor(r["counts"]), t(r["labels"]) if return_labels else None,
                     row_index if compact else None, t(r["labels_c"]) if compact else None, pos if compact else None,
                     torch.tensor(r["n_live"], dtype=torch.int32) if compact else None)


def fused_linear_cross_entropy(hidden, weight, shifted_labels, num_items_in_batch=None, scale=1.0,
                               ignore_index=-100, flags=0, chunk_cols=0, skip_ignored_rows=False, return_stats=False,
                               plan=None, norm_weight=None, norm_eps=1e-6):
    if norm_weight is not None:
        hidden = oracle.rms_norm(hidden, norm_weight, norm_eps)
    n = None if num_items_in_batch is None else float(num_items_in_batch)
    loss = oracle.lm_head_ce(hidden, weight, shifted_labels.long(), n, ignore_index) * scale
    if return_stats:
        nc, ent = oracle.token_stats(hidden.detach(), weight.detach(), shifted_labels, ignore_index)
        nv = int((shifted_labels != ignore_index).sum())
        return loss, {"n_valid": torch.tensor(nv), "n_correct": torch.tensor(nc), "entropy_sum": torch.tensor(ent)}
    return loss


def install(monkeypatch_or_none=None):
    from llm_fine_tune_distributed_b200 import ops

    if monkeypatch_or_none is not None:
        monkeypatch_or_none.setattr(ops, "label_prep", label_prep)
        monkeypatch_or_none.setattr(ops, "label_frontend", label_frontend)
        monkeypatch_or_none.setattr(ops, "fused_linear_cross_entropy", fused_linear_cross_entropy)
    else:
        ops.label_prep = label_prep
        ops.label_frontend = label_frontend
        ops.fused_linear_cross_entropy = fused_linear_cross_entropy
