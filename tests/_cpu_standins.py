"""CPU stand-ins for the two CUDA operators, built from the oracle, so that the host-side wiring
(FusedLMHead, loss_function patch, compute_loss mixin, DDP hook, trainer loop) can be exercised
without a GPU.  Test-only: the product raises on CPU tensors."""
import torch

from oracle import sft_oracle as oracle


def label_prep(labels, ignore_index=-100):
    shifted = oracle.shift_labels(labels, ignore_index)
    counts = torch.tensor([int(labels.ne(ignore_index).sum()), int(shifted.ne(ignore_index).sum())])
    return shifted.reshape(-1).to(torch.int32), counts


def fused_linear_cross_entropy(hidden, weight, shifted_labels, num_items_in_batch=None, scale=1.0,
                               ignore_index=-100, flags=0, chunk_cols=0, skip_ignored_rows=False, return_stats=False):
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
        monkeypatch_or_none.setattr(ops, "fused_linear_cross_entropy", fused_linear_cross_entropy)
    else:
        ops.label_prep = label_prep
        ops.fused_linear_cross_entropy = fused_linear_cross_entropy
