"""End-to-end parity of the compute_loss seam on the GPU: the same tiny SmolLM3 (BASELINE config 1
architecture: 2 layers, hidden 256, vocab 128,256, tied head) is trained for a few optimizer steps
twice from the same seed -- once through the fused kernels (head swap + FusedCELossMixin, valid-row
compaction on), once through the stock transformers path (lm_head -> ForCausalLMLoss) -- with the
same gradient-accumulation windows, num_items_in_batch, clip and AdamW.  Losses must track within
BASELINE.json's 1e-3 relative tolerance and the trained weights must stay aligned."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rows(n=48):
    return [{"full-question": f"For topic {i % 5}, how do I handle situation number {i}?",
             "answer": "Stay calm and work through it step by step. " * (i % 6 + 1)} for i in range(n)]


def _run(device, fused: bool, steps: int = 3):
    from llm_fine_tune_distributed_b200.loss import install_fused_head
    from llm_fine_tune_distributed_b200.training import (ByteChatTokenizer, MiniSFTTrainer, SFTSettings, ShardedBatches,
                                                         apply_freeze_policy, format_prompt, make_model)

    torch.manual_seed(0)
    model = make_model("tiny", device, attn="sdpa")
    apply_freeze_policy(model)
    tok = ByteChatTokenizer(model.config.vocab_size)
    tokens = [tok.apply_chat_template(format_prompt(r)["messages"]) for r in _rows()]
    args = SFTSettings(per_device_train_batch_size=4, gradient_accumulation_steps=2, max_steps=steps, max_seq_length=256,
                       gradient_checkpointing=False, learning_rate=2e-4, logging_steps=1)
    batches = ShardedBatches(tokens, 4, 0, 1, tok.pad_token_id, args.max_seq_length, seed=42)
    logs = []

    class Log:
        def on_log(self, l):
            logs.append(dict(l))

    loss_impl = None
    if fused:
        install_fused_head(model)
    else:
        def loss_impl(m, inputs, n):   # the reference path: model computes lm_head + ForCausalLMLoss itself
            return m(**inputs, num_items_in_batch=n).loss

    trainer = MiniSFTTrainer(model, args, batches, callbacks=[Log()], device=device, loss_impl=loss_impl)
    trainer.train()
    return [l["loss"] for l in logs], [l["grad_norm"] for l in logs], model.lm_head.weight.detach().float().clone()


def test_fused_seam_tracks_stock_path_over_optimizer_steps(cuda_device):
    from llm_fine_tune_distributed_b200 import ops

    ops.STATS.reset()
    loss_f, gn_f, w_f = _run(cuda_device, fused=True)
    launched = ops.STATS.kernel_launches
    loss_s, gn_s, w_s = _run(cuda_device, fused=False)
    assert launched > 0, "the fused path did not launch this library's kernels"
    assert len(loss_f) == len(loss_s) == 3
    for a, b in zip(loss_f, loss_s):
        assert abs(a - b) <= 1e-3 * abs(b), (loss_f, loss_s)
    for a, b in zip(gn_f, gn_s):
        assert abs(a - b) <= 2e-2 * abs(b), (gn_f, gn_s)
    cos = float((w_f.flatten().double() @ w_s.flatten().double()) / (w_f.norm().double() * w_s.norm().double()))
    assert cos >= 0.9999
