"""Generates tests/golden/*.npz by running the REAL third-party code the reference executes
(transformers 5.5.0 / torch 2.11.0 installed in the build image) on small seeded inputs.

The reference ships no tests or golden vectors, so these fixtures are what pins the oracle
(oracle/sft_oracle.py).  Run in the build container:  python tests/golden/make_golden.py
bf16 tensors are stored as float32 (exactly representable).
"""
import os
import sys
from types import SimpleNamespace

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))


def f32(t):
    return t.detach().float().cpu().numpy()


def make_causal_lm_loss():
    """transformers.loss.loss_utils.ForCausalLMLoss on logits = F.linear(hidden, weight), bf16."""
    from transformers.loss.loss_utils import ForCausalLMLoss

    g = torch.Generator().manual_seed(7)
    B, T, Hd, V = 3, 12, 32, 96
    hidden = torch.randn(B, T, Hd, generator=g).to(torch.bfloat16)
    weight = (torch.randn(V, Hd, generator=g) * 0.3).to(torch.bfloat16)
    labels = torch.randint(0, V, (B, T), generator=g)
    labels[0, :5] = -100
    labels[1, 8:] = -100
    labels[2, ::3] = -100
    out = {"hidden": f32(hidden), "weight": f32(weight), "labels": labels.numpy()}
    for tag, n in (("sum", 29), ("mean", None)):
        h = hidden.clone().requires_grad_(True)
        w = weight.clone().requires_grad_(True)
        logits = F.linear(h, w)
        loss = ForCausalLMLoss(logits=logits, labels=labels, vocab_size=V, num_items_in_batch=n)
        loss.backward()
        out[f"loss_{tag}"] = f32(loss)
        out[f"dhidden_{tag}"] = f32(h.grad)
        out[f"dweight_{tag}"] = f32(w.grad)
    out["num_items_sum"] = np.int64(29)
    np.savez_compressed(os.path.join(HERE, "causal_lm_loss.npz"), **out)


def tiny_model():
    from transformers import SmolLM3Config, SmolLM3ForCausalLM

    cfg = SmolLM3Config(num_hidden_layers=2, hidden_size=64, intermediate_size=128, num_attention_heads=4,
                        num_key_value_heads=1, vocab_size=512, max_position_embeddings=128, tie_word_embeddings=True,
                        pad_token_id=0, bos_token_id=1, eos_token_id=2)
    cfg._attn_implementation = "eager"
    torch.manual_seed(0)
    return SmolLM3ForCausalLM(cfg).to(torch.bfloat16)


def make_smollm3_and_trainer():
    """SmolLM3ForCausalLM.forward with labels, Trainer._get_num_items_in_batch and
    Trainer.compute_loss (both called unbound with a stub `self`; accelerate is not installed)."""
    from transformers.trainer import Trainer

    model = tiny_model()
    model.eval()
    g = torch.Generator().manual_seed(3)
    B, T, V = 2, 16, 512
    window = []
    for _ in range(2):  # a gradient-accumulation window of two micro-batches
        ids = torch.randint(0, V, (B, T), generator=g)
        labels = ids.clone()
        labels[:, : T // 4] = -100
        labels[1, -3:] = -100
        window.append({"input_ids": ids, "labels": labels})

    stub = SimpleNamespace(
        args=SimpleNamespace(average_tokens_across_devices=True, n_gpu=1, world_size=1),
        accelerator=SimpleNamespace(num_processes=2, gather=lambda t: t, parallelism_config=None,
                                    unwrap_model=lambda m: m),
        label_smoother=None, compute_loss_func=None, model_accepts_loss_kwargs=True, model=model)
    n_items = Trainer._get_num_items_in_batch(stub, window, torch.device("cpu"))

    captured = {}
    hook = model.model.norm.register_forward_hook(lambda m, i, o: captured.__setitem__("h", o.detach().clone()))
    out = {"num_items": np.int64(int(n_items)), "weight": f32(model.lm_head.weight)}
    for i, mb in enumerate(window):
        res = model(input_ids=mb["input_ids"], labels=mb["labels"], num_items_in_batch=n_items)
        loss_cl = Trainer.compute_loss(stub, model, dict(mb), num_items_in_batch=n_items)
        out[f"input_ids_{i}"] = mb["input_ids"].numpy()
        out[f"labels_{i}"] = mb["labels"].numpy()
        out[f"hidden_{i}"] = f32(captured["h"])
        out[f"model_loss_{i}"] = f32(res.loss)
        out[f"compute_loss_{i}"] = f32(loss_cl)      # = model loss * num_processes (2)
    res = model(input_ids=window[0]["input_ids"], labels=window[0]["labels"])  # num_items None -> mean
    out["model_loss_mean_0"] = f32(res.loss)
    hook.remove()
    np.savez_compressed(os.path.join(HERE, "smollm3_tiny_step.npz"), **out)


def _hook_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from torch.distributed.algorithms.ddp_comm_hooks import default_hooks

    g = torch.Generator().manual_seed(100 + rank)
    bucket = torch.randn(1000, generator=g).to(torch.bfloat16)
    before = bucket.clone()

    class FakeBucket:  # GradBucket cannot be constructed from Python; the hook only calls buffer()
        def buffer(self):
            return bucket

    fut = default_hooks.allreduce_hook(None, FakeBucket())
    res = fut.wait()
    ret[rank] = (f32(before), f32(res))
    dist.destroy_process_group()


def make_allreduce_hook():
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_hook_worker, args=(world, 29611, ret), nprocs=world, join=True)
    np.savez_compressed(os.path.join(HERE, "allreduce_hook.npz"), bucket_rank0=ret[0][0], bucket_rank1=ret[1][0],
                        result_rank0=ret[0][1], result_rank1=ret[1][1], world=np.int64(world))


if __name__ == "__main__":
    make_causal_lm_loss()
    make_smollm3_and_trainer()
    make_allreduce_hook()
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")
