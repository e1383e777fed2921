"""The driver with the reference's surface (`python -m llm_fine_tune_distributed_b200.training`,
configured only through the reference's environment variables) runs BASELINE config 1 -- tiny
random-init SmolLM3 (2 layers, hidden 256, vocab 128,256) on a QA parquet file -- on one GPU through
the fused kernels, and writes the artefacts training.py:309-335 writes."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _write_parquet(path, n=96):
    import pandas as pd

    pd.DataFrame({
        "full-question": [f"For topic {i % 7}, what should I do in situation {i}?" for i in range(n)],
        "answer": [("Check your surroundings, stay calm, and act deliberately. " * (i % 5 + 1)).strip() for i in range(n)],
    }).to_parquet(path, compression="snappy")


def _env(tmp_path, **extra):
    env = dict(os.environ)
    env.update(RANDOM_INIT="1", MODEL_CONFIG="tiny", ATTN_IMPL="sdpa", MAX_STEPS="3", BATCH_SIZE="4", GRAD_ACCUM="2",
               EVAL_STEPS="2",
               EPOCHS="1", LEARNING_RATE="1e-4", DATASET_PATH=str(tmp_path / "qa.parquet"),
               OUTPUT_DIR=str(tmp_path / "out"), PYTHONPATH=ROOT + os.pathsep + env.get("PYTHONPATH", ""))
    env.update(extra)
    return env


def _check(tmp_path, world):
    out = tmp_path / "out"
    summary = json.load(open(out / "training_summary.json"))
    history = json.load(open(out / "training_history.json"))
    assert summary["world_size"] == world and summary["distributed_training"] == (world > 1)
    assert summary["training_samples"] + summary["validation_samples"] == 96
    assert summary["trainable_params"] <= summary["total_params"]
    # logging_first_step=True, logging_steps=2 (training.py:266-267): 3 optimizer steps log at steps 1 and 2;
    # EVAL_STEPS=2 adds one evaluation entry after step 2 (eval_loss + eval_perplexity, training.py:224-234)
    train_logs = [h for h in history if "loss" in h]
    eval_logs = [h for h in history if "eval_loss" in h]
    assert len(train_logs) == 2 and all(8.0 < h["loss"] < 16.0 and "perplexity" in h for h in train_logs)
    assert len(eval_logs) == 1 and 8.0 < eval_logs[0]["eval_loss"] < 16.0 and "eval_perplexity" in eval_logs[0]
    assert os.path.isdir(out / "best_model") and os.listdir(out / "best_model")


def test_driver_single_gpu(cuda_device, tmp_path):
    _write_parquet(tmp_path / "qa.parquet")
    res = subprocess.run([sys.executable, "-m", "llm_fine_tune_distributed_b200.training"], cwd=ROOT,
                         env=_env(tmp_path, WORLD_SIZE="1", RANK="0", LOCAL_RANK="0"), capture_output=True, text=True,
                         timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    _check(tmp_path, 1)


def test_driver_padding_free_completion_only(cuda_device, tmp_path):
    """PACKING=1: padding-free [1, total] batches with restarting position_ids and FlashAttention varlen
    arguments; COMPLETION_ONLY=1: only the assistant's tokens carry labels, so most rows are compacted away."""
    _write_parquet(tmp_path / "qa.parquet")
    res = subprocess.run([sys.executable, "-m", "llm_fine_tune_distributed_b200.training"], cwd=ROOT,
                         env=_env(tmp_path, WORLD_SIZE="1", RANK="0", LOCAL_RANK="0", PACKING="1", COMPLETION_ONLY="1",
                                  ATTN_IMPL="flash_attention_2"), capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    _check(tmp_path, 1)


@pytest.mark.parametrize("comm", ["nccl", "nvlink", "nvlink+tied"])
def test_driver_two_gpus_ddp(cuda_device, tmp_path, comm):
    """nvlink+tied: TIED_EXCHANGE=1 also splits the tied lm_head / embed_tokens gradient (tied_embedding.py)."""
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    _write_parquet(tmp_path / "qa.parquet")
    env = _env(tmp_path, B200SFT_COMM=comm.split("+")[0], TIED_EXCHANGE="1" if comm.endswith("tied") else "0")
    for k in ("WORLD_SIZE", "RANK", "LOCAL_RANK"):
        env.pop(k, None)
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29761", "-m",
                          "llm_fine_tune_distributed_b200.training"], cwd=ROOT, env=env, capture_output=True, text=True,
                         timeout=400)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    _check(tmp_path, 2)
    summary = json.load(open(tmp_path / "out" / "training_summary.json"))
    assert summary.get("tied_exchange_steps") == (3 if comm.endswith("tied") else None)   # MAX_STEPS=3 optimizer steps


@pytest.mark.parametrize("extra", [dict(COMPLETION_ONLY="1"),
                                   dict(PACKING="1", COMPLETION_ONLY="1", ATTN_IMPL="flash_attention_2")],
                         ids=["padded", "packed"])
def test_driver_device_labels_same_run(cuda_device, tmp_path, extra):
    """DEVICE_LABELS=1: the collator ships raw masks and b200sft_label_frontend builds labels, counts and the
    compaction list on the GPU.  Same data, same seed -> the same losses as with host-made labels."""
    _write_parquet(tmp_path / "qa.parquet")
    histories = []
    for flag in ("0", "1"):
        res = subprocess.run([sys.executable, "-m", "llm_fine_tune_distributed_b200.training"], cwd=ROOT,
                             env=_env(tmp_path, WORLD_SIZE="1", RANK="0", LOCAL_RANK="0", DEVICE_LABELS=flag, **extra),
                             capture_output=True, text=True, timeout=300)
        assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
        _check(tmp_path, 1)
        histories.append(json.load(open(tmp_path / "out" / "training_history.json")))
    host, dev = ([h["loss"] for h in hist if "loss" in h] for hist in histories)
    assert host[0] == dev[0]  # first window: identical labels through identical kernels
    assert dev == pytest.approx(host, rel=1e-3)
    ev = [[h["eval_loss"] for h in hist if "eval_loss" in h] for hist in histories]
    assert ev[1] == pytest.approx(ev[0], rel=1e-3)
