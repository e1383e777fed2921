"""bench.py's host-side helpers (no GPU): synthetic workloads have the shapes and masks the contract
names, token accounting matches attention_mask.sum(), and the nvidia-smi sampler parses its lines."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_full_workload_shapes_and_token_counts():
    w = bench.synthetic_window(step=3, ga=2, B=8, T=1024, pin=False, workload="full")
    assert len(w) == 2 and all(mb["input_ids"].shape == (8, 1024) and mb["input_ids"].dtype == torch.int64 for mb in w)
    assert all(torch.equal(mb["labels"], mb["input_ids"]) for mb in w)
    assert bench.window_tokens(w) == (16384, 16384)
    again = bench.synthetic_window(step=3, ga=2, B=8, T=1024, pin=False, workload="full")
    assert torch.equal(w[0]["input_ids"], again[0]["input_ids"])           # seeded
    assert int(w[0]["input_ids"].max()) < bench.V


def test_padded_workload_matches_reference_collator_semantics():
    mb = bench.synthetic_window(step=1, ga=1, B=8, T=1024, pin=False, workload="padded")[0]
    mask = mb["attention_mask"].bool()
    assert bool((mb["labels"][~mask] == -100).all()) and torch.equal(mb["labels"][mask], mb["input_ids"][mask])
    assert bool((mask.long().diff(dim=1) <= 0).all())                    # right padding only
    trained, valid = bench.window_tokens([mb])
    assert trained == int(mask.sum()) == valid


def test_packed_workload_is_flattened_with_boundaries():
    mb = bench.synthetic_window(step=1, ga=1, B=2, T=4096, pin=False, workload="packed")[0]
    assert mb["input_ids"].shape == (1, 8192) and mb["position_ids"].shape == (1, 8192)
    cu = mb["cu_seq_lens_q"].tolist()
    assert cu[0] == 0 and cu[-1] == 8192 and mb["cu_seq_lens_q"].dtype == torch.int32
    starts = torch.tensor(cu[:-1])
    assert bool((mb["position_ids"][0, starts] == 0).all()) and bool((mb["labels"][0, starts] == -100).all())
    assert int((mb["labels"] == -100).sum()) == len(cu) - 1             # only the sample boundaries are masked
    assert mb["max_length_q"] == max(b - a for a, b in zip(cu[:-1], cu[1:]))


def test_clock_sampler_parsing():
    s = bench.ClockSampler(0)
    s.proc = type("P", (), {"terminate": lambda self: None, "wait": lambda self, timeout=None: 0, "kill": lambda self: None})()
    s.lines = ["0, 1905, 1965, 955.1, Not Active, Not Active, Not Active, Active",
               "0, 1890, 1965, 970.2, Not Active, Not Active, Not Active, Active",
               "0, 1935, 1965, 940.0, Not Active, Not Active, Not Active, Not Active", "garbage"]
    out = s.stop()
    assert out["sm_mhz"] == 1905 and out["sm_max_mhz"] == 1965 and out["reasons"] == ["sw_power_cap"]
    assert out["samples"] == 3 and abs(out["power_w_max"] - 970.2) < 1e-6


def test_peaks_come_from_measured_file_when_present():
    p = bench.read_peaks()
    assert p["burst"] > 1000 and p["sustained"] <= p["burst"] and p["hbm"] > 1000


def test_completion_only_workload_masks_the_prompt():
    mb = bench.synthetic_window(step=2, ga=1, B=4, T=1000, pin=False, workload="completion")[0]
    assert "attention_mask" not in mb                                   # full-length rows: every token is trained on
    assert bool((mb["labels"][:, :600] == -100).all()) and torch.equal(mb["labels"][:, 600:], mb["input_ids"][:, 600:])
    assert bench.window_tokens([mb]) == (4000, 1600)


def test_reference_arm_thread_count_ignores_torchruns_omp_setting(monkeypatch):
    """torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU arm (rank 0 alone) must still use every core it may
    run on, and the other ranks must leave without work."""
    monkeypatch.setenv("OMP_NUM_THREADS", "1")
    assert bench.host_threads() == len(os.sched_getaffinity(0)) >= 1
    monkeypatch.setenv("RANK", "3")
    monkeypatch.setenv("WORLD_SIZE", "8")
    assert bench.run_reference(type("A", (), {"steps": 5, "warmup": 3, "cpu_tokens": 16})()) == {}
