"""b200sft_flce_fwd_noalloc -- the forward entry point for callers that cannot allocate.

Who needs it: anything that binds the C ABI but has no way to obtain device scratch -- a thin
ctypes/cgo-style stub that only ever sees tensors owned by another framework, a callback that runs
when PyTorch's allocator has already claimed almost all of HBM (a raw cudaMalloc would fail), or
code on the step's critical path where cudaMalloc/cudaFree would synchronise the device.  The
variant keeps its scratch in a fixed 40 MiB array inside the library, so these tests drive it
through ctypes exactly as such a caller would and compare it -- bit for bit -- with the regular entry
point (same kernels, caller-provided scratch), and with the oracle, on two calls in a row: the second call reuses the array with the
first call's leftovers in it, with a different shape and different labels.

This file sorts last on purpose, so a run that stops at the first failure has already run
everything else.  It must also pass alone in a fresh process (the first call on a device does the
one-time address lookup).
"""
import pytest
import torch

from oracle import sft_oracle as oracle

pytestmark = pytest.mark.gpu

ERR_WORKSPACE = -3


def _inputs(N, H, V, masked, seed, device):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(N, H, generator=g).to(torch.bfloat16).to(device)
    w = (torch.randn(V, H, generator=g) * 0.05).to(torch.bfloat16).to(device)
    labels = torch.randint(0, V, (N,), generator=g, dtype=torch.int32)
    labels[torch.rand(N, generator=g) < masked] = -100
    return x, w, labels.to(device)


def _run(lib, noalloc, x, w, labels, denom):
    """One forward call through the C ABI; returns (loss, lse, n_valid)."""
    N, H = x.shape
    V = w.shape[0]
    dev = x.device
    loss = torch.full((), float("nan"), dtype=torch.float32, device=dev)
    lse = torch.full((N,), float("nan"), dtype=torch.float32, device=dev)
    n_valid = torch.full((), -1, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    head = (x.data_ptr(), w.data_ptr(), labels.data_ptr(), N, H, V, -100, 1.0, 1, denom.data_ptr(), 0)
    outs = (loss.data_ptr(), lse.data_ptr(), n_valid.data_ptr())
    if noalloc:
        rc = lib.b200sft_flce_fwd_noalloc(*head, *outs, stream)
    else:
        ws_bytes = lib.b200sft_flce_fwd_workspace_bytes(N, H, V)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        rc = lib.b200sft_flce_fwd(*head, None, *outs, ws.data_ptr(), ws_bytes, stream)   # None: all N rows live
    assert rc == 0, lib.b200sft_last_error().decode()
    torch.cuda.synchronize(dev)
    return loss, lse, n_valid


def test_noalloc_forward_matches_regular_entry_point_twice(cuda_device):
    from llm_fine_tune_distributed_b200 import _cabi

    lib = _cabi.load()
    with torch.cuda.device(cuda_device):
        # two calls in a row on the shared scratch: a larger problem first, then a smaller one with
        # other labels, so stale partials / target logits / row losses would show up if not reset
        for (N, H, V, masked, seed) in [(1024, 256, 8192, 0.3, 1), (384, 128, 1000, 0.7, 2)]:
            x, w, labels = _inputs(N, H, V, masked, seed, cuda_device)
            denom = torch.tensor(max(int((labels != -100).sum()), 1), dtype=torch.int64, device=cuda_device)
            ref_loss, ref_lse, ref_nv = _run(lib, False, x, w, labels, denom)
            loss, lse, nv = _run(lib, True, x, w, labels, denom)
            # Both entry points run the same deterministic kernels and differ only in where the scratch lives,
            # so the outputs must agree exactly -- anything less would mean leftovers in the shared array leaked in.
            assert torch.equal(loss, ref_loss) and torch.equal(lse, ref_lse) and torch.equal(nv, ref_nv)
            assert int(nv) == int((labels != -100).sum())
            # ... and against the oracle (the reference's lm_head + cross-entropy), with the tolerances of
            # tests/test_gpu_flce.py: loss 1e-3 relative, lse 1e-3 absolute
            want = oracle.lm_head_ce(x, w, labels.long(), float(denom))
            assert abs(loss.item() - want.item()) <= 1e-3 * abs(want.item())
            want_lse = torch.logsumexp(torch.nn.functional.linear(x, w).float(), dim=-1)
            assert torch.allclose(lse, want_lse, atol=1e-3, rtol=0)


def test_noalloc_on_a_second_device():
    """The scratch array exists once per GPU and its address is looked up per device: the first call on device 1
    must resolve device 1's copy, not reuse device 0's pointer."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from llm_fine_tune_distributed_b200 import _cabi

    lib = _cabi.load()
    for dev in (torch.device("cuda:0"), torch.device("cuda:1")):
        with torch.cuda.device(dev):
            x, w, labels = _inputs(512, 128, 4096, 0.4, 3, dev)
            denom = torch.tensor(max(int((labels != -100).sum()), 1), dtype=torch.int64, device=dev)
            ref = _run(lib, False, x, w, labels, denom)
            got = _run(lib, True, x, w, labels, denom)
            assert all(torch.equal(a, b) for a, b in zip(got, ref))


def test_noalloc_rejects_oversized_request(cuda_device):
    from llm_fine_tune_distributed_b200 import _cabi

    lib = _cabi.load()
    N, H, V = 65536, 2048, 128256   # needs ~251 MiB of scratch: beyond the built-in 40 MiB
    assert lib.b200sft_flce_fwd_workspace_bytes(N, H, V) > _cabi.NOALLOC_SCRATCH_BYTES
    with torch.cuda.device(cuda_device):
        one = torch.zeros(8, dtype=torch.int64, device=cuda_device)  # never dereferenced: the size check comes first
        rc = lib.b200sft_flce_fwd_noalloc(one.data_ptr(), one.data_ptr(), one.data_ptr(), N, H, V, -100, 1.0, 0, None,
                                          0, one.data_ptr(), one.data_ptr(), None,
                                          torch.cuda.current_stream(cuda_device).cuda_stream)
    assert rc == ERR_WORKSPACE
    msg = lib.b200sft_last_error().decode()
    assert "scratch" in msg and str(_cabi.NOALLOC_SCRATCH_BYTES) in msg
