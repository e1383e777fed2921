"""The C-ABI library builds, loads, and exports every symbol include/b200sft.h declares (no GPU
needed: nothing is launched).  Also checks the host-only size queries and argument validation."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200sft.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200sft_\w+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    from llm_fine_tune_distributed_b200 import _cabi, build

    build.build()
    return _cabi.load()


def test_header_and_binding_agree(lib):
    from llm_fine_tune_distributed_b200 import _cabi

    decl = _declared()
    assert len(decl) >= 12
    assert sorted(_cabi.SIGNATURES) == decl
    raw = ctypes.CDLL(str(_cabi.LIB_PATH))
    for name in decl:
        assert getattr(raw, name) is not None, f"{name} not exported"


def test_version_and_workspace_queries(lib):
    assert lib.b200sft_version() >= 100
    # forward: 501 vocabulary tiles x 8192 rows x 8 B of partials + two fp32 row vectors
    need = lib.b200sft_flce_fwd_workspace_bytes(8192, 2048, 128256)
    assert need == 501 * 8192 * 8 + 2 * 8192 * 4
    assert lib.b200sft_flce_fwd_workspace_bytes(0, 2048, 128256) == 0
    # backward: one [N, slab] bf16 dlogits slab + fp32 dX accumulators: one per K-split of the dX GEMM (4 when
    # that GEMM has fewer than 16 tiles per worker, else 1 -- and none at all for a single unsplit slab)
    one = lib.b200sft_flce_bwd_workspace_bytes(4096, 2048, 128256, 0)       # 1 GiB budget: single slab
    two = lib.b200sft_flce_bwd_workspace_bytes(8192, 2048, 128256, 0)       # two slabs of 64256 columns
    big = lib.b200sft_flce_bwd_workspace_bytes(65536, 2048, 128256, 0)      # 16 slabs of 8192, dX not split
    # ... + one fp32 128x256 tile per SM for the K pieces of the dW GEMM's tail tiles (148 SMs assumed without a GPU)
    tail = 148 * 128 * 256 * 4
    assert one == 4096 * 128256 * 2 + 4 * 4096 * 2048 * 4 + tail
    assert two == 8192 * 64256 * 2 + 4 * 8192 * 2048 * 4 + tail
    assert big == 65536 * 8192 * 2 + 65536 * 2048 * 4 + tail
    assert lib.b200sft_flce_bwd_workspace_bytes(8192, 2048, 128256, 128256) == 8192 * 128256 * 2 + 4 * 8192 * 2048 * 4 + tail


def test_argument_validation_without_launching(lib):
    # bad dims / null pointers are rejected before any CUDA call
    rc = lib.b200sft_flce_fwd(None, None, None, 128, 2048, 128256, -100, 1.0, 0, None, 0, None, None, None, None, None, 0, None)
    assert rc < 0 and b"null" in lib.b200sft_last_error()
    rc = lib.b200sft_flce_fwd(None, None, None, 128, 2047, 128256, -100, 1.0, 0, None, 0, None, None, None, None, None, 0, None)
    assert rc < 0 and b"multiples of 8" in lib.b200sft_last_error()
    rc = lib.b200sft_gemm_bf16(None, None, None, 1, 1, 1, 8, 8, 8, 0, 0, None)
    assert rc < 0


def test_python_ops_refuse_cpu_tensors():
    import torch

    from llm_fine_tune_distributed_b200 import ops
    from llm_fine_tune_distributed_b200._cabi import B200SftError

    with pytest.raises(B200SftError):
        ops.label_prep(torch.zeros(2, 4, dtype=torch.int64))
    with pytest.raises(B200SftError):
        ops.scale_(torch.zeros(8, dtype=torch.bfloat16), 0.5)
    with pytest.raises(B200SftError):
        ops.fused_linear_cross_entropy(torch.zeros(8, 8, dtype=torch.bfloat16), torch.zeros(8, 8, dtype=torch.bfloat16),
                                       torch.zeros(8, dtype=torch.int32))


def test_product_code_never_imports_the_oracle():
    """oracle/ is test infrastructure: only tests/, __graft_entry__.smoke() and bench.py's CPU legs may use
    it.  The package, the benchmarks and the examples must not."""
    import pathlib
    import re

    root = pathlib.Path(__file__).resolve().parent.parent
    pat = re.compile(r"^\s*(from\s+oracle\b|import\s+oracle\b)", re.M)
    offenders = [str(p.relative_to(root))
                 for d in ("llm_fine_tune_distributed_b200", "benchmarks", "examples")
                 for p in (root / d).rglob("*.py") if pat.search(p.read_text())]
    assert not offenders, offenders
    for src in (root / "llm_fine_tune_distributed_b200" / "csrc").glob("*.cu*"):
        assert "oracle" not in src.read_text(), src.name
