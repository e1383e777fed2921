"""ctypes binding of libb200sft.so -- the C-ABI boundary declared in include/b200sft.h.

This is the only place the shared library is loaded.  There is no CPU or PyTorch fallback: if the
library is missing or a call fails, the caller gets an exception.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_size_t, c_uint32, c_void_p
from pathlib import Path

LIB_PATH = Path(__file__).resolve().parent / "libb200sft.so"

FLAG_ROUND_LOGITS_BF16 = 1
DENOM_NONE, DENOM_DEVICE, DENOM_VALID = 0, 1, 2
NOALLOC_SCRATCH_BYTES = 40 << 20  # B200SFT_NOALLOC_SCRATCH_BYTES

# name -> (restype, argtypes); mirrors include/b200sft.h one to one (checked by tests/test_cabi_symbols.py)
SIGNATURES = {
    "b200sft_last_error": (c_char_p, []),
    "b200sft_version": (c_int, []),
    "b200sft_set_cta_group": (c_int, [c_int]),
    "b200sft_label_prep": (c_int, [c_void_p, c_int, c_int, c_int64, c_void_p, c_void_p, c_void_p]),
    "b200sft_flce_fwd_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "b200sft_flce_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_int,
                                 c_void_p, c_uint32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t,
                                 c_void_p]),
    "b200sft_flce_fwd_stats_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "b200sft_flce_fwd_stats": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_int,
                                       c_void_p, c_uint32, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                       c_void_p, c_void_p, c_size_t, c_void_p]),
    "b200sft_compact_rows": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                     c_void_p, c_void_p, c_void_p]),
    "b200sft_label_frontend": (c_int, [c_void_p] * 5 + [c_int, c_int, c_int64] + [c_void_p] * 8),
    "b200sft_compact_labels": (c_int, [c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "b200sft_gather_rows": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "b200sft_rmsnorm_gather": (c_int, [c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p,
                                       c_void_p]),
    "b200sft_rmsnorm_bwd_workspace_bytes": (c_size_t, [c_int]),
    "b200sft_rmsnorm_scatter_bwd": (c_int, [c_void_p] * 7 + [c_int, c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "b200sft_scatter_rows": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "b200sft_flce_fwd_noalloc": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_int,
                                         c_void_p, c_uint32, c_void_p, c_void_p, c_void_p, c_void_p]),
    "b200sft_flce_bwd_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "b200sft_flce_bwd_launch_count": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "b200sft_flce_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int,
                                 c_float, c_int, c_void_p, c_uint32, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                 c_size_t, c_void_p]),
    "b200sft_gemm_bf16": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                  c_int, c_void_p]),
    "b200sft_scale_bf16_inplace": (c_int, [c_void_p, c_size_t, c_float, c_void_p]),
    "b200sft_scale_f32_inplace": (c_int, [c_void_p, c_size_t, c_float, c_void_p]),
    "b200sft_scale_cast_f32_to_bf16": (c_int, [c_void_p, c_void_p, c_size_t, c_float, c_void_p]),
    "b200sft_cast_bf16_to_f32": (c_int, [c_void_p, c_void_p, c_size_t, c_float, c_void_p]),
    "b200sft_grad_norm_workspace_bytes": (c_size_t, [c_void_p, c_int]),
    "b200sft_grad_norm_bf16": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_size_t, c_void_p]),
    "b200sft_adamw_clip_bf16": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_float, c_float,
                                        c_float, c_float, c_float, c_int, c_float, c_void_p, c_void_p]),
    "b200sft_allreduce_scale_bf16_multimem": (c_int, [c_void_p, c_size_t, c_int, c_int, c_float, c_void_p]),
    "b200sft_allreduce_scale_bf16_p2p": (c_int, [c_void_p, c_size_t, c_int, c_int, c_float, c_void_p]),
    "b200sft_embed_grad_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "b200sft_embed_grad_rows": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_float, c_void_p,
                                        c_void_p, c_size_t, c_void_p]),
}

_lib = None


class B200SftError(RuntimeError):
    """A C-ABI call returned a negative status."""


def load() -> ctypes.CDLL:
    """Load libb200sft.so (built by build.py / __graft_entry__.build()); raises if it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise B200SftError(
            f"{LIB_PATH} is missing: run `python -m llm_fine_tune_distributed_b200.build` "
            "(or __graft_entry__.build()). There is no fallback path."
        )
    lib = ctypes.CDLL(str(LIB_PATH))
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    grp = os.environ.get("B200SFT_CTA_GROUP")
    if grp:
        check(lib.b200sft_set_cta_group(int(grp)), "b200sft_set_cta_group")
    return lib


def set_cta_group(cta_group: int) -> None:
    """1 = one CTA per 128x256 tile, 2 = two-CTA clusters on 256x256 tiles (process-wide)."""
    check(load().b200sft_set_cta_group(int(cta_group)), "b200sft_set_cta_group")


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().b200sft_last_error()
        raise B200SftError(f"{what} failed ({rc}): {msg.decode() if msg else 'unknown error'}")
