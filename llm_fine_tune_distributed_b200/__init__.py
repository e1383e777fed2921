"""B200 (sm_100a) implementation of the data-parallel hot path of the SFT step of
thesteve0/llm-fine-tune-distributed: fused LM-head + masked cross-entropy (forward and backward)
and the DDP gradient-bucket hook, behind the reference's own seams
(``Trainer.compute_loss`` and ``DistributedDataParallel.register_comm_hook``).

Layout:  csrc/ (CUDA kernels + C ABI, built into libb200sft.so), _cabi.py (ctypes binding),
ops.py (torch custom ops), loss.py (compute_loss seam), comm_hook.py (DDP hook seam),
optim.py (grad-norm clip + AdamW step tail), training.py (driver with the reference's env/SFTConfig surface).
"""
from . import _cabi  # noqa: F401
from .ops import (  # noqa: F401
    FLAG_ROUND_LOGITS_BF16,
    IGNORE_INDEX,
    cast_bf16_to_f32,
    flce_backward,
    flce_forward,
    fused_linear_cross_entropy,
    gather_rows,
    gemm_bf16,
    label_frontend,
    label_prep,
    scale_,
    scale_cast_f32_to_bf16,
)
from .loss import FusedCELossMixin, FusedLMHead, fused_causal_lm_loss, fused_head_armed, install_fused_head  # noqa: F401
from .comm_hook import B200HookState, b200_allreduce_hook, b200_bf16_compress_hook, register_b200_comm_hook  # noqa: F401

from .optim import B200AdamW  # noqa: F401

__version__ = "0.1.0"
