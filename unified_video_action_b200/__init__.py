"""unified_video_action_b200 — B200-native (sm_100a) per-token diffusion head of UVA.

Only the hot path of ``BASELINE.json:north_star`` lives here: ``DiffLoss`` / ``DiffActLoss``
(SimpleMLPAdaLN denoiser, training loss fwd/bwd, 100-step DDPM sampling with CFG) as hand-written
CUDA behind the C ABI of ``include/uva_head.h``.  Everything else of UVA stays the reference's
PyTorch; ``patch(mar_model)`` swaps the heads of a reference ``MAR`` in place.
"""
from .diffusion_loss import DiffLoss, SimpleMLPAdaLN, param_names  # noqa: F401
from .diffusion_action_loss import DiffActLoss  # noqa: F401
from .patch import patch, shard_rows, allreduce_head_grads, overlap_allreduce, sample_masked_tokens  # noqa: F401
from .optim import HeadAdamW, ema_decay_at  # noqa: F401
from .checkpoint import export_head_state_dict, head_state_dicts, load_reference_checkpoint  # noqa: F401

__all__ = ["DiffLoss", "DiffActLoss", "SimpleMLPAdaLN", "param_names", "patch", "shard_rows", "allreduce_head_grads", "overlap_allreduce", "sample_masked_tokens", "HeadAdamW", "ema_decay_at",
           "load_reference_checkpoint", "head_state_dicts", "export_head_state_dict"]
