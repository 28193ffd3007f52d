"""Glue for the callers of the head: in-place swap inside a reference ``MAR`` (MAR:269-344), row
sharding for multi-GPU sampling and the head-gradient all-reduce for data-parallel training
(SURVEY.md §8e)."""
from __future__ import annotations

from typing import Iterable, Optional, Tuple

import torch
import torch.nn as nn

from .diffusion_action_loss import DiffActLoss
from .diffusion_loss import DiffLoss

_HEAD_ATTRS = ("diffloss", "diffloss_wrist", "diffactloss", "diffproploss")


def _geometry(ref_head) -> dict:
    net = ref_head.net
    return dict(
        target_channels=ref_head.in_channels,
        z_channels=net.cond_embed.in_features,
        depth=len(net.res_blocks),
        width=net.cond_embed.out_features,
    )


def patch(mar_model: nn.Module, precision: Optional[str] = None) -> nn.Module:
    """Replace ``diffloss`` / ``diffloss_wrist`` / ``diffactloss`` / ``diffproploss`` of a reference
    MAR by the B200 heads, copying the weights (state_dict keys are identical)."""
    for attr in _HEAD_ATTRS:
        ref = getattr(mar_model, attr, None)
        if ref is None or isinstance(ref, (DiffLoss, DiffActLoss)):
            continue
        g = _geometry(ref)
        kw = dict(n_frames=getattr(ref, "n_frames", 4), language_emb_model=getattr(ref, "language_emb_model", None),
                  language_emb_model_type=getattr(ref, "language_emb_model_type", None), precision=precision)
        steps = str(ref.gen_diffusion.num_timesteps)
        if hasattr(ref, "act_model_type"):
            new = DiffActLoss(num_sampling_steps=steps, act_diff_training_steps=ref.train_diffusion.num_timesteps,
                              act_diff_testing_steps=steps, act_model_type=ref.act_model_type, **g, **kw)
        else:
            new = DiffLoss(num_sampling_steps=steps, **g, **kw)
        new.load_state_dict(ref.state_dict(), strict=True)
        p = next(ref.parameters())
        new.to(device=p.device)
        setattr(mar_model, attr, new)
    return mar_model


def shard_rows(n_rows: int, rank: int, world_size: int, pair_halves: bool = False) -> Tuple[int, int]:
    """[begin, end) of the rows rank ``rank`` owns.  Rows of the head are independent (DL:47-48), so
    inference shards by row with no collective.  With ``pair_halves`` (CFG: rows are cond || uncond)
    the split is over the first half; the caller takes the same range from both halves."""
    span = n_rows // 2 if pair_halves else n_rows
    base, rem = divmod(span, world_size)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def allreduce_head_grads(heads: Iterable[nn.Module], group=None, average: bool = True) -> int:
    """Data-parallel training: one all-reduce per head over a flat fp32 gradient buffer (NCCL over
    NVLink on GPUs, gloo in the CPU tests).  Returns the number of collectives issued."""
    import torch.distributed as dist

    n = 0
    world = dist.get_world_size(group)
    for head in heads:
        grads = [p.grad for p in head.parameters() if p.grad is not None]
        if not grads:
            continue
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat, group=group)
        if average:
            flat /= world
        off = 0
        for g in grads:
            g.copy_(flat[off:off + g.numel()].view_as(g))
            off += g.numel()
        n += 1
    return n
