"""Reference checkpoints <-> B200 heads.

The reference saves ``{"cfg", "state_dicts": {"model", "ema_model", "optimizer", ...}, "pickles"}`` with
``torch.save`` (workspace/base_workspace.py:33-77) and loads it back key by key, stripping the ``module.``
prefix that DDP / Accelerate add and falling back to ``ema_model`` when ``model`` is absent
(base_workspace.py:83-118).  The policy's state_dict holds the heads under ``model.diffloss.*``,
``model.diffloss_wrist.*``, ``model.diffactloss.*`` and ``model.diffproploss.*``
(policy/unified_video_action_policy.py:81, MAR:269-344).

The B200 heads keep the reference's parameter names, shapes and dtype (fp32 ``nn.Parameter``s), so the
round trip is lossless by construction: the packed bf16 operands of the CUDA library are derived data,
rebuilt by ``uva_head_load_weights`` when a parameter's version changes, and never stored.
"""
from __future__ import annotations

from typing import Dict, Mapping, Optional, Union

import torch
import torch.nn as nn

HEAD_ATTRS = ("diffloss", "diffloss_wrist", "diffactloss", "diffproploss")


def _payload(src: Union[str, Mapping]) -> Mapping:
    if isinstance(src, Mapping):
        return src
    try:  # the reference pickles its hydra cfg with dill (base_workspace.py:8, 72)
        import dill

        return torch.load(open(src, "rb"), pickle_module=dill, map_location="cpu", weights_only=False)
    except ImportError:
        return torch.load(open(src, "rb"), map_location="cpu", weights_only=False)


def _policy_state(payload: Mapping, which: Optional[str]) -> Mapping[str, torch.Tensor]:
    sd = payload["state_dicts"] if "state_dicts" in payload else payload
    if which is None:
        which = "model" if "model" in sd else "ema_model"  # base_workspace.py:107-116
    if which in sd:
        sd = sd[which]
    # else: `payload` already is a policy / MAR state_dict
    return {k.replace("module.", ""): v for k, v in sd.items()}  # base_workspace.py:96-101


def head_state_dicts(src: Union[str, Mapping], which: Optional[str] = None) -> Dict[str, Dict[str, torch.Tensor]]:
    """``{"diffloss": {...}, "diffactloss": {...}, ...}`` — the head entries of a reference checkpoint
    (path, payload dict, or a bare policy / MAR state_dict), keys relative to the head module."""
    flat = _policy_state(_payload(src), which)
    out: Dict[str, Dict[str, torch.Tensor]] = {}
    for k, v in flat.items():
        parts = k.split(".")
        for i, p in enumerate(parts[:-1]):
            if p in HEAD_ATTRS and (i == 0 or parts[i - 1] == "model"):
                out.setdefault(p, {})[".".join(parts[i + 1:])] = v
                break
    return out


def _mar_of(target: nn.Module) -> nn.Module:
    if any(hasattr(target, a) for a in HEAD_ATTRS):
        return target
    inner = getattr(target, "model", None)  # UnifiedVideoActionPolicy.model is the MAR
    if inner is not None and any(hasattr(inner, a) for a in HEAD_ATTRS):
        return inner
    raise ValueError("target has none of the head attributes " + ", ".join(HEAD_ATTRS))


def load_reference_checkpoint(target: nn.Module, src: Union[str, Mapping], which: Optional[str] = None,
                              strict: bool = True) -> Dict[str, int]:
    """Load the head weights of a reference checkpoint into the heads of ``target`` (a MAR, or a policy
    whose ``.model`` is one; the heads may be the B200 modules or the reference's own).  Returns the
    number of tensors loaded per head.  ``strict``: every head present on ``target`` must be in the
    checkpoint with exactly its keys."""
    mar = _mar_of(target)
    found = head_state_dicts(src, which)
    n = {}
    for attr in HEAD_ATTRS:
        head = getattr(mar, attr, None)
        if head is None:
            continue
        if attr not in found:
            if strict:
                raise KeyError(f"checkpoint has no '{attr}' entries")
            continue
        head.load_state_dict(found[attr], strict=strict)
        n[attr] = len(found[attr])
    return n


def export_head_state_dict(target: nn.Module, prefix: str = "model.") -> Dict[str, torch.Tensor]:
    """The heads of ``target`` under the reference's key names (``model.diffloss.net...``), ready to be
    merged into ``payload["state_dicts"]["model"]`` of a reference checkpoint."""
    mar = _mar_of(target)
    out = {}
    for attr in HEAD_ATTRS:
        head = getattr(mar, attr, None)
        if head is None:
            continue
        for k, v in head.state_dict().items():
            out[f"{prefix}{attr}.{k}"] = v.detach().cpu()
    return out
