"""Reference checkpoint format <-> B200 heads (base_workspace.py:33-118).  CPU only: the
heads' parameters are plain fp32 nn.Parameters under the reference's names, so this is host logic."""
import types

import pytest
import torch

from unified_video_action_b200 import DiffActLoss, DiffLoss
from unified_video_action_b200.checkpoint import export_head_state_dict, head_state_dicts, load_reference_checkpoint

KW = dict(n_frames=4, language_emb_model=None, language_emb_model_type=None)


def _mar(seed):
    torch.manual_seed(seed)
    m = torch.nn.Module()
    m.diffloss = DiffLoss(16, 64, 2, 256, "100", **KW)
    m.diffactloss = DiffActLoss(2, 64, 2, 256, "100", **KW)
    m.encoder = torch.nn.Linear(8, 8)  # stands for the rest of MAR
    with torch.no_grad():
        for p in m.parameters():
            p.copy_(torch.randn_like(p))
    return m


def test_reference_payload_round_trip(tmp_path):
    src, dst = _mar(0), _mar(1)
    policy_sd = {"module.model." + k: v.clone() for k, v in src.state_dict().items()}  # DDP-wrapped policy keys
    policy_sd["module.normalizer.params_dict.action.scale"] = torch.ones(2)
    payload = {"cfg": None, "state_dicts": {"ema_model": policy_sd, "optimizer": {"state": {}}}, "pickles": {}}
    path = tmp_path / "latest.ckpt"
    torch.save(payload, path.open("wb"))

    heads = head_state_dicts(str(path))
    assert set(heads) == {"diffloss", "diffactloss"}
    assert set(heads["diffloss"]) == set(src.diffloss.state_dict())
    policy = types.SimpleNamespace(model=dst)  # UnifiedVideoActionPolicy.model is the MAR
    n = load_reference_checkpoint(policy, str(path))
    assert n == {"diffloss": len(src.diffloss.state_dict()), "diffactloss": len(src.diffactloss.state_dict())}
    for a in ("diffloss", "diffactloss"):
        for (k, p), (_, q) in zip(getattr(src, a).state_dict().items(), getattr(dst, a).state_dict().items()):
            assert torch.equal(p, q), (a, k)
    assert not torch.equal(src.encoder.weight, dst.encoder.weight)  # only the heads are touched

    # and back: bit-identical tensors under the reference's key names
    out = export_head_state_dict(dst)
    for k, v in out.items():
        assert torch.equal(v, policy_sd["module." + k]), k
    assert len(out) == n["diffloss"] + n["diffactloss"]


def test_strict_and_fallbacks():
    src, dst = _mar(0), _mar(1)
    sd = {"model." + k: v for k, v in src.state_dict().items() if "diffactloss" not in k}
    with pytest.raises(KeyError):
        load_reference_checkpoint(dst, {"state_dicts": {"model": sd}})
    assert load_reference_checkpoint(dst, {"state_dicts": {"model": sd}}, strict=False).keys() == {"diffloss"}
    bad = dict(sd)
    bad.pop("model.diffloss.net.cond_embed.weight")
    with pytest.raises(RuntimeError):
        load_reference_checkpoint(dst, {"state_dicts": {"model": bad}}, strict=True)
    # a bare MAR state_dict (no payload wrapper, no "model." prefix) is accepted as well
    assert set(head_state_dicts(src.state_dict())) == {"diffloss", "diffactloss"}
