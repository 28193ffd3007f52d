"""patch(mar) — the call INTEGRATION.md tells a maintainer to make (replaces MAR:269-344's head construction).

A stub MAR holding the REFERENCE's own heads (imported from /root/reference, which exists in the build container
only: skipped elsewhere) in the Libero multi-view layout (BASELINE config 4: diffloss + diffloss_wrist +
diffactloss with act_dim 10 + diffproploss with 9 channels) is patched; the replacement heads must carry the same
state_dict keys and tensors, the same geometry and diffusion lengths, and the side-stream wiring."""
import os
import sys

import pytest
import torch

REF = "/root/reference"


def _ref_heads():
    if not os.path.isdir(os.path.join(REF, "unified_video_action")):
        pytest.skip("reference sources are not on this box")
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from unified_video_action.model.autoregressive.diffusion_action_loss import DiffActLoss as RefAct
    from unified_video_action.model.autoregressive.diffusion_loss import DiffLoss as RefLoss

    return RefLoss, RefAct


def test_patch_swaps_reference_heads_in_place():
    RefLoss, RefAct = _ref_heads()
    from unified_video_action_b200 import DiffActLoss, DiffLoss, patch

    kw = dict(n_frames=4, language_emb_model=None, language_emb_model_type=None)

    class StubMAR(torch.nn.Module):  # the attribute names of MAR:269-344
        def __init__(self):
            super().__init__()
            torch.manual_seed(0)
            self.diffloss = RefLoss(target_channels=16, z_channels=768, depth=6, width=1024, num_sampling_steps="100",
                                    grad_checkpointing=False, **kw)
            self.diffloss_wrist = RefLoss(target_channels=16, z_channels=768, depth=6, width=1024,
                                          num_sampling_steps="100", grad_checkpointing=False, **kw)
            self.diffactloss = RefAct(target_channels=10, z_channels=768, depth=6, width=1024, num_sampling_steps="100",
                                      grad_checkpointing=False, act_diff_training_steps=1000,
                                      act_diff_testing_steps="100", act_model_type="conv_fc", **kw)
            self.diffproploss = RefAct(target_channels=9, z_channels=768, depth=6, width=1024, num_sampling_steps="100",
                                       grad_checkpointing=False, act_diff_training_steps=1000,
                                       act_diff_testing_steps="100", act_model_type="conv_fc", **kw)
            self.other = torch.nn.Linear(3, 3)  # anything else on the MAR is left alone

    mar = StubMAR()
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for p in mar.parameters():  # the zero-initialised tensors must survive the copy as what they are trained to
            if float(p.abs().max()) == 0.0:
                p.copy_(0.02 * torch.randn(p.shape, generator=g))
    before = {a: {k: v.clone() for k, v in getattr(mar, a).state_dict().items()}
              for a in ("diffloss", "diffloss_wrist", "diffactloss", "diffproploss")}
    other = mar.other
    assert patch(mar) is mar
    assert mar.other is other
    for a, want in before.items():
        new = getattr(mar, a)
        assert isinstance(new, DiffActLoss if a in ("diffactloss", "diffproploss") else DiffLoss), a
        sd = new.state_dict()
        assert list(sd.keys()) == list(want.keys()), a
        for k in want:
            assert sd[k].dtype == want[k].dtype and torch.equal(sd[k], want[k]), (a, k)
        assert new.net.in_channels == {"diffloss": 16, "diffloss_wrist": 16, "diffactloss": 10, "diffproploss": 9}[a]
        assert new.num_sampling_steps == 100 and new.num_timesteps == 1000
    # the small heads run beside the video head (MAR.forward_loss calls diffloss first with the same z)
    assert mar.diffactloss._fork_from() is mar.diffloss and mar.diffproploss._fork_from() is mar.diffloss
    assert mar.diffloss._record_entry
    # idempotent: a second patch() leaves the B200 heads in place
    ids = [id(getattr(mar, a)) for a in before]
    patch(mar)
    assert ids == [id(getattr(mar, a)) for a in before]
    # a MAR without the optional heads (PushT: no wrist view, no proprioception)
    class Small(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.diffloss = RefLoss(target_channels=16, z_channels=768, depth=6, width=1024, num_sampling_steps="100",
                                    grad_checkpointing=False, **kw)
    small = patch(Small())
    assert isinstance(small.diffloss, DiffLoss) and not hasattr(small, "diffactloss")
