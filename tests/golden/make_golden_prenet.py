"""Golden vectors of the DiffActLoss "conv_fc" pre-net (diffusion_action_loss.py:41-61, 112-124) from the UNMODIFIED
reference -> tests/golden/prenet.npz.  The reference module is built from a seed (the replacement module built from
the same seed holds bit-identical weights: same constructor order, checked through the digests below), its forward
runs on a seeded z [2, 1024, 768]; the pre-net output is taken from a forward hook on `refine` (the last pre-net
module, DAL:124) and differentiated against a seeded cotangent: d z and every pre-net parameter gradient.
Run here once:  python tests/golden/make_golden_prenet.py"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, "/root/reference")
SEED, B = 11, 2
NAMES = ("conv.0.weight", "conv.0.bias", "fc.0.weight", "fc.0.bias", "fc.2.weight", "fc.2.bias",
         "interpolate.weight", "interpolate.bias", "refine.0.weight", "refine.0.bias", "refine.2.weight", "refine.2.bias")


def digest(t):
    f = t.detach().double().flatten()
    return np.array([float(f.sum()), float(f.abs().sum()), float(f.norm())])


def main():
    from unified_video_action.model.autoregressive.diffusion_action_loss import DiffActLoss

    torch.manual_seed(SEED)
    ref = DiffActLoss(target_channels=2, z_channels=768, depth=6, width=1024, num_sampling_steps="100", grad_checkpointing=False,
                      n_frames=4, act_diff_training_steps=1000, act_diff_testing_steps="100", act_model_type="conv_fc",
                      language_emb_model=None, language_emb_model_type=None)
    g = torch.Generator().manual_seed(SEED + 1)
    z = torch.randn(B, 1024, 768, generator=g).requires_grad_(True)
    target = torch.rand(B, 16, 2, generator=g) * 2 - 1
    cot = torch.randn(B, 16, 768, generator=g)
    box = {}
    ref.refine.register_forward_hook(lambda m, i, o: box.__setitem__("out", o))
    torch.manual_seed(0)
    ref(target, z)  # the reference's own forward (DAL:109-166); the hook keeps the pre-net output
    out = box["out"]
    (out * cot).sum().backward()
    sd = dict(ref.named_parameters())
    data = {"seed": np.array([SEED, B]), "out": out.detach().numpy(), "dz_digest": digest(z.grad)}
    gi = torch.Generator().manual_seed(3)
    idx = torch.randperm(z.grad.numel(), generator=gi)[:4096]
    data["dz_idx"], data["dz_val"] = idx.numpy(), z.grad.flatten()[idx].numpy()
    for n in NAMES:
        p = sd[n]
        data["wdigest." + n] = digest(p)
        data["gdigest." + n] = digest(p.grad)
        k = min(2048, p.numel())
        idx = torch.randperm(p.numel(), generator=gi)[:k]
        data["gidx." + n], data["gval." + n] = idx.numpy(), p.grad.flatten()[idx].numpy()
    np.savez_compressed(os.path.join(HERE, "prenet.npz"), **data)
    print("wrote prenet.npz; |out| max", float(out.abs().max()))


if __name__ == "__main__":
    main()
