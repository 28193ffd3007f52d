"""Golden vectors for the other channel counts the reference builds heads with (run in the build container only).

    PYTHONPATH=/root/repo python tests/golden/make_golden_dims.py

The action head has ``target_channels = act_dim``: 2 on PushT (covered by action_head.npz) and 10 on Libero, UMI and
ToolHang (config/task/{libero10,umi,toolhang}.yaml ``shape_meta.action.shape``); the proprioception heads have 6 and 9
(MAR:317, 332).  For A in {6, 9, 10} this runs the UNMODIFIED reference ``DiffActLoss`` on CPU (same method as
make_golden.py: the reference's own ``p_sample`` step by step, its ``training_losses`` + autograd) and writes
``tests/golden/head_dims.npz``; parameters come from ``oracle.make_params(A, ..., seed=20 + A)``.
"""
import os

import numpy as np
import torch

import make_golden as G  # sets sys.path for the reference and the repo
from make_golden import DiffActLoss, KW, O, grad_digest, load_net, ref_sample

OUT = os.path.dirname(os.path.abspath(__file__))
Z, W, D = 768, 1024, 6


def main():
    out = {}
    for A in (6, 9, 10):
        act = DiffActLoss(target_channels=A, z_channels=Z, depth=D, width=W, num_sampling_steps="100", **KW)
        load_net(act, O.make_params(A, Z, W, D, seed=20 + A))
        g = torch.Generator().manual_seed(100 + A)
        N = 16  # one predict_action call (B = 1)
        zr = torch.randn(N, Z, generator=g)
        torch.manual_seed(200 + A)
        x_init = torch.randn(N, A)
        fin, states, noises = ref_sample(act, zr, x_init, cfg=1.0, temperature=0.95, clip=True)  # DAL:219-227
        p = f"a{A}."
        out.update({p + "s_z": zr.numpy(), p + "s_x_init": x_init.numpy(), p + "s_noise": noises.numpy(),
                    p + "s_states": states.numpy(), p + "s_final": fin.numpy()})
        # training rows (DAL:145-166): 24 rows so that row counts that are no multiple of 16 are seen too
        M = 24
        zt = torch.randn(M, Z, generator=g)
        tgt = torch.rand(M, A, generator=g) * 2 - 1
        tt = torch.randint(0, 1000, (M,), generator=g)
        tt[0] = 0  # the decoder-NLL branch (GD:739-743)
        nz = torch.randn(M, A, generator=g)
        z2 = zt.clone().requires_grad_(True)
        act.zero_grad()
        terms = act.train_diffusion.training_losses(act.net, tgt, tt, dict(c=z2), noise=nz)
        loss = terms["loss"].mean()
        loss.backward()
        out.update({p + "tr_target": tgt.numpy(), p + "tr_t": tt.numpy(), p + "tr_noise": nz.numpy(), p + "tr_z": zt.numpy(),
                    p + "tr_loss": np.array(loss.item(), dtype=np.float64), p + "tr_rowloss": terms["loss"].detach().numpy(),
                    p + "tr_dz": z2.grad.numpy()})
        for k, p_ in act.net.named_parameters():
            dg, samp, idx = grad_digest(p_.grad)
            out[p + f"tr_gdigest.{k}"] = dg
            out[p + f"tr_gsample.{k}"] = samp
            out[p + f"tr_gidx.{k}"] = idx
    path = os.path.join(OUT, "head_dims.npz")
    np.savez_compressed(path, **out)
    print("head_dims", os.path.getsize(path))


if __name__ == "__main__":
    assert G.REF == "/root/reference"
    main()
