"""Generate golden vectors from the UNMODIFIED reference (run in the build container only).

    PYTHONPATH=/root/repo python tests/golden/make_golden.py

Imports ``unified_video_action.model.autoregressive.{diffusion_loss,diffusion_action_loss}``
from /root/reference, loads deterministic non-degenerate parameters (oracle.make_params, numpy
PCG64 -> platform independent), runs the reference on seeded inputs on CPU and writes the
input/output vectors to ``tests/golden/*.npz``.  The reference cannot travel to the GPU box, the
``.npz`` files do.

``DiffLoss.sample`` / ``DiffActLoss.sample`` hard-code ``.cuda()`` (DL:71,76; DAL:210,215;
GD:518,528) so on CPU the loop GD:519-541 is mirrored here by calling the reference's own
``gen_diffusion.p_sample`` (GD:395-444) step by step; every arithmetic op is still the reference's.
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from unified_video_action.model.autoregressive.diffusion_loss import DiffLoss  # noqa: E402
from unified_video_action.model.autoregressive.diffusion_action_loss import DiffActLoss  # noqa: E402

from oracle import uva_oracle as O  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
KW = dict(n_frames=4, language_emb_model=None, language_emb_model_type=None)
torch.set_num_threads(8)


def load_net(head, params):
    sd = {k: torch.from_numpy(v.copy()) for k, v in params.items()}
    head.net.load_state_dict(sd, strict=True)


def ref_sample(head, z, x_init, cfg, temperature, clip):
    """GD:494-541 loop on CPU; records the randn_like drawn inside p_sample each step."""
    diff = head.gen_diffusion
    if cfg != 1.0:
        model_kwargs = dict(c=z, cfg_scale=cfg)
        fn = head.net.forward_with_cfg
    else:
        model_kwargs = dict(c=z)
        fn = head.net.forward
    img = x_init
    states, noises = [], []
    for i in list(range(diff.num_timesteps))[::-1]:
        t = torch.tensor([i] * x_init.shape[0])
        st = torch.get_rng_state()
        noises.append(torch.randn_like(img))  # what p_sample will draw (GD:431)
        torch.set_rng_state(st)
        with torch.no_grad():
            out = diff.p_sample(fn, img, t, clip_denoised=clip, model_kwargs=model_kwargs, temperature=temperature)
        img = out["sample"]
        states.append(img.clone())
    return img, torch.stack(states), torch.stack(noises)


def grad_digest(g: torch.Tensor):
    f = g.detach().double().flatten()
    idx = torch.linspace(0, f.numel() - 1, steps=min(64, f.numel())).long()
    return np.array([f.sum().item(), f.abs().sum().item(), f.norm().item()]), f[idx].numpy(), idx.numpy()


def main():
    # ---------------- schedules: full float64 tables ----------------
    h = DiffLoss(target_channels=4, z_channels=8, depth=1, width=16, num_sampling_steps="100", **KW)
    sched = {}
    for tag, d in (("gen", h.gen_diffusion), ("train", h.train_diffusion)):
        for name in (
            "betas alphas_cumprod alphas_cumprod_prev sqrt_alphas_cumprod sqrt_one_minus_alphas_cumprod "
            "sqrt_recip_alphas_cumprod sqrt_recipm1_alphas_cumprod posterior_variance "
            "posterior_log_variance_clipped posterior_mean_coef1 posterior_mean_coef2"
        ).split():
            sched[f"{tag}.{name}"] = np.asarray(getattr(d, name), dtype=np.float64)
        sched[f"{tag}.timestep_map"] = np.asarray(d.timestep_map, dtype=np.int64)
    np.savez_compressed(os.path.join(OUT, "schedule.npz"), **sched)

    # ---------------- small geometry (weights travel inside the file) ----------------
    Cs, Zs, Ws, Ds = 4, 32, 64, 2
    ps = O.make_params(Cs, Zs, Ws, Ds, seed=11)
    head = DiffLoss(target_channels=Cs, z_channels=Zs, depth=Ds, width=Ws, num_sampling_steps="100", **KW)
    load_net(head, ps)
    g = torch.Generator().manual_seed(5)
    N = 12
    x = torch.randn(N, Cs, generator=g)
    z = torch.randn(N, Zs, generator=g)
    t = torch.tensor([0, 1, 10, 50, 61, 500, 998, 999, 3, 250, 750, 999])
    with torch.no_grad():
        out = head.net(x, t, z)
    torch.manual_seed(21)
    x_init = torch.randn(N, Cs)
    fin, states, noises = ref_sample(head, z, x_init, cfg=1.0, temperature=0.9, clip=False)
    small = {f"param.{k}": v for k, v in ps.items()}
    small.update(dict(net_x=x.numpy(), net_t=t.numpy(), net_c=z.numpy(), net_out=out.numpy(),
                      s_z=z.numpy(), s_x_init=x_init.numpy(), s_noise=noises.numpy(), s_states=states.numpy(),
                      s_final=fin.numpy(), geometry=np.array([Cs, Zs, Ws, Ds])))
    np.savez_compressed(os.path.join(OUT, "small_head.npz"), **small)

    # ---------------- full geometry, video head (C=16) ----------------
    C, Z, W, D = 16, 768, 1024, 6
    pv = O.make_params(C, Z, W, D, seed=1)
    head = DiffLoss(target_channels=C, z_channels=Z, depth=D, width=W, num_sampling_steps="100", **KW)
    load_net(head, pv)
    assert sum(p.numel() for p in head.parameters()) == 35749920
    g = torch.Generator().manual_seed(7)
    N = 8
    x = torch.randn(N, C, generator=g)
    z = torch.randn(N, Z, generator=g)
    t = torch.tensor([0, 10, 61, 500, 999, 989, 3, 250])
    with torch.no_grad():
        out = head.net(x, t, z)
    full = dict(net_x=x.numpy(), net_t=t.numpy(), net_c=z.numpy(), net_out=out.numpy())

    # sampling, cfg = 1, clip False (DiffLoss.sample, DL:75-88)
    torch.manual_seed(31)
    x_init = torch.randn(N, C)
    fin, states, noises = ref_sample(head, z, x_init, cfg=1.0, temperature=0.95, clip=False)
    full.update(s1_z=z.numpy(), s1_x_init=x_init.numpy(), s1_noise=noises.numpy(), s1_states=states.numpy(), s1_final=fin.numpy())

    # sampling with CFG = 3 (DL:70-74): z rows cond || uncond, noise duplicated
    torch.manual_seed(32)
    half = torch.randn(N // 2, C)
    x_init = torch.cat([half, half], dim=0)
    fin, states, noises = ref_sample(head, z, x_init, cfg=3.0, temperature=0.95, clip=False)
    full.update(s2_z=z.numpy(), s2_x_init=x_init.numpy(), s2_noise=noises.numpy(), s2_states=states.numpy(), s2_final=fin.numpy())

    # training: DiffLoss.forward + backward (DL:44-66), bsz=2, seq=8
    bsz, seq = 2, 8
    g = torch.Generator().manual_seed(9)
    target = 0.5 * torch.randn(bsz, seq, C, generator=g)
    zt = torch.randn(bsz, seq, Z, generator=g).requires_grad_(True)
    mask = (torch.rand(bsz, seq, generator=g) < 0.75).float()
    torch.manual_seed(41)
    st = torch.get_rng_state()
    t_draw = torch.randint(0, 1000, (bsz * seq,))
    n_draw = torch.randn(bsz * seq, C)
    torch.set_rng_state(st)
    head.zero_grad()
    loss = head(target, zt, mask)
    loss.backward()
    full.update(tr_target=target.numpy(), tr_z=zt.detach().numpy(), tr_mask=mask.numpy(), tr_t=t_draw.numpy(),
                tr_noise=n_draw.numpy(), tr_loss=np.array(loss.item(), dtype=np.float64), tr_dz=zt.grad.numpy())
    for k, p_ in head.net.named_parameters():
        dg, samp, idx = grad_digest(p_.grad)
        full[f"tr_gdigest.{k}"] = dg
        full[f"tr_gsample.{k}"] = samp
        full[f"tr_gidx.{k}"] = idx
    # forced t == 0 rows (decoder-NLL branch, GD:739-743) -> per-row losses from training_losses
    t0 = torch.tensor([0, 0, 1, 999, 0, 5, 0, 2])
    x0 = torch.tensor(np.linspace(-1.2, 1.2, 8 * C).reshape(8, C), dtype=torch.float32)
    nz = torch.randn(8, C, generator=g)
    with torch.no_grad():
        terms = head.train_diffusion.training_losses(head.net, x0, t0, dict(c=z), noise=nz)
    full.update(t0_x=x0.numpy(), t0_t=t0.numpy(), t0_noise=nz.numpy(), t0_c=z.numpy(),
                t0_loss=terms["loss"].numpy(), t0_mse=terms["mse"].numpy(), t0_vb=terms["vb"].numpy())
    np.savez_compressed(os.path.join(OUT, "video_head.npz"), **full)

    # ---------------- full geometry, action head (A=2, PushT) ----------------
    A = 2
    act = DiffActLoss(target_channels=A, z_channels=Z, depth=D, width=W, num_sampling_steps="100", **KW)
    pa = O.make_params(A, Z, W, D, seed=2)
    load_net(act, pa)
    assert sum(p.numel() for p in act.parameters()) == 52225876
    g = torch.Generator().manual_seed(13)
    N = 16
    zr = torch.randn(N, Z, generator=g)  # rows after the pre-net
    torch.manual_seed(51)
    x_init = torch.randn(N, A)
    fin, states, noises = ref_sample(act, zr, x_init, cfg=1.0, temperature=0.95, clip=True)  # DAL:219-227
    a = dict(s_z=zr.numpy(), s_x_init=x_init.numpy(), s_noise=noises.numpy(), s_states=states.numpy(), s_final=fin.numpy())
    # training on post-pre-net rows through the reference's training_losses (DAL:145-166)
    tgt = torch.rand(N, A, generator=g) * 2 - 1
    tt = torch.randint(0, 1000, (N,), generator=g)
    nz = torch.randn(N, A, generator=g)
    zr2 = zr.clone().requires_grad_(True)
    act.zero_grad()
    terms = act.train_diffusion.training_losses(act.net, tgt, tt, dict(c=zr2), noise=nz)
    loss = terms["loss"].mean()
    loss.backward()
    a.update(tr_target=tgt.numpy(), tr_t=tt.numpy(), tr_noise=nz.numpy(), tr_z=zr.numpy(),
             tr_loss=np.array(loss.item(), dtype=np.float64), tr_rowloss=terms["loss"].detach().numpy(), tr_dz=zr2.grad.numpy())
    for k, p_ in act.net.named_parameters():
        dg, samp, idx = grad_digest(p_.grad)
        a[f"tr_gdigest.{k}"] = dg
        a[f"tr_gsample.{k}"] = samp
        a[f"tr_gidx.{k}"] = idx
    np.savez_compressed(os.path.join(OUT, "action_head.npz"), **a)
    for f in ("schedule", "small_head", "video_head", "action_head"):
        print(f, os.path.getsize(os.path.join(OUT, f + ".npz")))


if __name__ == "__main__":
    main()
