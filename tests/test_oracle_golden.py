"""The CPU oracle (oracle/uva_oracle.py) against the golden vectors produced by the UNMODIFIED
reference (tests/golden/make_golden.py) and the float64 known-answer constants of SURVEY.md §8(c).
This is what pins the oracle; the GPU parity tests then compare the CUDA path with the oracle."""
import os

import numpy as np
import pytest
import torch

from oracle import uva_oracle as O

torch.set_num_threads(max(1, min(8, os.cpu_count() or 1)))


def _load(golden_dir, name):
    return np.load(os.path.join(golden_dir, name + ".npz"))


def _T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def test_schedule_tables_bit_exact(golden_dir):
    g = _load(golden_dir, "schedule")
    for tag, S in (("gen", O.create_schedule("100")), ("train", O.create_schedule(""))):
        for k in g.files:
            if not k.startswith(tag + "."):
                continue
            name = k.split(".", 1)[1]
            mine = getattr(S, name)
            assert np.array_equal(np.asarray(mine, dtype=g[k].dtype), g[k]), (tag, name)


def test_schedule_known_answers():
    S = O.create_schedule("100")
    assert list(S.timestep_map[:8]) == [0, 10, 20, 30, 40, 50, 61, 71]
    assert list(S.timestep_map[-4:]) == [969, 979, 989, 999]
    kat = {
        ("betas", 0): 4.128422482196914e-05, ("betas", 1): 6.798400685055217e-04,
        ("betas", 50): 3.417170133661673e-02, ("betas", 99): 0.9999899991985922,
        ("alphas_cumprod", 50): 0.48449452061639064, ("alphas_cumprod", 99): 2.4287669070228034e-09,
        ("posterior_log_variance_clipped", 0): -10.15394511055059,
        ("posterior_log_variance_clipped", 1): -10.15394511055059,
        ("posterior_log_variance_clipped", 50): -3.410175196467124,
        ("posterior_log_variance_clipped", 99): -2.5288514517641907e-04,
        ("posterior_mean_coef1", 1): 0.9427674281100751, ("posterior_mean_coef1", 99): 0.015583721365743979,
        ("posterior_mean_coef2", 50): 0.9500863770173814,
        ("sqrt_recip_alphas_cumprod", 99): 20291.169634711496,
        ("sqrt_recipm1_alphas_cumprod", 0): 0.006425412771141183,
        ("sqrt_recipm1_alphas_cumprod", 99): 20291.169610070236,
    }
    for (name, i), v in kat.items():
        assert getattr(S, name)[i] == pytest.approx(v, rel=1e-13), (name, i)
    Tr = O.create_schedule("")
    assert Tr.betas[999] == pytest.approx(0.999, rel=1e-12)
    assert Tr.sqrt_alphas_cumprod[500] == pytest.approx(0.7016303673935465, rel=1e-13)
    assert Tr.sqrt_one_minus_alphas_cumprod[500] == pytest.approx(0.7125411058677225, rel=1e-13)
    assert Tr.posterior_log_variance_clipped[0] == pytest.approx(-10.734082532465003, rel=1e-13)


def test_param_counts():
    n = sum(int(np.prod(s)) for _, s in O.param_specs(16, 768, 1024, 6))
    assert n == 35749920
    n = sum(int(np.prod(s)) for _, s in O.param_specs(2, 768, 1024, 6))
    assert n == 35706884


def test_small_head_forward_and_chain(golden_dir):
    g = _load(golden_dir, "small_head")
    P = O.to_torch_params({k[6:]: g[k] for k in g.files if k.startswith("param.")})
    out = O.net_forward(P, _T(g["net_x"]), _T(g["net_t"]), _T(g["net_c"]))
    np.testing.assert_allclose(out.numpy(), g["net_out"], rtol=2e-5, atol=2e-6)
    S = O.create_schedule("100")
    fin, states = O.sample_loop(P, S, _T(g["s_z"]), _T(g["s_x_init"]), _T(g["s_noise"]), temperature=0.9,
                                cfg=1.0, clip_denoised=False, trace=True)
    np.testing.assert_allclose(states.numpy(), g["s_states"], rtol=2e-4, atol=2e-5)
    np.testing.assert_allclose(fin.numpy(), g["s_final"], rtol=2e-4, atol=2e-5)


def test_video_head_forward(golden_dir):
    g = _load(golden_dir, "video_head")
    P = O.to_torch_params(O.make_params(16, 768, 1024, 6, seed=1))
    out = O.net_forward(P, _T(g["net_x"]), _T(g["net_t"]), _T(g["net_c"]))
    np.testing.assert_allclose(out.numpy(), g["net_out"], rtol=2e-5, atol=2e-6)


def test_video_head_teacher_forced_steps(golden_dir):
    """Per-step parity: feed the reference's state x_k to the oracle, compare x_{k+1}."""
    g = _load(golden_dir, "video_head")
    P = O.to_torch_params(O.make_params(16, 768, 1024, 6, seed=1))
    S = O.create_schedule("100")
    for tag, cfg in (("s1", 1.0), ("s2", 3.0)):
        states = _T(g[tag + "_states"])
        teacher = torch.cat([_T(g[tag + "_x_init"])[None], states[:-1]], dim=0)
        # run all 100 steps teacher-forced in one call
        _, tr = O.sample_loop(P, S, _T(g[tag + "_z"]), _T(g[tag + "_x_init"]), _T(g[tag + "_noise"]), temperature=0.95,
                              cfg=cfg, clip_denoised=False, trace=True, teacher=teacher)
        ref = states.numpy()
        err = np.abs(tr.numpy() - ref).reshape(100, -1).max(axis=1)
        scale = np.abs(ref).reshape(100, -1).max(axis=1)
        assert (err <= 1e-4 * scale + 1e-6).all(), (tag, float((err / scale).max()))


def test_action_head_chain_clip(golden_dir):
    g = _load(golden_dir, "action_head")
    P = O.to_torch_params(O.make_params(2, 768, 1024, 6, seed=2))
    S = O.create_schedule("100")
    states = _T(g["s_states"])
    teacher = torch.cat([_T(g["s_x_init"])[None], states[:-1]], dim=0)
    _, tr = O.sample_loop(P, S, _T(g["s_z"]), _T(g["s_x_init"]), _T(g["s_noise"]), temperature=0.95, cfg=1.0,
                          clip_denoised=True, trace=True, teacher=teacher)
    np.testing.assert_allclose(tr.numpy(), states.numpy(), rtol=1e-4, atol=1e-5)
    # free-running: clip_denoised keeps the chain bounded, so the end of the chain is comparable too
    fin = O.sample_loop(P, S, _T(g["s_z"]), _T(g["s_x_init"]), _T(g["s_noise"]), temperature=0.95, cfg=1.0,
                        clip_denoised=True)
    np.testing.assert_allclose(fin.numpy(), g["s_final"], rtol=1e-3, atol=1e-4)


def test_training_loss_and_grads(golden_dir):
    g = _load(golden_dir, "video_head")
    P = O.to_torch_params(O.make_params(16, 768, 1024, 6, seed=1), requires_grad=True)
    S = O.create_schedule("")
    z = _T(g["tr_z"]).reshape(-1, 768).requires_grad_(True)
    loss = O.diffloss_forward(P, S, _T(g["tr_target"]).reshape(-1, 16), z, _T(g["tr_mask"]).reshape(-1),
                              _T(g["tr_t"]), _T(g["tr_noise"]))
    loss.backward()
    assert loss.item() == pytest.approx(float(g["tr_loss"]), rel=1e-5)
    np.testing.assert_allclose(z.grad.numpy().reshape(g["tr_dz"].shape), g["tr_dz"], rtol=1e-3, atol=1e-7)
    for k, p in P.items():
        dg = g["tr_gdigest." + k]
        f = p.grad.double().flatten()
        mine = np.array([f.sum().item(), f.abs().sum().item(), f.norm().item()])
        np.testing.assert_allclose(mine[1:], dg[1:], rtol=1e-4, err_msg=k)
        np.testing.assert_allclose(f[_T(g["tr_gidx." + k])].numpy(), g["tr_gsample." + k], rtol=2e-3,
                                   atol=1e-6 * max(1.0, float(dg[2])), err_msg=k)


def test_t0_decoder_nll_rows(golden_dir):
    g = _load(golden_dir, "video_head")
    P = O.to_torch_params(O.make_params(16, 768, 1024, 6, seed=1))
    S = O.create_schedule("")
    with torch.no_grad():
        terms = O.training_losses(P, S, _T(g["t0_x"]), _T(g["t0_t"]), _T(g["t0_noise"]), _T(g["t0_c"]))
    np.testing.assert_allclose(terms["mse"].numpy(), g["t0_mse"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(terms["vb"].numpy(), g["t0_vb"], rtol=1e-4, atol=1e-6)
    np.testing.assert_allclose(terms["loss"].numpy(), g["t0_loss"], rtol=1e-4, atol=1e-6)


def test_action_training_rows(golden_dir):
    g = _load(golden_dir, "action_head")
    P = O.to_torch_params(O.make_params(2, 768, 1024, 6, seed=2))
    S = O.create_schedule("")
    with torch.no_grad():
        terms = O.training_losses(P, S, _T(g["tr_target"]), _T(g["tr_t"]), _T(g["tr_noise"]), _T(g["tr_z"]))
    np.testing.assert_allclose(terms["loss"].numpy(), g["tr_rowloss"], rtol=1e-4, atol=1e-6)
    assert terms["loss"].mean().item() == pytest.approx(float(g["tr_loss"]), rel=1e-5)


def test_philox_known_answers_and_moments():
    """The sampler's optional in-kernel noise generator.  Philox4x32-10 is pinned by the
    known-answer vectors published with Random123 (kat_vectors: "philox4x32 10 ..."), the normal transform
    by its moments."""
    def kat(c, k):
        return [int(v) for v in O.philox4x32_10(*[np.uint64(x) for x in c], k[0], k[1])]

    assert kat([0, 0, 0, 0], [0, 0]) == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    assert kat([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2) == [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]
    assert kat([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0]) == \
        [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]
    x = O.philox_normal(seed=1234, offset=0, step0=0, per_step=1 << 18, steps=3)
    assert abs(float(x.mean())) < 5e-3 and abs(float(x.std()) - 1.0) < 5e-3
    assert abs(float(np.mean(x.astype(np.float64) ** 4)) - 3.0) < 0.1  # kurtosis of a normal
    assert abs(float(np.corrcoef(x[0], x[1])[0, 1])) < 1e-2  # steps are independent streams
    assert not np.array_equal(x[0], O.philox_normal(1234, 1, 0, 1 << 18, 1)[0])  # the call offset matters


@pytest.mark.parametrize("A", [6, 9, 10])
def test_other_channel_counts(golden_dir, A):
    """Heads the reference builds with act_dim = 10 (Libero / UMI / ToolHang) and with 6 / 9 proprioception channels
    (MAR:298, 317, 332): sampling chain with clip_denoised and the training rows + gradients, against
    tests/golden/head_dims.npz (make_golden_dims.py)."""
    g = _load(golden_dir, "head_dims")
    p = f"a{A}."
    Pn = O.make_params(A, 768, 1024, 6, seed=20 + A)
    P = O.to_torch_params(Pn)
    S = O.create_schedule("100")
    states = _T(g[p + "s_states"])
    teacher = torch.cat([_T(g[p + "s_x_init"])[None], states[:-1]], dim=0)
    _, tr = O.sample_loop(P, S, _T(g[p + "s_z"]), _T(g[p + "s_x_init"]), _T(g[p + "s_noise"]), temperature=0.95, cfg=1.0,
                          clip_denoised=True, trace=True, teacher=teacher)
    np.testing.assert_allclose(tr.numpy(), states.numpy(), rtol=1e-4, atol=1e-5)
    Pg = O.to_torch_params(Pn, requires_grad=True)
    St = O.create_schedule("")
    z = _T(g[p + "tr_z"]).requires_grad_(True)
    loss = O.diffactloss_forward(Pg, St, _T(g[p + "tr_target"]), z, _T(g[p + "tr_t"]), _T(g[p + "tr_noise"]))
    loss.backward()
    assert loss.item() == pytest.approx(float(g[p + "tr_loss"]), rel=1e-5)
    np.testing.assert_allclose(z.grad.numpy(), g[p + "tr_dz"], rtol=1e-3, atol=1e-7)
    for k, q in Pg.items():
        dg = g[p + "tr_gdigest." + k]
        f = q.grad.double().flatten()
        np.testing.assert_allclose(np.array([f.abs().sum().item(), f.norm().item()]), dg[1:], rtol=1e-4, err_msg=k)


def test_decoder_tail_restatement_vs_golden(golden_dir):
    """oracle.decoder_tail (MAR:702-724) against the statements of the reference executed on nn.LayerNorm
    (tests/golden/make_golden_tail.py): bit-exact, same ATen ops."""
    g = np.load(os.path.join(golden_dir, "decoder_tail.npz"))
    z = O.decoder_tail(torch.from_numpy(g["x_dec"]), torch.from_numpy(g["ln_w"]), torch.from_numpy(g["ln_b"]),
                       float(g["eps"][0]), torch.from_numpy(g["temporal"]), torch.from_numpy(g["spatial"]),
                       skip=int(g["meta"][4]))
    assert z.shape == g["z"].shape
    assert torch.equal(z, torch.from_numpy(g["z"]))


PRENET_NAMES = ("conv.0.weight", "conv.0.bias", "fc.0.weight", "fc.0.bias", "fc.2.weight", "fc.2.bias",
                "interpolate.weight", "interpolate.bias", "refine.0.weight", "refine.0.bias", "refine.2.weight", "refine.2.bias")


def prenet_case(golden_dir):
    """The seeded module / inputs of tests/golden/make_golden_prenet.py, rebuilt without the reference."""
    from unified_video_action_b200 import DiffActLoss

    g = np.load(os.path.join(golden_dir, "prenet.npz"))
    seed, B = [int(v) for v in g["seed"]]
    torch.manual_seed(seed)
    mod = DiffActLoss(2, 768, 6, 1024, "100", n_frames=4, language_emb_model=None, language_emb_model_type=None)
    sd = dict(mod.named_parameters())
    for n in PRENET_NAMES:  # same constructor order as the reference => same weights from the same seed
        f = sd[n].detach().double().flatten()
        np.testing.assert_allclose([float(f.sum()), float(f.abs().sum()), float(f.norm())], g["wdigest." + n], rtol=1e-12)
    gen = torch.Generator().manual_seed(seed + 1)
    z = torch.randn(B, 1024, 768, generator=gen)
    torch.rand(B, 16, 2, generator=gen)
    cot = torch.randn(B, 16, 768, generator=gen)
    return g, mod, z, cot


def test_prenet_restatement_vs_reference_golden(golden_dir):
    """oracle.prenet_conv_fc (DAL:112-124) against the unmodified reference's pre-net output, d z and parameter
    gradients (tests/golden/prenet.npz): fp32, same ATen ops."""
    g, mod, z, cot = prenet_case(golden_dir)
    Pp = {n: p.detach().clone().requires_grad_(True) for n, p in mod.named_parameters() if n in PRENET_NAMES}
    zz = z.clone().requires_grad_(True)
    out = O.prenet_conv_fc(Pp, zz)
    assert torch.equal(out.detach(), torch.from_numpy(g["out"]))
    (out * cot).sum().backward()
    np.testing.assert_allclose(zz.grad.flatten()[torch.from_numpy(g["dz_idx"])].numpy(), g["dz_val"], rtol=1e-5, atol=1e-8)
    for n in PRENET_NAMES:
        np.testing.assert_allclose(Pp[n].grad.flatten()[torch.from_numpy(g["gidx." + n])].numpy(), g["gval." + n], rtol=1e-4, atol=1e-7)
    # the drop-in module's PyTorch pre-net (fp32 parity mode / CPU) is the same op sequence
    with torch.no_grad():
        assert torch.equal(mod.prenet_torch(z), torch.from_numpy(g["out"]))
