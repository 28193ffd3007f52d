"""GPU parity tests of the training path (loss forward + hand-written backward) through the C ABI.

Reference: DiffLoss.forward / DiffActLoss.forward (DL:44-66, DAL:145-166) -> training_losses
(GD:746-818) + autograd.  Checked against (i) the golden vectors produced by the unmodified reference
(loss, dz, per-parameter gradient digests and samples), (ii) the CPU oracle's autograd on seeded inputs
with ragged row counts, masks with zeros, t = 0 rows (decoder NLL branch) and the action-head shape.

Tolerances: fp32 mode <= 1e-4 relative (max-norm per tensor, against the tensor's own max); bf16 mode:
loss <= 1e-2 relative, gradients <= 2e-2 relative L2 per tensor (bf16 operands in forward and backward
GEMMs, fp32 accumulation; the reference trains under bf16 autocast the same way).
"""
import os

import numpy as np
import pytest
import torch

from oracle import uva_oracle as O

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-4
BF16_LOSS_TOL = 1e-2
BF16_GRAD_L2 = 2e-2


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)


def _T(a, dev=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    return t.to(dev) if dev is not None else t


def _head(C, Z, W, D, params, mode, dev):
    from unified_video_action_b200 import _lib

    names = [n for n, _ in O.param_specs(C, Z, W, D)]
    with torch.cuda.device(dev):
        h = _lib.Head(C, Z, W, D, 1000, 100, mode)
        h.load_weights([_T(params[n], dev) for n in names])
    return h, names


def _run(h, names, params, dev, target, z, mask, t, noise, dloss=1.0):
    grads = [torch.zeros(params[n].shape, dtype=torch.float32, device=dev) for n in names]
    m = None if mask is None else mask.to(dev)
    loss, rows = h.loss_fwd(target.to(dev), z.to(dev), m, t.to(dev), noise.to(dev), want_rows=True)
    dz = h.loss_bwd(torch.tensor([dloss], device=dev), target.to(dev), z.to(dev), m, t.to(dev), noise.to(dev), grads)
    torch.cuda.synchronize()
    return loss.cpu(), rows.cpu(), dz.cpu(), {n: g.cpu() for n, g in zip(names, grads)}


def _oracle(P_np, target, z, mask, t, noise, act=False):
    P = O.to_torch_params(P_np, requires_grad=True)
    S = O.create_schedule("")
    zz = z.clone().requires_grad_(True)
    if act:
        loss = O.diffactloss_forward(P, S, target, zz, t, noise)
    else:
        loss = O.diffloss_forward(P, S, target, zz, mask, t, noise)
    loss.backward()
    with torch.no_grad():
        rows = O.training_losses(P, S, target, t, noise, z)["loss"]
    return loss.detach(), rows, zz.grad, {k: p.grad for k, p in P.items()}


def _check(mine, ref, mode, what):
    mine, ref = mine.double(), ref.double()
    scale = float(ref.abs().max())
    if scale == 0.0:
        assert float(mine.abs().max()) == 0.0, what
        return
    if mode == 0:
        err = float((mine - ref).abs().max()) / scale
        assert err <= FP32_TOL, (what, err)
    else:
        err = float((mine - ref).norm() / ref.norm())
        assert err <= BF16_GRAD_L2, (what, err)


@pytest.fixture(scope="module")
def video(golden_dir):
    dev = _cuda()
    g = np.load(os.path.join(golden_dir, "video_head.npz"))
    P = O.make_params(16, 768, 1024, 6, seed=1)
    return dev, g, P


@pytest.mark.parametrize("mode", [0, 1], ids=["fp32", "bf16"])
def test_video_loss_and_grads_vs_reference_golden(video, mode):
    """16 rows, masked mean: loss, dz and every parameter gradient against the unmodified reference."""
    dev, g, P = video
    h, names = _head(16, 768, 1024, 6, P, mode, dev)
    loss, _, dz, grads = _run(h, names, P, dev, _T(g["tr_target"]).reshape(-1, 16), _T(g["tr_z"]).reshape(-1, 768),
                              _T(g["tr_mask"]).reshape(-1), _T(g["tr_t"]), _T(g["tr_noise"]))
    tol = FP32_TOL if mode == 0 else BF16_LOSS_TOL
    assert abs(float(loss) - float(g["tr_loss"])) <= tol * abs(float(g["tr_loss"]))
    _check(dz, _T(g["tr_dz"]).reshape(-1, 768), mode, "dz")
    for k in names:
        dg = g["tr_gdigest." + k]  # [sum, sum|.|, l2] of the reference gradient (float64)
        f = grads[k].double().flatten()
        rel = FP32_TOL * 10 if mode == 0 else BF16_GRAD_L2
        assert abs(float(f.norm()) - dg[2]) <= rel * max(dg[2], 1e-12), (k, float(f.norm()), dg[2])
        idx = _T(g["tr_gidx." + k])
        ref = _T(g["tr_gsample." + k]).double()
        err = float((f[idx] - ref).abs().max())
        bound = (FP32_TOL if mode == 0 else BF16_GRAD_L2) * max(float(ref.abs().max()), dg[2] / max(1.0, f.numel() ** 0.5))
        assert err <= 4 * bound, (k, err, bound)


@pytest.mark.parametrize("mode", [0, 1], ids=["fp32", "bf16"])
def test_t0_decoder_nll_rows_vs_reference_golden(video, mode):
    """Rows with t = 0 take the discretised-Gaussian NLL branch (DU:47-73)."""
    dev, g, P = video
    h, names = _head(16, 768, 1024, 6, P, mode, dev)
    x, t, nz, c = _T(g["t0_x"]), _T(g["t0_t"]), _T(g["t0_noise"]), _T(g["t0_c"])
    _, rows, _, _ = _run(h, names, P, dev, x, c, None, t, nz)
    tol = 2e-4 if mode == 0 else 2e-2
    np.testing.assert_allclose(rows.numpy(), g["t0_loss"], rtol=tol, atol=tol * float(np.abs(g["t0_loss"]).max()))


@pytest.mark.parametrize("mode", [0, 1], ids=["fp32", "bf16"])
def test_ragged_masked_rows_vs_oracle(video, mode):
    """333 rows (not a tile multiple), mask with zeros, t = 0 and t = 999 rows included."""
    dev, _, P = video
    gen = torch.Generator().manual_seed(11)
    N = 333
    target = torch.randn(N, 16, generator=gen) * 0.5
    target[5, 3] = -1.2  # beyond the +-0.999 thresholds of the decoder NLL (DU:66-72)
    target[6, 1] = 1.3
    z = torch.randn(N, 768, generator=gen)
    mask = (torch.rand(N, generator=gen) > 0.3).float()
    t = torch.randint(0, 1000, (N,), generator=gen)
    t[:12] = 0
    t[12] = 999
    noise = torch.randn(N, 16, generator=gen)
    h, names = _head(16, 768, 1024, 6, P, mode, dev)
    loss, rows, dz, grads = _run(h, names, P, dev, target, z, mask, t, noise, dloss=0.7)
    rl, rr, rdz, rg = _oracle(P, target, z, mask, t, noise)
    tol = FP32_TOL if mode == 0 else BF16_LOSS_TOL
    assert abs(float(loss) - float(rl)) <= tol * abs(float(rl))
    _check(rows, rr, mode, "row_loss")
    _check(dz, 0.7 * rdz, mode, "dz")
    for k in names:
        _check(grads[k], 0.7 * rg[k], mode, k)


@pytest.mark.parametrize("mode", [0, 1], ids=["fp32", "bf16"])
def test_action_head_plain_mean_vs_oracle(golden_dir, mode):
    """DiffActLoss shape: C = 2, 256 rows (PushT B=16), unmasked mean (DAL:162-164)."""
    dev = _cuda()
    P = O.make_params(2, 768, 1024, 6, seed=2)
    gen = torch.Generator().manual_seed(5)
    N = 256
    target = torch.rand(N, 2, generator=gen) * 2 - 1
    z = torch.randn(N, 768, generator=gen)
    t = torch.randint(0, 1000, (N,), generator=gen)
    noise = torch.randn(N, 2, generator=gen)
    h, names = _head(2, 768, 1024, 6, P, mode, dev)
    loss, rows, dz, grads = _run(h, names, P, dev, target, z, None, t, noise)
    rl, rr, rdz, rg = _oracle(P, target, z, None, t, noise, act=True)
    tol = FP32_TOL if mode == 0 else BF16_LOSS_TOL
    assert abs(float(loss) - float(rl)) <= tol * abs(float(rl))
    _check(dz, rdz, mode, "dz")
    for k in names:
        _check(grads[k], rg[k], mode, k)
    g = np.load(os.path.join(golden_dir, "action_head.npz"))
    h2, names2 = _head(2, 768, 1024, 6, P, mode, dev)
    _, rows2, _, _ = _run(h2, names2, P, dev, _T(g["tr_target"]), _T(g["tr_z"]), None, _T(g["tr_t"]), _T(g["tr_noise"]))
    rt = 2e-4 if mode == 0 else 2e-2
    np.testing.assert_allclose(rows2.numpy(), g["tr_rowloss"], rtol=rt, atol=rt * float(np.abs(g["tr_rowloss"]).max()))


def test_small_geometry_fp32_vs_oracle():
    """Generic-width kernels (W = 64, depth 2, C = 3, Z = 32)."""
    dev = _cuda()
    C, Z, W, D = 3, 32, 64, 2
    P = O.make_params(C, Z, W, D, seed=3)
    gen = torch.Generator().manual_seed(8)
    N = 70
    target = torch.randn(N, C, generator=gen)
    z = torch.randn(N, Z, generator=gen)
    mask = (torch.rand(N, generator=gen) > 0.2).float()
    t = torch.randint(0, 1000, (N,), generator=gen)
    t[0] = 0
    noise = torch.randn(N, C, generator=gen)
    h, names = _head(C, Z, W, D, P, 0, dev)
    loss, _, dz, grads = _run(h, names, P, dev, target, z, mask, t, noise)
    rl, _, rdz, rg = _oracle(P, target, z, mask, t, noise)
    assert abs(float(loss) - float(rl)) <= FP32_TOL * abs(float(rl))
    _check(dz, rdz, 0, "dz")
    for k in names:
        _check(grads[k], rg[k], 0, k)


@pytest.mark.parametrize("mode", [0, 1], ids=["fp32", "bf16"])
def test_full_size_properties(video, mode):
    """BASELINE size (PushT train, 16384 rows per GPU): size-independent properties.
    (a) gradients are linear in dloss; (b) the backward ACCUMULATES into the gradient tensors;
    (c) rows with mask = 0 contribute nothing: zeroing their inputs' influence leaves loss unchanged;
    (d) the loss of a row permutation is the same."""
    dev, _, P = video
    N = 16384 if mode == 1 else 2048
    gen = torch.Generator().manual_seed(21)
    target = torch.randn(N, 16, generator=gen) * 0.5
    z = torch.randn(N, 768, generator=gen)
    mask = (torch.rand(N, generator=gen) > 0.25).float()
    t = torch.randint(0, 1000, (N,), generator=gen)
    noise = torch.randn(N, 16, generator=gen)
    h, names = _head(16, 768, 1024, 6, P, mode, dev)
    l1, rows1, dz1, g1 = _run(h, names, P, dev, target, z, mask, t, noise, dloss=1.0)
    l2, _, dz2, g2 = _run(h, names, P, dev, target, z, mask, t, noise, dloss=2.0)
    assert torch.isfinite(dz1).all() and float(l1) == float(l2)
    rel = 1e-4 if mode == 0 else 2e-2  # split-K atomics reorder fp32 sums; bf16 rounds dnet-scaled operands
    assert float((dz2 - 2 * dz1).norm() / (2 * dz1).norm()) <= rel
    for k in names:
        assert float((g2[k] - 2 * g1[k]).norm() / (2 * g1[k]).norm().clamp_min(1e-20)) <= rel, k
    # masked-out rows: dz is exactly zero there, and the loss is the masked mean of the row losses
    assert float(dz1[mask == 0].abs().max()) == 0.0
    ref = float((rows1.double() * mask.double()).sum() / mask.double().sum())
    assert abs(float(l1) - ref) <= 1e-5 * abs(ref)
    # permutation invariance of the loss
    perm = torch.randperm(N, generator=gen)
    l3, _, _, _ = _run(h, names, P, dev, target[perm], z[perm], mask[perm], t[perm], noise[perm])
    assert abs(float(l3) - float(l1)) <= 1e-5 * abs(float(l1))
    # accumulate semantics
    from unified_video_action_b200 import _lib  # noqa: F401

    grads = [torch.zeros(P[n].shape, dtype=torch.float32, device=dev) for n in names]
    args = (target.to(dev), z.to(dev), mask.to(dev), t.to(dev), noise.to(dev))
    for _ in range(2):
        h.loss_fwd(*args)
        h.loss_bwd(torch.tensor([1.0], device=dev), *args, grads)
    torch.cuda.synchronize()
    for n, gacc in zip(names, grads):
        assert float((gacc.cpu() - 2 * g1[n]).norm() / (2 * g1[n]).norm().clamp_min(1e-20)) <= rel, n


def test_wgrad_side_stream_equals_in_order(video, monkeypatch):
    """loss_bwd with more than 1024 rows runs its W x W weight-gradient GEMMs on the handle's side stream beside the
    data-gradient chain (UVA_WGRAD_STREAM, default on): same kernels on the same operands, so every gradient equals the
    in-order run up to the order of the split-K atomics, call after call (a buffer rewritten too early would show up)."""
    dev, _, P = video
    N = 3000
    gen = torch.Generator().manual_seed(5)
    target = torch.randn(N, 16, generator=gen) * 0.5
    z = torch.randn(N, 768, generator=gen)
    mask = (torch.rand(N, generator=gen) > 0.25).float()
    t = torch.randint(0, 1000, (N,), generator=gen)
    noise = torch.randn(N, 16, generator=gen)
    res = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("UVA_WGRAD_STREAM", flag)
        h, names = _head(16, 768, 1024, 6, P, 1, dev)
        runs = [_run(h, names, P, dev, target, z, mask, t, noise) for _ in range(3)]
        for l, _, dz, g in runs[1:]:
            assert float(l) == float(runs[0][0]) and torch.equal(dz, runs[0][2])
            for k in names:
                assert float((g[k] - runs[0][3][k]).norm() / runs[0][3][k].norm().clamp_min(1e-20)) <= 1e-5, (flag, k)
        res[flag] = runs[0]
    assert float(res["0"][0]) == float(res["1"][0]) and torch.equal(res["0"][2], res["1"][2])
    for k in names:
        assert float((res["1"][3][k] - res["0"][3][k]).norm() / res["0"][3][k].norm().clamp_min(1e-20)) <= 1e-5, k


@pytest.mark.parametrize("switch,tol", [("UVA_FUSE_COLSUM", 5e-3), ("UVA_EPI_COLSUM", 5e-3), ("UVA_FINAL_TC", BF16_GRAD_L2),
                                        ("UVA_FUSE_GATE_LN", 1e-5)])
def test_backward_switches_off_equal_default(video, monkeypatch, switch, tol):
    """The fused forms of the backward keep their unfused fallbacks behind switches: bias
    gradients summed inside k_gate_bwd / k_lnmod_bwd_split vs separate k_colsum passes (UVA_FUSE_COLSUM), inside the SiLU'
    GEMM epilogue vs a pass (UVA_EPI_COLSUM), the gated-residual backward inside the LayerNorm-modulate backward of the
    block above vs its own kernel (UVA_FUSE_GATE_LN: same values, only the order of two bias sums differs), final-layer
    backward on tcgen05 vs CUDA cores (UVA_FINAL_TC: the tcgen05 form
    rounds d loss / d net to bf16, hence the bf16 gradient tolerance).  Same gradients either way; the fused bias sums
    add the fp32 values BEFORE they are rounded to the bf16 operand the separate pass reads, so the two differ by that
    rounding (measured 1.4e-3 on mlp.2.bias at 2500 rows; bound 5e-3)."""
    dev, _, P = video
    N = 2500
    gen = torch.Generator().manual_seed(9)
    target = torch.randn(N, 16, generator=gen) * 0.5
    z = torch.randn(N, 768, generator=gen)
    mask = (torch.rand(N, generator=gen) > 0.25).float()
    t = torch.randint(0, 1000, (N,), generator=gen)
    noise = torch.randn(N, 16, generator=gen)
    res = {}
    for flag in ("1", "0"):
        monkeypatch.setenv(switch, flag)
        h, names = _head(16, 768, 1024, 6, P, 1, dev)
        res[flag] = _run(h, names, P, dev, target, z, mask, t, noise)
    assert float(res["0"][0]) == float(res["1"][0])
    assert float((res["0"][2] - res["1"][2]).norm() / res["1"][2].norm()) <= tol
    for k in names:
        a, b = res["0"][3][k], res["1"][3][k]
        assert float((a - b).norm() / b.norm().clamp_min(1e-20)) <= tol, (switch, k)


def test_module_forward_backward_matches_oracle(video):
    """The drop-in nn.Module: DiffLoss.forward under autograd with the reference's RNG call order
    (randint for t, then randn_like for the noise; DL:51, GD:761)."""
    dev, _, P = video
    from unified_video_action_b200 import DiffLoss

    head = DiffLoss(16, 768, 6, 1024, "100", n_frames=4, language_emb_model=None, language_emb_model_type=None,
                    precision="fp32")
    head.net.load_state_dict({k: _T(v) for k, v in P.items()})
    head = head.to(dev)
    gen = torch.Generator().manual_seed(3)
    B, S = 3, 50
    target = torch.randn(B, S, 16, generator=gen).to(dev)
    z = torch.randn(B, S, 768, generator=gen).to(dev).requires_grad_(True)
    mask = (torch.rand(B, S, generator=gen) > 0.3).float().to(dev)
    torch.manual_seed(77)
    loss = head(target, z, mask)
    loss.backward()
    torch.manual_seed(77)
    t = torch.randint(0, 1000, (B * S,), device=dev)
    noise = torch.randn(B * S, 16, device=dev)
    rl, _, rdz, rg = _oracle(P, target.reshape(-1, 16).cpu(), z.detach().reshape(-1, 768).cpu(), mask.reshape(-1).cpu(),
                             t.cpu(), noise.cpu())
    assert loss.dim() == 0
    assert abs(float(loss.detach()) - float(rl)) <= FP32_TOL * abs(float(rl))
    _check(z.grad.reshape(-1, 768).cpu(), rdz, 0, "dz")
    for k, p in head.net.named_parameters():
        assert p.grad is not None, k
        _check(p.grad.cpu(), rg[k], 0, k)
    # the fused backward hands out views of one flat buffer; autograd keeps them as .grad, which is what
    # lets allreduce_head_grads reduce the head's gradients in place with one collective
    from unified_video_action_b200.patch import _shared_flat

    flat = _shared_flat([p.grad for p in head.net.parameters()])
    assert flat is not None and flat.numel() >= sum(p.numel() for p in head.net.parameters())
    # a second backward accumulates into the same tensors (no aliasing surprises)
    g0 = head.net.cond_embed.weight.grad.clone()
    torch.manual_seed(77)
    head(target, z, mask).backward()
    assert float((head.net.cond_embed.weight.grad - 2 * g0).abs().max()) <= 1e-5 * float(g0.abs().max())


def test_action_head_side_stream_equals_serial(video, monkeypatch):
    """The joint step of MAR.forward_loss (MAR:770-784): video loss + action loss, one backward.  The action
    head (pre-net, fused loss and, through autograd, their backward) runs on a side stream forked from the
    entry of the video head; losses, dz and every gradient must equal the single-stream run (same RNG
    draws: the generator advances in host call order, not in stream order)."""
    dev, _, P = video
    import copy

    from unified_video_action_b200 import DiffActLoss, DiffLoss

    kw = dict(n_frames=4, language_emb_model=None, language_emb_model_type=None, precision="fp32")
    vid = DiffLoss(16, 768, 6, 1024, "100", **kw)
    vid.net.load_state_dict({k: _T(v) for k, v in P.items()})
    act = DiffActLoss(2, 768, 6, 1024, "100", **kw)
    g0 = torch.Generator().manual_seed(11)
    with torch.no_grad():
        for prm in act.net.parameters():  # un-zero the zero-initialised tensors (DL:250-259)
            if float(prm.abs().max()) == 0.0:
                prm.copy_(0.02 * torch.randn(prm.shape, generator=g0))
    vid, act = vid.to(dev), act.to(dev)
    keys = set(act.state_dict().keys())
    act.overlap_with(vid)
    assert set(act.state_dict().keys()) == keys and not any(m is vid for m in act.modules())
    assert isinstance(copy.deepcopy(act), DiffActLoss)
    B = 2
    gen = torch.Generator().manual_seed(5)
    target = torch.randn(B, 1024, 16, generator=gen).to(dev)
    mask = (torch.rand(B, 1024, generator=gen) > 0.2).float().to(dev)
    actions = (torch.rand(B, 16, 2, generator=gen) * 2 - 1).to(dev)
    z0 = torch.randn(B, 1024, 768, generator=gen).to(dev)

    def run(flag):
        monkeypatch.setenv("UVA_ACT_OVERLAP", flag)
        for m in (vid, act):
            for prm in m.parameters():
                prm.grad = None
        z = z0.clone().requires_grad_(True)
        torch.manual_seed(99)
        lv = vid(target, z, mask)
        la = act(actions, z)
        (lv + la).backward()
        torch.cuda.synchronize()
        return (float(lv), float(la), z.grad.clone(),
                {k: prm.grad.clone() for m, tag in ((vid, "v."), (act, "a.")) for k, prm in ((tag + n, q) for n, q in m.named_parameters())})

    lv0, la0, dz0, g_ser = run("0")
    vid.net.eager_backward = True  # opt-in: the big head's backward is enqueued with its forward (d loss = 1), scaled later
    lvn, lan, dzn, g_non = run("0")
    assert lvn == lv0 and lan == la0
    _check(dzn.cpu(), dz0.cpu(), 0, "dz eager vs autograd-time backward")
    for k in g_ser:
        _check(g_non[k].cpu(), g_ser[k].cpu(), 0, k)
    # d loss != 1 (loss scaling): the eager gradients are scaled afterwards
    for prm in vid.parameters():
        prm.grad = None
    zz = z0.clone().requires_grad_(True)
    torch.manual_seed(99)
    (vid(target, zz, mask) * 3.0).backward()
    _check(vid.net.cond_embed.weight.grad.cpu(), 3.0 * g_ser["v.net.cond_embed.weight"].cpu(), 0, "scaled eager grads")
    with torch.no_grad():  # no graph, no eager work
        assert vid(target, z0, mask).requires_grad is False
    vid.net.eager_backward = False
    for _ in range(3):  # repeated: a missing stream dependency shows up as a flaky mismatch
        lv1, la1, dz1, g_par = run("1")
        assert lv1 == lv0 and la1 == la0
        _check(dz1.cpu(), dz0.cpu(), 0, "dz")
        assert g_par.keys() == g_ser.keys()
        for k in g_ser:
            _check(g_par[k].cpu(), g_ser[k].cpu(), 0, k)


def test_error_paths(video):
    dev, _, P = video
    from unified_video_action_b200 import _lib

    h, names = _head(16, 768, 1024, 6, P, 1, dev)
    x = torch.zeros(4, 16, device=dev)
    z = torch.zeros(4, 768, device=dev)
    t = torch.zeros(4, dtype=torch.int64, device=dev)
    ws = torch.empty(1024, dtype=torch.uint8, device=dev)
    loss = torch.empty(1, device=dev)
    rc = _lib.lib().uva_head_loss_fwd(h._h, x.data_ptr(), z.data_ptr(), None, t.data_ptr(), x.data_ptr(), 4,
                                      loss.data_ptr(), None, ws.data_ptr(), ws.numel(), None)
    assert rc == -4 and b"workspace" in _lib.lib().uva_last_error()
    rc = _lib.lib().uva_head_loss_fwd(h._h, None, z.data_ptr(), None, t.data_ptr(), x.data_ptr(), 4,
                                      loss.data_ptr(), None, ws.data_ptr(), ws.numel(), None)
    assert rc == -1


@pytest.mark.parametrize("A", [6, 9, 10])
@pytest.mark.parametrize("mode", [0, 1], ids=["fp32", "bf16"])
def test_other_channel_counts_vs_reference_golden(golden_dir, A, mode):
    """Training rows, d z and parameter gradients of the act_dim = 10 and the 6 / 9 channel heads (MAR:298, 317,
    332) against the reference's training_losses + autograd (tests/golden/head_dims.npz); 24 rows, one with t = 0."""
    dev = _cuda()
    g = np.load(os.path.join(golden_dir, "head_dims.npz"))
    p = f"a{A}."
    P = O.make_params(A, 768, 1024, 6, seed=20 + A)
    h, names = _head(A, 768, 1024, 6, P, mode, dev)
    loss, rows, dz, grads = _run(h, names, P, dev, _T(g[p + "tr_target"]), _T(g[p + "tr_z"]), None, _T(g[p + "tr_t"]), _T(g[p + "tr_noise"]))
    rt = 2e-4 if mode == 0 else 2e-2
    np.testing.assert_allclose(rows.numpy(), g[p + "tr_rowloss"], rtol=rt, atol=rt * float(np.abs(g[p + "tr_rowloss"]).max()))
    assert abs(float(loss) - float(g[p + "tr_loss"])) <= (FP32_TOL if mode == 0 else BF16_LOSS_TOL) * abs(float(g[p + "tr_loss"]))
    _check(dz, _T(g[p + "tr_dz"]), mode, "dz")
    for k in names:
        dg = g[p + "tr_gdigest." + k]
        f = grads[k].double().flatten()
        tol = 2e-4 if mode == 0 else BF16_GRAD_L2
        assert abs(float(f.norm()) - dg[2]) <= tol * max(dg[2], 1e-12), (k, float(f.norm()), dg[2])
        samp = f[_T(g[p + "tr_gidx." + k])]
        ref = _T(g[p + "tr_gsample." + k])
        if float(ref.abs().max()) > 0:
            err = float((samp - ref).norm() / ref.norm())
            assert err <= (1e-3 if mode == 0 else 4 * BF16_GRAD_L2), (k, err)
