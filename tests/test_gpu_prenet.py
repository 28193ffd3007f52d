"""the DiffActLoss "conv_fc" pre-net fused into the library (implicit-GEMM convolution on the CTA-pair
tcgen05 kernel with bias + ReLU + 4x4 average pool in the epilogue, tcgen05 Linears, hand-written backward).
Checked against the UNMODIFIED reference's outputs / gradients (tests/golden/prenet.npz) and the oracle restatement;
bf16 operands with fp32 accumulation (what the reference trains with under bf16 autocast): output <= 1e-2 relative L2
against the fp32 reference, <= 3e-3 against the oracle with bf16 rounding at the same points.  Gradients: <= 2e-2 against
the oracle's autograd through the SAME bf16-rounded forward (same ReLU masks); against the fp32 reference only the
gradient norms are held to 2e-2 and the entries to 1e-1, because the pre-net has two ReLUs right behind GEMMs: an output
within 3e-3 of zero (0.4 % of the conv outputs here) can land on the other side of the ReLU in bf16 arithmetic, and
every such mask flip is a full-size error in the entries behind it (sqrt(0.004) = 6 % in L2) - a property of bf16
training of this pre-net, the reference under autocast included, not of the kernels.  The same mechanism, much
weaker, is why the emulation bound is 3e-2 and not 1e-3: the convolution agrees with the oracle to 6e-6 (pooled
maps, scripts/prenet_diag.py), but a value that sits on a bf16 rounding boundary rounds the other way (2e-4 after
fc.0, 1.5e-3 at the output), which flips a few ReLU masks of the later layers."""
import os

import numpy as np
import pytest
import torch

from oracle import uva_oracle as O
from test_oracle_golden import PRENET_NAMES, prenet_case

pytestmark = pytest.mark.gpu
OUT_TOL, EMU_TOL, GRAD_TOL, GRAD_TOL_FP32_MASKS, BWD_TOL = 1e-2, 3e-3, 3e-2, 1e-1, 1e-2


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)


def _l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


def test_fused_prenet_forward_backward_vs_reference_golden(golden_dir, record_metric):
    dev = _cuda()
    g, mod, z, cot = prenet_case(golden_dir)
    Pp = {n: p.detach().clone() for n, p in mod.named_parameters() if n in PRENET_NAMES}
    mod = mod.to(dev)
    zz = z.to(dev).requires_grad_(True)
    out = mod.prenet(zz)
    assert out.shape == (z.shape[0], 16, 768) and out.dtype == torch.float32
    ref = torch.from_numpy(g["out"])
    e = _l2(out.detach().cpu(), ref)
    emu = O.prenet_conv_fc(Pp, z, emulate_bf16=True)
    e2 = _l2(out.detach().cpu(), emu)
    record_metric("prenet", "out_l2_vs_fp32_reference", e, OUT_TOL)
    record_metric("prenet", "out_l2_vs_bf16_emulation", e2, EMU_TOL)
    assert e <= OUT_TOL and e2 <= EMU_TOL, (e, e2)
    (out * cot.to(dev)).sum().backward()
    dz = zz.grad.cpu()
    dg = g["dz_digest"]
    assert abs(float(dz.double().norm()) - dg[2]) <= GRAD_TOL * dg[2]
    idx = torch.from_numpy(g["dz_idx"])
    e = _l2(dz.flatten()[idx], torch.from_numpy(g["dz_val"]))
    record_metric("prenet", "dz_l2_vs_fp32_reference", e, GRAD_TOL_FP32_MASKS)
    assert e <= GRAD_TOL_FP32_MASKS, e
    # the same gradients against the oracle's autograd through the bf16-rounded forward (same ReLU masks): tight
    Pe = {n: p.clone().requires_grad_(True) for n, p in Pp.items()}
    ze = z.clone().requires_grad_(True)
    (O.prenet_conv_fc(Pe, ze, emulate_bf16=True) * cot).sum().backward()
    e = _l2(dz, ze.grad)
    record_metric("prenet", "dz_l2_vs_bf16_emulation", e, GRAD_TOL)
    assert e <= GRAD_TOL, e
    worst = ("", 0.0)
    for n, p in mod.named_parameters():
        if n not in PRENET_NAMES:
            continue
        assert p.grad is not None, n
        gd = g["gdigest." + n]
        f = p.grad.detach().cpu().double().flatten()
        assert abs(float(f.norm()) - gd[2]) <= GRAD_TOL * gd[2], (n, float(f.norm()), gd[2])
        e = _l2(f[torch.from_numpy(g["gidx." + n])], torch.from_numpy(g["gval." + n]))
        record_metric("prenet", "grad_l2_vs_fp32_reference." + n, e, GRAD_TOL_FP32_MASKS)
        assert e <= GRAD_TOL_FP32_MASKS, (n, e)
        e = _l2(p.grad.detach().cpu(), Pe[n].grad)
        record_metric("prenet", "grad_l2_vs_bf16_emulation." + n, e, GRAD_TOL)
        if e > worst[1]:
            worst = (n, e)
    assert worst[1] <= GRAD_TOL, worst


def _saved_masks(head, B, Z=768, T=4, J=16):
    """ReLU masks the library's forward kept for its backward (training workspace layout of carve_prenet in
    csrc/head_prenet.inc: zb, pooled, pooled_b, acc1 [24][128][Z], h1, f2, itp, r1, y, ...)."""
    ws = head._ws[2]
    BT, BJ, PX = B * T, B * J, B * T * 256
    off = [0]

    def take(n):
        r = off[0]
        off[0] += (n + 1023) // 1024 * 1024
        return r

    take(PX * Z * 2); take(BT * 16 * Z * 4); take(BT * 16 * Z * 2); take(24 * 128 * Z * 4)
    o_h1 = take(BT * Z * 2); take(BT * Z * 4); take(BJ * Z * 2)
    o_r1 = take(BJ * Z * 2)
    o_y = take(PX * Z * 2)
    view = lambda o, n, shape: ws[o:o + n * 2].view(torch.bfloat16).reshape(shape).float().cpu()
    return view(o_y, PX * Z, (BT, 16, 16, Z)) > 0, view(o_h1, BT * Z, (BT, Z)) > 0, view(o_r1, BJ * Z, (B, J, Z)) > 0


def _masked_linear_prenet(Pp, z, my, mh, mr):
    """The pre-net with its three ReLUs replaced by GIVEN masks and bf16-rounded GEMM operands: linear in z, so its
    autograd gives the exact gradients for those masks - what a correct backward must reproduce."""
    import torch.nn.functional as F

    r = O._bf16r
    B, S, Z = z.shape
    x = z.reshape(B * 4, 16, 16, Z).permute(0, 3, 1, 2)
    x = F.conv2d(r(x), r(Pp["conv.0.weight"]), Pp["conv.0.bias"], padding=1) * my.permute(0, 3, 1, 2)
    x = F.adaptive_avg_pool2d(x, (4, 4)).reshape(B * 4, -1)
    x = F.linear(r(x), r(Pp["fc.0.weight"]), Pp["fc.0.bias"]) * mh
    x = F.linear(r(x), r(Pp["fc.2.weight"]), Pp["fc.2.bias"])
    x = x.reshape(B, 4, Z).permute(0, 2, 1)
    x = F.linear(x, Pp["interpolate.weight"], Pp["interpolate.bias"]).permute(0, 2, 1)
    x = F.linear(r(x), r(Pp["refine.0.weight"]), Pp["refine.0.bias"]) * mr
    return F.linear(r(x), r(Pp["refine.2.weight"]), Pp["refine.2.bias"])


@pytest.mark.parametrize("B", [1, 3, 16])
def test_fused_prenet_vs_oracle_other_batches(B, record_metric):
    """predict_action (B = 1), a ragged batch, the training batch of BASELINE config 2 (B = 16).  Forward against the
    oracle (bf16 rounding at the same points: 3e-3; plain fp32: 1e-2).  Backward (d z and all 12 parameter gradients)
    against the autograd of the oracle evaluated with the library's OWN ReLU masks: with the masks fixed the pre-net is
    a linear map, so this isolates the backward kernels from the mask flips described in the module docstring - held
    to 1e-2 (bf16 rounding of the backward operands).  Gradients accumulate."""
    dev = _cuda()
    from unified_video_action_b200 import DiffActLoss

    torch.manual_seed(20 + B)
    mod = DiffActLoss(2, 768, 6, 1024, "100", n_frames=4, language_emb_model=None, language_emb_model_type=None)
    Pp = {n: p.detach().clone().requires_grad_(True) for n, p in mod.named_parameters() if n in PRENET_NAMES}
    gen = torch.Generator().manual_seed(B)
    z = torch.randn(B, 1024, 768, generator=gen)
    cot = torch.randn(B, 16, 768, generator=gen)
    mod = mod.to(dev)
    with torch.no_grad():
        o_inf = mod.prenet(z.to(dev))  # inference call (no activations kept)
        ref_emu, ref_f32 = O.prenet_conv_fc(Pp, z, emulate_bf16=True), O.prenet_conv_fc(Pp, z)
    zz = z.to(dev).requires_grad_(True)
    out = mod.prenet(zz)
    assert torch.equal(out.detach(), o_inf)  # the forward is bit-reproducible (fc.0 split-K planes are reduced in a fixed order)
    e = _l2(out.detach().cpu(), ref_emu)
    record_metric("prenet", f"B{B}_out_l2_vs_bf16_emulation", e, EMU_TOL)
    assert e <= EMU_TOL, e
    e = _l2(out.detach().cpu(), ref_f32)
    record_metric("prenet", f"B{B}_out_l2_vs_fp32", e, OUT_TOL)
    assert e <= OUT_TOL, e
    my, mh, mr = _saved_masks(mod.net.head(), B)
    (out * cot.to(dev)).sum().backward()
    zr = z.clone().requires_grad_(True)
    lin = _masked_linear_prenet(Pp, zr, my, mh, mr)
    assert _l2(lin.detach(), out.detach().cpu()) <= EMU_TOL  # same masks; what is left are values on bf16 rounding boundaries
    (lin * cot).sum().backward()
    e = _l2(zz.grad.cpu(), zr.grad)
    record_metric("prenet", f"B{B}_dz_l2_same_masks", e, BWD_TOL)
    assert e <= BWD_TOL, e
    for n, p in mod.named_parameters():
        if n in PRENET_NAMES:
            e = _l2(p.grad.cpu(), Pp[n].grad)
            record_metric("prenet", f"B{B}_grad_l2_same_masks.{n}", e, BWD_TOL)
            assert e <= BWD_TOL, (n, e)
    g1 = mod.conv[0].weight.grad.clone()
    out2 = mod.prenet(zz)
    with torch.no_grad():
        mod.prenet(torch.randn_like(zz))  # an inference call between a training forward and its backward shares the buffers
    (out2 * cot.to(dev)).sum().backward()
    assert _l2(mod.conv[0].weight.grad, 2 * g1) <= 1e-3  # accumulate semantics (split-K atomics reorder the sums)


def test_fused_prenet_inside_the_module_calls(monkeypatch):
    """DiffActLoss.forward / .sample in bf16 mode use the fused pre-net; results track the PyTorch pre-net path
    (UVA_FUSED_PRENET=0) within the bf16 tolerance, and an optimizer step re-packs the pre-net operands."""
    dev = _cuda()
    from unified_video_action_b200 import DiffActLoss, HeadAdamW

    torch.manual_seed(3)
    act = DiffActLoss(2, 768, 6, 1024, "100", n_frames=4, language_emb_model=None, language_emb_model_type=None)
    g0 = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for p in act.net.parameters():
            if float(p.abs().max()) == 0.0:
                p.copy_(0.02 * torch.randn(p.shape, generator=g0))
    act = act.to(dev)
    z = torch.randn(2, 1024, 768, generator=g0).to(dev).requires_grad_(True)
    target = (torch.rand(2, 16, 2, generator=g0) * 2 - 1).to(dev)

    def run(flag):
        monkeypatch.setenv("UVA_FUSED_PRENET", flag)
        for p in act.parameters():
            p.grad = None
        z.grad = None
        torch.manual_seed(9)
        loss = act(target, z)
        loss.backward()
        torch.manual_seed(9)
        s = act.sample(z.detach(), 0.95, 1.0)
        return float(loss.detach()), z.grad.clone(), act.fc[0].weight.grad.clone(), s

    l1, dz1, gw1, s1 = run("1")
    l0, dz0, gw0, s0 = run("0")
    assert abs(l1 - l0) <= 1e-2 * abs(l0)
    assert _l2(dz1, dz0) <= GRAD_TOL_FP32_MASKS and _l2(gw1, gw0) <= GRAD_TOL_FP32_MASKS  # fp32 PyTorch pre-net: other ReLU masks
    assert _l2(s1, s0) <= 3e-2
    monkeypatch.setenv("UVA_FUSED_PRENET", "1")
    opt = HeadAdamW([act], lr=1e-2, weight_decay=0.0)
    before = act.prenet(z.detach()).clone()
    torch.manual_seed(9)
    act(target, z).backward()
    opt.step()
    after = act.prenet(z.detach())
    assert _l2(after, before) > 1e-3  # the packed conv / Linear operands follow the in-place parameter update
