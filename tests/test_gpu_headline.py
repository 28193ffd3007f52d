"""Parity of the BENCHMARKED code path against the oracle, directly ("headline-size parity is
transitive"): above 2048 rows the bf16 sampler takes the CTA-pair tcgen05 kernels with the fused epilogues
(AdaLN GEMM + LayerNorm-modulate, GEMM1 + SiLU, GEMM2 + gated residual, final linear + CFG mix + DDPM update) and,
from two chunks on, two lanes.  These tests compare exactly that path with the CPU oracle

* at 4096 rows (CFG 3.0) and 6000 rows (ragged, no CFG): 3 free-running and 3 teacher-forced steps against the
  oracle with bf16 rounding at the same points (<= 6e-3 max-relative) and against the plain fp32 oracle (<= 1e-2
  relative L2, the bf16 tolerance of BASELINE.json);
* at the bench size, 131 072 rows (cond || uncond, CFG 3.0, two lanes): 2 steps, a strided 1024-row subset;
* training at the C2 size (16 384 rows + mask) against the oracle's autograd: loss, dz, all 60 gradients;
* a 30-step AdamW loss curve (HeadAdamW, bf16 mode) against the oracle trained with torch.optim.AdamW on the same
  data and draws ("matching loss curves", GD:746-818).
"""
import numpy as np
import pytest
import torch

from oracle import uva_oracle as O

pytestmark = pytest.mark.gpu

BF16_TOL = 1e-2       # relative L2 per step vs the fp32 oracle (BASELINE.json: bf16 mode <= 1e-2 relative)
BF16_EMU_TOL = 6e-3   # max|err| / max|ref| per step vs the oracle with bf16 rounding at the same points
BF16_LOSS_TOL = 1e-2
BF16_GRAD_L2 = 2e-2   # per-tensor relative L2 of bf16-mode gradients vs the fp32 oracle autograd


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)


def _T(a, dev=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    return t.to(dev) if dev is not None else t


def _head(C, P, mode, dev):
    from unified_video_action_b200 import _lib

    names = [n for n, _ in O.param_specs(C, 768, 1024, 6)]
    with torch.cuda.device(dev):
        h = _lib.Head(C, 768, 1024, 6, 1000, 100, mode)
        h.load_weights([_T(P[n], dev) for n in names])
    return h, names


def _max_rel(mine, ref):
    T = ref.shape[0]
    err = (mine - ref).abs().reshape(T, -1).max(dim=1).values
    return float((err / ref.abs().reshape(T, -1).max(dim=1).values.clamp_min(1e-6)).max())


def _l2_rel(mine, ref):
    T = ref.shape[0]
    return float(((mine - ref).reshape(T, -1).norm(dim=1) / ref.reshape(T, -1).norm(dim=1).clamp_min(1e-6)).max())


@pytest.fixture(scope="module")
def P16():
    return O.make_params(16, 768, 1024, 6, seed=1)


@pytest.mark.parametrize("N,cfg", [(4096, 3.0), (6000, 1.0)])
def test_pair_kernel_path_vs_oracle(P16, record_metric, N, cfg):
    """One chunk above the latency-path limit: every GEMM of the stack runs on the CTA-pair kernel with its fused
    epilogue, the final linear carries the CFG mix and the DDPM update."""
    dev = _cuda()
    h, _ = _head(16, P16, 1, dev)
    S = O.create_schedule("100")
    gen = torch.Generator().manual_seed(N)
    steps = 3
    z = torch.randn(N, 768, generator=gen)
    if cfg != 1.0:
        half = torch.randn(N // 2, 16, generator=gen)
        x0 = torch.cat([half, half])
    else:
        x0 = torch.randn(N, 16, generator=gen)
    nz = torch.randn(steps, N, 16, generator=gen)
    Pt = O.to_torch_params(P16)
    kw = dict(temperature=0.95, cfg=cfg, clip_denoised=False)
    _, emu = O.sample_loop(Pt, S, z, x0, nz, emulate_bf16=True, trace=True, num_steps=steps, **kw)
    _, f32 = O.sample_loop(Pt, S, z, x0, nz, trace=True, num_steps=steps, **kw)
    args = (z.to(dev), x0.to(dev), nz.to(dev))
    out, tr = h.sample(*args, trace=True, k_begin=0, k_end=steps, **kw)  # a trace always takes the chunked path
    tr = tr.cpu()
    assert torch.equal(out.cpu(), tr[-1])
    e_emu, e_f32 = _max_rel(tr, emu), _l2_rel(tr, f32)
    record_metric("pair_path_free", f"N{N}_cfg{cfg}_emu_max", e_emu, BF16_EMU_TOL)
    record_metric("pair_path_free", f"N{N}_cfg{cfg}_fp32_l2", e_f32, BF16_TOL)
    assert e_emu <= BF16_EMU_TOL, e_emu
    assert e_f32 <= BF16_TOL, e_f32
    # teacher-forced per step: every step starts from the oracle's own state
    for ref, name, tol, fn in ((emu, "emu_max", BF16_EMU_TOL, _max_rel), (f32, "fp32_l2", BF16_TOL, _l2_rel)):
        teacher = torch.cat([x0[None], ref[:-1]], dim=0)
        _, tt = h.sample(*args, teacher=teacher.to(dev), trace=True, k_begin=0, k_end=steps, **kw)
        e = fn(tt.cpu(), ref)
        record_metric("pair_path_teacher", f"N{N}_cfg{cfg}_{name}", e, tol)
        assert e <= tol, (name, e)


def test_bench_size_cfg_two_lanes_subset_vs_oracle(P16, record_metric):
    """The exact BENCH configuration: 131 072 rows (65 536 cond || 65 536 uncond), CFG 3.0, temperature 0.95, two
    chunk lanes; 2 denoise steps; 512 strided cond / uncond row pairs against the oracle."""
    dev = _cuda()
    h, _ = _head(16, P16, 1, dev)
    S = O.create_schedule("100")
    N, half, steps = 131072, 65536, 2
    g = torch.Generator(device=dev).manual_seed(77)
    z = torch.randn(N, 768, generator=g, device=dev)
    xh = torch.randn(half, 16, generator=g, device=dev)
    x0 = torch.cat([xh, xh])
    nz = torch.randn(steps, N, 16, generator=g, device=dev)
    kw = dict(temperature=0.95, cfg=3.0, clip_denoised=False)
    out = h.sample(z, x0, nz, k_begin=0, k_end=steps, **kw)
    assert torch.isfinite(out).all()
    sel = torch.arange(37, half, 128, device=dev)[:512]  # pairs from both lanes (chunks of 32 768 pairs each)
    rows = torch.cat([sel, sel + half])
    zs, xs, ns = z[rows].cpu(), x0[rows].cpu(), nz[:, rows].cpu()
    Pt = O.to_torch_params(P16)
    emu = O.sample_loop(Pt, S, zs, xs, ns, emulate_bf16=True, num_steps=steps, **kw)
    f32 = O.sample_loop(Pt, S, zs, xs, ns, num_steps=steps, **kw)
    mine = out[rows].cpu()
    e_emu, e_f32 = _max_rel(mine[None], emu[None]), _l2_rel(mine[None], f32[None])
    record_metric("bench_size_subset", "emu_max", e_emu, BF16_EMU_TOL)
    record_metric("bench_size_subset", "fp32_l2", e_f32, BF16_TOL)
    assert e_emu <= BF16_EMU_TOL, e_emu
    assert e_f32 <= BF16_TOL, e_f32


def _oracle_grads(P_np, target, z, mask, t, noise):
    P = O.to_torch_params(P_np, requires_grad=True)
    S = O.create_schedule("")
    zz = z.clone().requires_grad_(True)
    loss = O.diffloss_forward(P, S, target, zz, mask, t, noise)
    loss.backward()
    return loss.detach(), zz.grad, {k: p.grad for k, p in P.items()}


def test_training_at_c2_size_vs_oracle_autograd(P16, record_metric):
    """BASELINE config 2 per-GPU size: 16 384 rows with a mask (zeros included), t = 0 and t = 999 rows present;
    bf16 mode (the pair kernel's training epilogues, split-K weight gradients) against the oracle's fp32 autograd."""
    dev = _cuda()
    N = 16384
    gen = torch.Generator().manual_seed(16384)
    target = torch.randn(N, 16, generator=gen) * 0.5
    z = torch.randn(N, 768, generator=gen)
    mask = (torch.rand(N, generator=gen) < 0.85).float()
    t = torch.randint(0, 1000, (N,), generator=gen)
    t[:16] = 0
    t[16] = 999
    noise = torch.randn(N, 16, generator=gen)
    h, names = _head(16, P16, 1, dev)
    grads = [torch.zeros(P16[n].shape, dtype=torch.float32, device=dev) for n in names]
    a = (target.to(dev), z.to(dev), mask.to(dev), t.to(dev), noise.to(dev))
    loss = h.loss_fwd(*a)
    dz = h.loss_bwd(torch.tensor([1.0], device=dev), *a, grads)
    torch.cuda.synchronize()
    rl, rdz, rg = _oracle_grads(P16, target, z, mask, t, noise)
    e = abs(float(loss) - float(rl)) / abs(float(rl))
    record_metric("train_c2", "loss_rel", e, BF16_LOSS_TOL)
    assert e <= BF16_LOSS_TOL, e
    worst = ("", 0.0)
    for k, mine, ref in [("dz", dz.cpu(), rdz)] + [(n, g.cpu(), rg[n]) for n, g in zip(names, grads)]:
        err = float((mine.double() - ref.double()).norm() / ref.double().norm().clamp_min(1e-30))
        record_metric("train_c2", "grad_l2." + k, err, BF16_GRAD_L2)
        if err > worst[1]:
            worst = (k, err)
    assert worst[1] <= BF16_GRAD_L2, worst
    assert float(dz[a[2] == 0].abs().max()) == 0.0  # masked-out rows get no gradient (DL:64-66)


def test_loss_curve_matches_oracle_adamw(record_metric):
    """"Matching loss curves" (BASELINE.json north_star): 30 optimizer steps of the drop-in DiffLoss in bf16 mode
    with the fused HeadAdamW against the oracle (fp32 torch autograd) with torch.optim.AdamW, same initial weights,
    same batches, same t / noise draws.  Every loss of the curve within 1e-2 relative."""
    dev = _cuda()
    from unified_video_action_b200 import DiffLoss, HeadAdamW

    P = O.make_params(16, 768, 1024, 6, seed=4)
    head = DiffLoss(16, 768, 6, 1024, "100", n_frames=4, language_emb_model=None, language_emb_model_type=None,
                    precision="bf16")
    head.net.load_state_dict({k: _T(v) for k, v in P.items()})
    head = head.to(dev)
    lr, betas, wd, steps, B, Sq = 2e-4, (0.9, 0.95), 0.02, 30, 2, 192
    opt = HeadAdamW([head], lr=lr, betas=betas, weight_decay=wd)
    Pt = O.to_torch_params(P, requires_grad=True)
    decay = [p for k, p in Pt.items() if p.dim() > 1 and not k.endswith(".bias")]  # add_weight_decay (POL:325-341)
    rest = [p for k, p in Pt.items() if not (p.dim() > 1 and not k.endswith(".bias"))]
    ropt = torch.optim.AdamW([{"params": decay, "weight_decay": wd}, {"params": rest, "weight_decay": 0.0}], lr=lr, betas=betas)
    S = O.create_schedule("")
    gen = torch.Generator().manual_seed(9)
    mine, ref = [], []
    for it in range(steps):
        target = 0.5 * torch.randn(B, Sq, 16, generator=gen)
        z = torch.randn(B, Sq, 768, generator=gen)
        mask = (torch.rand(B, Sq, generator=gen) < 0.85).float()
        torch.manual_seed(1000 + it)
        opt.zero_grad()
        loss = head(target.to(dev), z.to(dev), mask.to(dev))
        loss.backward()
        opt.step()
        mine.append(float(loss.detach()))
        torch.manual_seed(1000 + it)  # the module's own draws (DL:51, GD:761), re-drawn for the oracle
        t = torch.randint(0, 1000, (B * Sq,), device=dev).cpu()
        noise = torch.randn(B * Sq, 16, device=dev).cpu()
        ropt.zero_grad()
        rl = O.diffloss_forward(Pt, S, target.reshape(-1, 16), z.reshape(-1, 768), mask.reshape(-1), t, noise)
        rl.backward()
        ropt.step()
        ref.append(float(rl.detach()))
    mine, ref = np.array(mine), np.array(ref)
    rel = np.abs(mine - ref) / np.abs(ref)
    record_metric("loss_curve", "max_rel", rel.max(), BF16_LOSS_TOL)
    record_metric("loss_curve", "first_loss", ref[0])
    record_metric("loss_curve", "last_loss", ref[-1])
    assert ref[-5:].mean() < 0.9 * ref[:5].mean(), "the curve must actually move for the comparison to mean something"
    assert rel.max() <= BF16_LOSS_TOL, (rel.max(), mine, ref)


def test_two_forwards_before_one_backward():
    """`(l1 + l2).backward()` through ONE head - each fused-loss node differentiates its own forward
    although the head keeps a single activation workspace (the node whose activations were overwritten recomputes
    them from its saved inputs)."""
    dev = _cuda()
    from unified_video_action_b200 import DiffLoss

    P = O.make_params(16, 768, 1024, 6, seed=4)
    head = DiffLoss(16, 768, 6, 1024, "100", n_frames=4, language_emb_model=None, language_emb_model_type=None,
                    precision="fp32")
    head.net.load_state_dict({k: _T(v) for k, v in P.items()})
    head = head.to(dev)
    gen = torch.Generator().manual_seed(2)
    B, Sq = 2, 40
    ta, tb = torch.randn(B, Sq, 16, generator=gen).to(dev), torch.randn(B, 2 * Sq, 16, generator=gen).to(dev)
    za = torch.randn(B, Sq, 768, generator=gen).to(dev).requires_grad_(True)
    zb = torch.randn(B, 2 * Sq, 768, generator=gen).to(dev).requires_grad_(True)
    ma, mb = torch.ones(B, Sq, device=dev), torch.ones(B, 2 * Sq, device=dev)

    def grads_of(fn):
        for p in head.parameters():
            p.grad = None
        za.grad = zb.grad = None
        torch.manual_seed(5)
        fn()
        return za.grad.clone(), zb.grad.clone(), {k: p.grad.clone() for k, p in head.net.named_parameters()}

    def separate():
        head(ta, za, ma).backward()
        head(tb, zb, mb).backward()  # second forward has MORE rows: the workspace is regrown in between

    def joint():
        l1 = head(ta, za, ma)
        l2 = head(tb, zb, mb)
        with torch.no_grad():
            head(ta, za, ma)  # an eval forward between the training forwards and their backward
        (l1 + l2).backward()

    a1, b1, g1 = grads_of(separate)
    a2, b2, g2 = grads_of(joint)
    assert float((a1 - a2).abs().max()) <= 1e-6 * float(a1.abs().max())
    assert float((b1 - b2).abs().max()) <= 1e-6 * float(b1.abs().max())
    for k in g1:
        assert float((g1[k] - g2[k]).abs().max()) <= 1e-5 * float(g1[k].abs().max()), k


def test_c4_multiview_call_pattern_vs_oracle():
    """BASELINE config 4 (Libero10 / multi-view, MAR:1037-1041, 1101-1103, 1120-1122): one sample_tokens iteration
    calls diffactloss.sample (A = 10, clip_denoised) on z [B,1024,768], then diffloss.sample and
    diffloss_wrist.sample on the SAME flattened z rows (cfg 1, shipped configs).  The three modules' outputs against
    the oracle on the reference's RNG call order (fp32 mode, 20-step respacing keeps the free-running chain tame)."""
    dev = _cuda()
    from unified_video_action_b200 import DiffActLoss, DiffLoss

    kw = dict(n_frames=4, language_emb_model=None, language_emb_model_type=None, precision="fp32")
    Pv, Pw, Pa = (O.make_params(16, 768, 1024, 6, seed=s) for s in (1, 7, 30))
    Pa = O.make_params(10, 768, 1024, 6, seed=30)
    vid, wrist = DiffLoss(16, 768, 6, 1024, "20", **kw), DiffLoss(16, 768, 6, 1024, "20", **kw)
    act = DiffActLoss(10, 768, 6, 1024, "20", act_diff_testing_steps="20", **kw)
    for m, P in ((vid, Pv), (wrist, Pw), (act, Pa)):
        m.net.load_state_dict({k: _T(v) for k, v in P.items()})
    vid, wrist, act = vid.to(dev), wrist.to(dev), act.to(dev)
    S = O.create_schedule("20")
    B = 2
    gen = torch.Generator().manual_seed(44)
    z = torch.randn(B, 1024, 768, generator=gen).to(dev)
    torch.manual_seed(123)
    a = act.sample(z, 0.95, cfg=1.0)                      # MAR:1039
    rows = z.reshape(B * 1024, 768)                       # z[mask_to_pred.nonzero()] with num_iter = 1 (MAR:1090)
    v = vid.sample(rows, 0.95, 1.0)                       # MAR:1101
    w = wrist.sample(rows, 0.95, 1.0)                     # MAR:1120
    assert a.shape == (B, 16, 10) and v.shape == (B * 1024, 16) and w.shape == (B * 1024, 16)
    torch.manual_seed(123)
    with torch.no_grad():
        arows = act.prenet(z).reshape(B * 16, 768).cpu()
    x0 = torch.randn(B * 16, 10)
    nz = torch.stack([torch.randn(B * 16, 10, device=dev) for _ in range(20)]).cpu()
    ra = O.sample_loop(O.to_torch_params(Pa), S, arows, x0, nz, temperature=0.95, cfg=1.0, clip_denoised=True)
    assert float((a.reshape(-1, 10).cpu() - ra).abs().max()) <= 2e-3 * max(1.0, float(ra.abs().max()))
    sub = torch.arange(0, B * 1024, 16)  # the oracle follows 128 of the 2048 rows (rows are independent chains)
    for mine, P in ((v, Pv), (w, Pw)):
        x0 = torch.randn(B * 1024, 16)
        nz = torch.stack([torch.randn(B * 1024, 16, device=dev) for _ in range(20)]).cpu()
        ref = O.sample_loop(O.to_torch_params(P), S, rows[sub].cpu(), x0[sub], nz[:, sub], temperature=0.95, cfg=1.0,
                            clip_denoised=False)
        assert float((mine[sub].cpu() - ref).abs().max()) <= 2e-3 * float(ref.abs().max())


@pytest.mark.parametrize("layout,M,N,K", [(24, 300, 512, 1024), (25, 4101, 1024, 1024), (25, 256, 256, 64), (33, 2500, 1024, 1024)])
def test_pair_gemm_bf16_output_epilogues(layout, M, N, K):
    """The TMA-store epilogues of the CTA-pair kernel in isolation: layout 24 = bf16(acc + bias) (EPI_STORE_ACT),
    25 / 33 = bf16(SiLU(acc + bias)) (EPI_SILU_ACT; 33 on clusters of two pairs) against fp32 matmul."""
    dev = _cuda()
    from unified_video_action_b200 import _lib

    g = torch.Generator().manual_seed(layout + M)
    A = torch.randn(M, K, generator=g).to(dev).bfloat16()
    B = (torch.randn(N, K, generator=g) / K ** 0.5).to(dev).bfloat16()
    bias = torch.randn(N, generator=g).to(dev)
    D = _lib.gemm_bf16_selftest(A, B, bias, 256, layout=layout)
    torch.cuda.synchronize()
    mine = D.view(torch.bfloat16).reshape(-1)[: M * N].reshape(M, N).float()
    ref = A.float() @ B.float().t() + bias
    if layout != 24:
        ref = torch.nn.functional.silu(ref)
    # bf16 output rounding (2^-9 relative) + the one-MUFU tanh.approx SiLU (<= 2^-10.5 absolute on tanh)
    err = (mine - ref).abs()
    assert float((err / (ref.abs() + 0.25)).max()) <= 8e-3, float((err / (ref.abs() + 0.25)).max())


@pytest.mark.parametrize("N,cfg,group", [(300, 1.0, 0), (6000, 1.0, 1), (6000, 3.0, 0), (20000, 1.0, 16), (70000, 3.0, 0), (70000, 1.0, 5),
                                         (3000, 1.0, 0), (5200, 2.0, 0)])
def test_layer_chain_kernel_equals_per_layer_launches(P16, monkeypatch, N, cfg, group):
    """The persistent layer-chain kernel (stack_kernel.cuh: all 19 tcgen05 GEMMs of a denoise step in one launch,
    tiles ordered panel-major, layers chained through completion counters in global memory) runs the same MMAs in
    the same order and the same epilogue arithmetic as the per-layer launches (UVA_STACK=0): bit-equal samples, for
    every item order (0: the default diagonal wavefront; n > 0: groups of n tiles, 1 = the strictest dependency pattern) and across repeated calls
    (a missed dependency shows up as a flaky mismatch)."""
    dev = _cuda()
    gen = torch.Generator().manual_seed(N + group)
    steps = 2
    z = torch.randn(N, 768, generator=gen).to(dev)
    if cfg != 1.0:
        half = torch.randn(N // 2, 16, generator=gen)
        x0 = torch.cat([half, half]).to(dev)
    else:
        x0 = torch.randn(N, 16, generator=gen).to(dev)
    nz = torch.randn(steps, N, 16, generator=gen).to(dev)
    kw = dict(temperature=0.95, cfg=cfg, clip_denoised=False, k_begin=0, k_end=steps)
    monkeypatch.setenv("UVA_SMALL_ROWS", "0")
    monkeypatch.setenv("UVA_STACK", "0")
    h_ref, _ = _head(16, P16, 1, dev)
    ref = h_ref.sample(z, x0, nz, **kw)
    l_ref = h_ref.launch_count()
    monkeypatch.setenv("UVA_STACK", "1")
    monkeypatch.setenv("UVA_STACK_ROWS", "1000000")  # also where the default policy prefers the per-layer launches
    monkeypatch.setenv("UVA_STACK_GROUP", str(group))
    h, _ = _head(16, P16, 1, dev)
    for rep in range(4):
        out = h.sample(z, x0, nz, **kw)
        assert torch.equal(out, ref), (rep, float((out - ref).abs().max()))
    assert h.launch_count() / 4 < 0.5 * l_ref  # 4 kernels per step and chunk instead of 22


def test_sample_tokens_glue_matches_the_reference_gather_scatter():
    """patch.sample_masked_tokens (MAR:1086-1116 around diffloss.sample): with every token masked (num_iter = 1, all
    shipped configs) the sync-free form gives exactly what the reference's nonzero-gather / sample / nonzero-scatter
    sequence gives on the same RNG state; with a partial mask it follows the reference's sequence itself."""
    dev = _cuda()
    from unified_video_action_b200 import DiffLoss, sample_masked_tokens

    P = O.make_params(16, 768, 1024, 6, seed=1)
    head = DiffLoss(16, 768, 6, 1024, "20", n_frames=4, language_emb_model=None, language_emb_model_type=None)
    head.net.load_state_dict({k: _T(v) for k, v in P.items()})
    head = head.to(dev)
    B, S = 2, 64
    g = torch.Generator().manual_seed(6)
    z = torch.randn(B, S, 768, generator=g).to(dev)
    cur = torch.randn(B, S, 16, generator=g).to(dev)
    for mask in (torch.ones(B, S, dtype=torch.bool), torch.rand(B, S, generator=g) > 0.5):
        mask = mask.to(dev)
        torch.manual_seed(77)
        ref = cur.clone()
        ref[mask.nonzero(as_tuple=True)] = head.sample(z[mask.nonzero(as_tuple=True)], 0.95, 1.0)   # MAR:1090, 1101, 1112
        torch.manual_seed(77)
        out = sample_masked_tokens(head, z, mask, cur, 0.95, 1.0)
        assert torch.equal(out, ref)
        assert torch.equal(out[~mask], cur[~mask])
    torch.manual_seed(77)
    full = sample_masked_tokens(head, z, torch.ones(B, S, dtype=torch.bool, device=dev), cur, 0.95, 1.0, all_masked=True)
    torch.manual_seed(77)
    assert torch.equal(full, head.sample(z.reshape(-1, 768), 0.95, 1.0).reshape(B, S, 16))
