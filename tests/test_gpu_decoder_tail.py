"""the MAR decoder tail (decoder_norm + diffusion pos-embed, mar_con_unified.py:702-724) fused into
the head as the prologue of cond_embed.  (i) the rows the kernel produces against the reference golden
(tests/golden/decoder_tail.npz, fp32 1e-5: LayerNorm statistics are summed in another order than ATen's);
(ii) sampling from the decoder output == sampling on the materialised z (same kernels after the prologue:
bit-equal in bf16 mode on every path - chain kernel, graph path, chunked path, CFG, row subsets);
(iii) the module API DiffLoss.sample_from_decoder against the oracle on the reference's RNG order."""
import os

import numpy as np
import pytest
import torch

from oracle import uva_oracle as O

pytestmark = pytest.mark.gpu


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)


def _T(a, dev=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    return t.to(dev) if dev is not None else t


def _tail_inputs(dev, B, T, S, skip, seed):
    g = torch.Generator().manual_seed(seed)
    D = 768
    x_dec = (2.0 * torch.randn(B, skip + T * S, D, generator=g) + 0.3).to(dev)
    ln_w = (1.0 + 0.1 * torch.randn(D, generator=g)).to(dev)
    ln_b = (0.1 * torch.randn(D, generator=g)).to(dev)
    te = (0.02 * torch.randn(1, T, D, generator=g)).to(dev)
    se = (0.02 * torch.randn(1, S, D, generator=g)).to(dev)
    return x_dec, ln_w, ln_b, te, se


def test_decoder_tail_rows_vs_reference_golden(golden_dir):
    dev = _cuda()
    from unified_video_action_b200 import _lib

    g = np.load(os.path.join(golden_dir, "decoder_tail.npz"))
    B, T, S, D, skip = [int(v) for v in g["meta"]]
    tail = _lib.DecoderTail(_T(g["ln_w"], dev), _T(g["ln_b"], dev), float(g["eps"][0]), _T(g["temporal"], dev),
                            _T(g["spatial"], dev), skip=skip)
    z = _lib.decoder_tail_rows(_T(g["x_dec"], dev), tail).cpu()
    ref = _T(g["z"]).reshape(-1, D)
    assert float((z - ref).abs().max()) <= 1e-5 * float(ref.abs().max())
    idx = torch.tensor([5, 0, 47, 23, 24, 5])  # a gathered subset (MAR:1090), repeats allowed
    zi = _lib.decoder_tail_rows(_T(g["x_dec"], dev), tail, idx).cpu()
    assert torch.equal(zi, z[idx])


@pytest.mark.parametrize("B,T,S,skip,cfg,k_end,mode", [
    (1, 4, 4, 0, 1.0, 100, 1),     # 16 rows: persistent chain kernel
    (2, 4, 64, 64, 1.0, 7, 1),     # 512 rows, text tokens stripped: graph path
    (2, 4, 64, 0, 3.0, 5, 1),      # CFG on the graph path
    (3, 4, 256, 64, 1.0, 3, 1),    # 3072 rows: chunked path, two lanes
    (4, 4, 256, 0, 2.0, 2, 1),     # 4096 rows cond || uncond: chunked path with the CFG pairing
    (1, 4, 64, 8, 1.0, 4, 0),      # fp32 mode
])
def test_sample_from_decoder_equals_sample_on_materialised_z(B, T, S, skip, cfg, k_end, mode):
    dev = _cuda()
    from unified_video_action_b200 import _lib

    P = O.make_params(16, 768, 1024, 6, seed=1)
    names = [n for n, _ in O.param_specs(16, 768, 1024, 6)]
    h = _lib.Head(16, 768, 1024, 6, 1000, 100, mode)
    h.load_weights([_T(P[n], dev) for n in names])
    x_dec, ln_w, ln_b, te, se = _tail_inputs(dev, B, T, S, skip, seed=B * 100 + S)
    tail = _lib.DecoderTail(ln_w, ln_b, 1e-6, te, se, skip=skip)
    z = _lib.decoder_tail_rows(x_dec, tail)
    ref_z = O.decoder_tail(x_dec.cpu(), ln_w.cpu(), ln_b.cpu(), 1e-6, te.cpu(), se.cpu(), skip).reshape(-1, 768)
    assert float((z.cpu() - ref_z).abs().max()) <= 1e-5 * float(ref_z.abs().max())
    N = B * T * S
    g = torch.Generator().manual_seed(N)
    if cfg != 1.0:
        half = torch.randn(N // 2, 16, generator=g)
        x0 = torch.cat([half, half]).to(dev)
    else:
        x0 = torch.randn(N, 16, generator=g).to(dev)
    nz = torch.randn(k_end, N, 16, generator=g).to(dev)
    kw = dict(temperature=0.95, cfg=cfg, clip_denoised=False, k_begin=0, k_end=k_end)
    for _ in range(3):  # eager, graph capture, graph replay on the latency path
        a = h.sample(z, x0, nz, **kw)
        b = h.sample_from_decoder(x_dec, tail, x0, nz, **kw)
        assert torch.equal(a, b)
    # a row subset through row_index (cfg needs paired halves: only without cfg here)
    if cfg == 1.0 and N >= 64:
        idx = torch.randperm(N, generator=g)[: N // 2].to(dev)
        a = h.sample(z[idx], x0[idx], nz[:, idx], **kw)
        b = h.sample_from_decoder(x_dec, tail, x0[idx], nz[:, idx], row_index=idx, **kw)
        assert torch.equal(a, b)


def test_module_sample_from_decoder_vs_oracle():
    """DiffLoss.sample_from_decoder(x_dec, decoder_norm, temporal, spatial) == oracle sampling on
    oracle.decoder_tail(x_dec) with the reference's RNG call order (fp32 mode, 20-step respacing)."""
    dev = _cuda()
    from unified_video_action_b200 import DiffLoss

    P = O.make_params(16, 768, 1024, 6, seed=1)
    head = DiffLoss(16, 768, 6, 1024, "20", n_frames=4, language_emb_model=None, language_emb_model_type=None,
                    precision="fp32")
    head.net.load_state_dict({k: _T(v) for k, v in P.items()})
    head = head.to(dev)
    B, T, S, skip = 1, 4, 8, 64
    x_dec, ln_w, ln_b, te, se = _tail_inputs(dev, B, T, S, skip, seed=3)
    norm = torch.nn.LayerNorm(768, eps=1e-6).to(dev)
    with torch.no_grad():
        norm.weight.copy_(ln_w)
        norm.bias.copy_(ln_b)
    torch.manual_seed(31)
    out = head.sample_from_decoder(x_dec, norm, te, se, temperature=0.95, cfg=1.0, skip_tokens=skip)
    N = B * T * S
    assert out.shape == (N, 16)
    torch.manual_seed(31)
    x0 = torch.randn(N, 16)
    nz = torch.stack([torch.randn(N, 16, device=dev) for _ in range(20)]).cpu()
    z = O.decoder_tail(x_dec.cpu(), ln_w.cpu(), ln_b.cpu(), 1e-6, te.cpu(), se.cpu(), skip).reshape(-1, 768)
    ref = O.sample_loop(O.to_torch_params(P), O.create_schedule("20"), z, x0, nz, temperature=0.95, cfg=1.0)
    assert float((out.cpu() - ref).abs().max()) <= 2e-3 * float(ref.abs().max())
    # and it equals the two-step form a caller of the reference would write
    torch.manual_seed(31)
    two = head.sample(torch.from_numpy(z.numpy()).to(dev), temperature=0.95, cfg=1.0)
    assert float((out - two).abs().max()) <= 1e-4 * float(two.abs().max())
