"""Guard-band checks of every buffer the library writes (compute-sanitizer is closed on this GPU pool -
profiles/r02_sanitizer_closed.log - so out-of-bounds WRITES are hunted with sentinels of our own): outputs and the
caller-allocated workspace are carved out of larger allocations whose surroundings hold a sentinel pattern; the
workspace is exactly as large as uva_head_workspace_bytes asks for.  After the call (latency path with the chain
kernel and with the graph path, layer-chain kernel, per-layer chunked path with two lanes, fp32 mode, training forward
+ backward, decoder-tail source) every sentinel must be intact and nothing inside may be NaN."""
import ctypes

import numpy as np
import pytest
import torch

from oracle import uva_oracle as O

pytestmark = pytest.mark.gpu
GUARD = 1 << 16  # floats on each side
SENT = 0x7FC0DEAD  # a NaN payload nobody computes


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)


class Guarded:
    """A float32 / byte region with sentinel-filled guard bands on both sides."""

    def __init__(self, numel, dev, dtype=torch.float32):
        self.words = (numel * torch.empty(0, dtype=dtype).element_size() + 3) // 4
        pad = (-self.words) % 64  # keep the region's start 256-byte aligned like a fresh allocation would be
        self.buf = torch.full((GUARD + self.words + pad + GUARD,), SENT, dtype=torch.int32, device=dev)
        self.inner = self.buf[GUARD:GUARD + self.words]
        self.inner.zero_()
        self.t = self.inner.view(dtype)[:numel]

    def check(self, what):
        lo, hi = self.buf[:GUARD], self.buf[GUARD + self.words + ((-self.words) % 64):]
        pad = self.buf[GUARD + self.words:GUARD + self.words + ((-self.words) % 64)]
        assert bool((lo == SENT).all()), f"{what}: write BEFORE the buffer"
        assert bool((hi == SENT).all()), f"{what}: write AFTER the buffer"
        assert bool((pad == SENT).all()), f"{what}: write just past the end of the buffer"


def _head(C, mode, dev, seed=1):
    from unified_video_action_b200 import _lib

    P = O.make_params(C, 768, 1024, 6, seed=seed)
    names = [n for n, _ in O.param_specs(C, 768, 1024, 6)]
    h = _lib.Head(C, 768, 1024, 6, 1000, 100, mode)
    h.load_weights([torch.from_numpy(P[n]).to(dev) for n in names])
    return h, P, names


@pytest.mark.parametrize("N,cfg,mode,env", [
    (16, 1.0, 1, {}),                                   # persistent chain kernel
    (27, 1.0, 1, {"UVA_CHAIN": "0"}),                   # ragged rows on the graph path
    (300, 3.0, 1, {}),                                  # graph path with CFG
    (1311, 1.0, 1, {}),                                 # layer-chain kernel, ragged last tile
    (5000, 3.0, 1, {}),                                 # layer-chain kernel with CFG pairs
    (5000, 1.0, 1, {"UVA_STACK": "0"}),                 # per-layer launches, two lanes
    (40000, 3.0, 1, {"UVA_CHUNK_ROWS": "8192"}),        # several chunk rounds
    (333, 1.0, 0, {}),                                  # fp32 mode
])
def test_sampler_writes_stay_inside_its_buffers(monkeypatch, N, cfg, mode, env):
    dev = _cuda()
    from unified_video_action_b200 import _lib

    for k, v in env.items():
        monkeypatch.setenv(k, v)
    h, _, _ = _head(16, mode, dev)
    L = _lib.lib()
    g = torch.Generator().manual_seed(N)
    steps = 3
    z = torch.randn(N, 768, generator=g).to(dev)
    x0 = torch.randn(N, 16, generator=g).to(dev)
    if cfg != 1.0:
        x0[N // 2:] = x0[:N // 2]
    nz = torch.randn(steps, N, 16, generator=g).to(dev)
    out, tr = Guarded(N * 16, dev), Guarded(steps * N * 16, dev)
    ws = Guarded(int(L.uva_head_workspace_bytes(h._h, N, 0)), dev, torch.uint8)
    for use_trace in (False, True, False):  # eager / graph capture / replay on the latency path; trace = chunked path
        rc = L.uva_head_sample(h._h, z.data_ptr(), x0.data_ptr(), nz.data_ptr(), None, tr.t.data_ptr() if use_trace else None,
                               0.95, cfg, 0, 0, steps, N, out.t.data_ptr(), ws.t.data_ptr(), ws.t.numel(),
                               torch.cuda.current_stream().cuda_stream)
        assert rc == 0, L.uva_last_error()
        torch.cuda.synchronize()
        for name, gb in (("out", out), ("trace", tr), ("workspace", ws)):
            gb.check(f"sample N={N} {name}")
        assert torch.isfinite(out.t).all()


@pytest.mark.parametrize("N,C,mode", [(333, 16, 1), (333, 16, 0), (256, 2, 1), (2500, 16, 1)])
def test_training_writes_stay_inside_their_buffers(N, C, mode):
    dev = _cuda()
    from unified_video_action_b200 import _lib

    h, P, names = _head(C, mode, dev, seed=2)
    L = _lib.lib()
    g = torch.Generator().manual_seed(N + C)
    target = torch.randn(N, C, generator=g).to(dev)
    z = torch.randn(N, 768, generator=g).to(dev)
    mask = (torch.rand(N, generator=g) > 0.3).float().to(dev)
    t = torch.randint(0, 1000, (N,), generator=g).to(dev)
    noise = torch.randn(N, C, generator=g).to(dev)
    loss, rows, dz = Guarded(1, dev), Guarded(N, dev), Guarded(N * 768, dev)
    grads = [Guarded(int(np.prod(P[n].shape)), dev) for n in names]
    ws = Guarded(int(L.uva_head_workspace_bytes(h._h, N, 1)), dev, torch.uint8)
    st = torch.cuda.current_stream().cuda_stream
    one = torch.ones(1, device=dev)
    arr = (ctypes.c_void_p * len(grads))(*[gr.t.data_ptr() for gr in grads])
    for _ in range(3):  # eager, graph capture, replay for the small sizes
        rc = L.uva_head_loss_fwd(h._h, target.data_ptr(), z.data_ptr(), mask.data_ptr(), t.data_ptr(), noise.data_ptr(), N,
                                 loss.t.data_ptr(), rows.t.data_ptr(), ws.t.data_ptr(), ws.t.numel(), st)
        assert rc == 0, L.uva_last_error()
        rc = L.uva_head_loss_bwd(h._h, one.data_ptr(), target.data_ptr(), z.data_ptr(), mask.data_ptr(), t.data_ptr(),
                                 noise.data_ptr(), N, dz.t.data_ptr(), arr, ws.t.data_ptr(), ws.t.numel(), st)
        assert rc == 0, L.uva_last_error()
        torch.cuda.synchronize()
        for name, gb in [("loss", loss), ("rows", rows), ("dz", dz), ("workspace", ws)] + list(zip(names, grads)):
            gb.check(f"train N={N} {name}")
    assert torch.isfinite(loss.t).all() and torch.isfinite(dz.t).all()


def test_decoder_tail_source_writes_stay_inside(monkeypatch):
    dev = _cuda()
    from unified_video_action_b200 import _lib

    h, _, _ = _head(16, 1, dev)
    L = _lib.lib()
    B, T, S, skip = 3, 4, 100, 7
    N = B * T * S
    g = torch.Generator().manual_seed(5)
    x_dec = torch.randn(B, skip + T * S, 768, generator=g).to(dev)
    tail = _lib.DecoderTail(torch.ones(768, device=dev), torch.zeros(768, device=dev), 1e-6,
                            0.02 * torch.randn(1, T, 768, generator=g).to(dev), 0.02 * torch.randn(1, S, 768, generator=g).to(dev), skip=skip)
    zrows = Guarded(N * 768, dev)
    ts = tail.struct(x_dec.shape[1])
    rc = L.uva_decoder_tail_rows(x_dec.data_ptr(), ctypes.byref(ts), None, N, 768, zrows.t.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.uva_last_error()
    torch.cuda.synchronize()
    zrows.check("decoder_tail_rows")
    x0 = torch.randn(N, 16, generator=g).to(dev)
    nz = torch.randn(2, N, 16, generator=g).to(dev)
    out = Guarded(N * 16, dev)
    ws = Guarded(int(L.uva_head_workspace_bytes(h._h, N, 0)), dev, torch.uint8)
    rc = L.uva_head_sample_from_decoder(h._h, x_dec.data_ptr(), ctypes.byref(ts), None, x0.data_ptr(), nz.data_ptr(), None, None,
                                        0.95, 1.0, 0, 0, 2, N, out.t.data_ptr(), ws.t.data_ptr(), ws.t.numel(),
                                        torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.uva_last_error()
    torch.cuda.synchronize()
    out.check("sample_from_decoder out")
    ws.check("sample_from_decoder workspace")
    assert torch.isfinite(out.t).all()
