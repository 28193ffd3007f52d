"""Every path-selecting environment switch of INTEGRATION.md section 4 that has no test of its own: the flipped
setting must give the default's result.  Switches that only change scheduling (graphs, programmatic dependent launch,
stream / SM partitioning, grid sizes, tile widths with the same k order) are held to bit equality; switches that
change where a value is rounded or which SiLU form an epilogue uses are held to the bf16 tolerance of the other tests.
"""
import pytest
import torch

from oracle import uva_oracle as O

pytestmark = pytest.mark.gpu

BF16_EMU_TOL = 6e-3   # tests/test_gpu_infer.py: two bf16 evaluations of the same chain, max|diff| / max|ref|


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)


@pytest.fixture(scope="module")
def params():
    dev = _cuda()
    C, Z, W, D = 16, 768, 1024, 6
    P = O.make_params(C, Z, W, D, seed=1)
    names = [n for n, _ in O.param_specs(C, Z, W, D)]
    return dev, P, names


def _head(dev, P, names, mode=1):
    from unified_video_action_b200 import _lib

    with torch.cuda.device(dev):
        h = _lib.Head(16, 768, 1024, 6, 1000, 100, mode)
        h.load_weights([torch.as_tensor(P[n]).to(dev) for n in names])
    return h


def _sample(dev, P, names, N, cfg, k_end, calls=2):
    gen = torch.Generator().manual_seed(N)
    z = torch.randn(N, 768, generator=gen).to(dev)
    x0 = torch.randn(N, 16, generator=gen)
    if cfg != 1.0:
        x0[N // 2:] = x0[:N // 2]
    x0 = x0.to(dev)
    nz = torch.randn(k_end, N, 16, generator=gen).to(dev)
    h = _head(dev, P, names)
    outs = [h.sample(z, x0, nz, temperature=0.95, cfg=cfg, k_begin=0, k_end=k_end).clone() for _ in range(calls)]
    torch.cuda.synchronize()
    for o in outs[1:]:
        assert torch.equal(o, outs[0])  # eager call == graph replay / second call
    assert torch.isfinite(outs[0]).all()
    return outs[0]


# (switch, flipped value, rows, cfg, denoise steps, tolerance: 0 = bit equal)
SAMPLING = [
    ("UVA_GRAPH", "0", 300, 1.0, 6, 0.0),          # latency path without the CUDA graph
    ("UVA_PDL", "0", 300, 1.0, 6, 0.0),            # ... without programmatic dependent launch
    ("UVA_FUSE_INPROJ", "0", 300, 1.0, 6, BF16_EMU_TOL),  # input_proj and the first LayerNorm-modulate as two kernels (the row statistics are summed in another order: 1.7e-3 measured)
    ("UVA_BN32", "0", 300, 1.0, 6, BF16_EMU_TOL),  # CTA-pair 256 x 256 tiles instead of the one-wave narrow tiles (the pair kernel's SiLU epilogue is the one-MUFU form: 1.2e-3 measured)
    ("UVA_HOIST_SMS", "20", 640, 1.0, 40, 0.0),    # another SM share for the overlapped modulation hoist
    ("UVA_CHAIN_ROWS", "0", 16, 1.0, 6, BF16_EMU_TOL),  # 16 rows on the graph path instead of the persistent chain kernel
    ("UVA_PDL_BIG", "0", 40000, 3.0, 3, 0.0),      # chunked path without programmatic dependent launch
    ("UVA_LANE_SPLIT", "0", 140000, 3.0, 2, 0.0),  # two chunk lanes, each on the whole chip
    ("UVA_OVERLAP", "0", 140000, 3.0, 2, 0.0),     # one lane
    ("UVA_GRID_MULT", "2", 40000, 3.0, 3, 0.0),    # GEMM grids of twice the co-resident clusters
    ("UVA_QUAD", "1", 40000, 3.0, 3, 0.0),         # clusters of two CTA pairs with multicast A boxes
    ("UVA_PAIR", "0", 40000, 1.0, 3, BF16_EMU_TOL),  # single-CTA 128 x 256 kernel instead of the CTA-pair kernel
]


@pytest.mark.parametrize("switch,value,N,cfg,k_end,tol", SAMPLING, ids=[f"{s[0]}={s[1]}" for s in SAMPLING])
def test_sampling_switch_equals_default(params, monkeypatch, switch, value, N, cfg, k_end, tol):
    dev, P, names = params
    monkeypatch.delenv(switch, raising=False)
    ref = _sample(dev, P, names, N, cfg, k_end)
    monkeypatch.setenv(switch, value)
    out = _sample(dev, P, names, N, cfg, k_end)
    if tol == 0.0:
        assert torch.equal(out, ref), (float((out - ref).abs().max()), float(ref.abs().max()))
    else:
        assert float((out - ref).abs().max() / ref.abs().max()) <= tol


TRAINING = [
    ("UVA_TRAIN_GRAPH", "0", 256),       # small batch eagerly instead of the captured graph
    ("UVA_TRAIN_GRAPH_ROWS", "4096", 3000),  # a large batch through the graph path (no side-stream weight gradients)
    ("UVA_PDL_TRAIN", "1", 3000),        # programmatic dependent launch between the GEMMs of the training pass
]


@pytest.mark.parametrize("switch,value,N", TRAINING, ids=[f"{s[0]}={s[1]}" for s in TRAINING])
def test_training_switch_equals_default(params, monkeypatch, switch, value, N):
    """Same kernels on the same operands: loss and d z bit-equal, parameter gradients equal up to the order of the
    split-K / column-sum atomics (1e-5), on three calls in a row (eager, capture, replay)."""
    dev, P, names = params
    gen = torch.Generator().manual_seed(N)
    target = (torch.randn(N, 16, generator=gen) * 0.5).to(dev)
    z = torch.randn(N, 768, generator=gen).to(dev)
    mask = (torch.rand(N, generator=gen) > 0.25).float().to(dev)
    t = torch.randint(0, 1000, (N,), generator=gen).to(dev)
    noise = torch.randn(N, 16, generator=gen).to(dev)
    one = torch.ones(1, device=dev)
    res = {}
    for flag in (None, value):
        if flag is None:
            monkeypatch.delenv(switch, raising=False)
        else:
            monkeypatch.setenv(switch, flag)
        h = _head(dev, P, names)
        grads = [torch.zeros(P[n].shape, dtype=torch.float32, device=dev) for n in names]
        runs = []
        for _ in range(3):
            for g in grads:
                g.zero_()
            loss = h.loss_fwd(target, z, mask, t, noise)
            dz = h.loss_bwd(one, target, z, mask, t, noise, grads)
            torch.cuda.synchronize()
            runs.append((float(loss), dz.clone(), [g.clone() for g in grads]))
        for l, dz, gr in runs[1:]:
            assert l == runs[0][0] and torch.equal(dz, runs[0][1])
            for a, b, n in zip(gr, runs[0][2], names):
                assert float((a - b).norm() / b.norm().clamp_min(1e-20)) <= 1e-5, (flag, n)
        res[flag] = runs[0]
    assert res[value][0] == res[None][0] and torch.equal(res[value][1], res[None][1])
    for a, b, n in zip(res[value][2], res[None][2], names):
        assert float((a - b).norm() / b.norm().clamp_min(1e-20)) <= 1e-5, n
