"""GPU test of `uva_head_net_forward_static` (include/uva_head.h): the net forward on the library's static device
scratch, without a workspace and without any allocation.  Collected last on purpose (file name):
a failure here must not hide another test.

The new entry point is compared, through the C ABI, with the existing `uva_head_net_forward` on the same inputs, on two
calls in a row (the second with other inputs and another row count: the scratch keeps the first call's contents), in both
arithmetic modes.  Same kernels, same order: the tolerance is the one the other tests use for each mode (fp32 1e-4
max-relative; bf16 6e-3 max-relative between two bf16 runs), the achieved difference is expected to be ~0.
"""
import pytest
import torch

from oracle import uva_oracle as O

pytestmark = pytest.mark.gpu

FP32_TOL = 1e-4   # tests/test_gpu_infer.py
BF16_TOL = 6e-3   # tests/test_gpu_infer.py BF16_EMU_TOL


def _cuda():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)


def _head(mode, dev, C=16, Z=768, W=1024, D=6, seed=3):
    from unified_video_action_b200 import _lib

    P = O.make_params(C, Z, W, D, seed=seed)
    names = [n for n, _ in O.param_specs(C, Z, W, D)]
    with torch.cuda.device(dev):
        h = _lib.Head(C, Z, W, D, 1000, 100, mode)
        h.load_weights([torch.as_tensor(P[n]).to(dev) for n in names])
    return h


def _inputs(N, C, Z, seed, dev):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(N, C, generator=g).to(dev)
    c = torch.randn(N, Z, generator=g).to(dev)
    t = torch.randint(0, 1000, (N,), generator=g, dtype=torch.int64).to(dev)
    return x, t, c


@pytest.mark.parametrize("mode", [0, 1], ids=["fp32", "bf16"])
def test_static_scratch_net_forward_equals_workspace_entry_point(mode):
    dev = _cuda()
    from unified_video_action_b200 import _lib

    h = _head(mode, dev)
    C, Z = 16, 768
    max_rows = h.static_max_rows()
    assert max_rows == (128 if mode == 0 else 256)
    tol = FP32_TOL if mode == 0 else BF16_TOL
    with torch.cuda.device(dev):
        # two calls in a row on the shared scratch: first the maximum batch, then a smaller ragged one with other inputs
        for call, N in enumerate((max_rows, 37)):
            x, t, c = _inputs(N, C, Z, 100 + call, dev)
            ref = h.net_forward(x, t, c)                      # existing entry point, caller workspace
            out = torch.full((N, 2 * C), float("nan"), device=dev)
            torch.cuda.synchronize(dev)
            before = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
            h.net_forward_static(x, t, c, out)               # new entry point, no workspace
            after = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
            torch.cuda.synchronize(dev)
            assert after == before                            # (torch's allocator was not asked for anything either)
            assert torch.isfinite(out).all()
            err = float((out - ref).abs().max() / ref.abs().max().clamp_min(1e-6))
            assert err <= tol, (mode, N, err)
        # the same call again gives the same answer (nothing left over from the previous call matters)
        out2 = torch.empty_like(out)
        h.net_forward_static(x, t, c, out2)
        torch.cuda.synchronize(dev)
        assert float((out2 - ref).abs().max() / ref.abs().max().clamp_min(1e-6)) <= tol

        # beyond the maximum: a clear error, nothing launched
        N = max_rows + 1
        x, t, c = _inputs(N, C, Z, 7, dev)
        out = torch.zeros(N, 2 * C, device=dev)
        L = _lib.lib()
        rc = L.uva_head_net_forward_static(h._h, x.data_ptr(), t.data_ptr(), c.data_ptr(), N, out.data_ptr(), None)
        assert rc == -1
        msg = L.uva_last_error().decode()
        assert "static" in msg and str(max_rows) in msg
        with pytest.raises(RuntimeError):
            h.net_forward_static(x, t, c, out)
        torch.cuda.synchronize(dev)
        assert float(out.abs().max()) == 0.0
