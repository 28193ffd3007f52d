"""Host-side checks that need no GPU: the C-ABI library builds, loads, exports every symbol the
header declares, its float64 schedule tables equal the reference's, and compute entry points fail
loudly (no CPU fallback) when there is no CUDA device."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from unified_video_action_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    src = open(os.path.join(ROOT, "include", "uva_head.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(uva_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_header_symbols():
    L = _lib.lib()
    names = _header_functions()
    assert len(names) >= 14
    assert sorted(names) == sorted(_lib.SYMBOLS)
    for n in names:
        assert hasattr(L, n), n
    assert b"sm_100a" in L.uva_build_info()


def test_param_table_matches_reference_shapes():
    from oracle import uva_oracle as O

    L = _lib.lib()
    for (C, Z, W, D) in ((16, 768, 1024, 6), (2, 768, 1024, 6), (4, 32, 64, 2)):
        specs = O.param_specs(C, Z, W, D)
        assert L.uva_head_num_params(D) == len(specs)
        d = _lib._Desc(C, Z, W, D, 1000, 100, 0)
        for i, (_, shape) in enumerate(specs):
            assert L.uva_head_param_numel(ctypes.byref(d), i) == int(np.prod(shape))
        assert L.uva_head_param_numel(ctypes.byref(d), len(specs)) == -1


def test_param_names_match_oracle_order():
    from oracle import uva_oracle as O
    from unified_video_action_b200.diffusion_loss import param_names

    assert param_names(6) == [n for n, _ in O.param_specs(16, 768, 1024, 6)]


def test_schedule_tables_equal_reference(golden_dir):
    g = np.load(os.path.join(golden_dir, "schedule.npz"))
    for which, tag in ((0, "gen"), (1, "train")):
        for k in g.files:
            if not k.startswith(tag + "."):
                continue
            name = k.split(".", 1)[1]
            mine = _lib.schedule_table(name, which)
            np.testing.assert_allclose(mine, g[k].astype(np.float64), rtol=1e-14, atol=0, err_msg=k)


def test_schedule_rejects_bad_arguments():
    L = _lib.lib()
    buf = (ctypes.c_double * 8)()
    assert L.uva_schedule_table(1000, 100, 0, b"no_such_table", buf, 8) < 0
    assert b"unknown table" in L.uva_last_error()
    assert L.uva_schedule_table(10, 100, 0, b"betas", buf, 8) < 0


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    with pytest.raises(RuntimeError):
        _lib.Head(16, 768, 1024, 6)
    from unified_video_action_b200 import DiffLoss

    head = DiffLoss(16, 768, 6, 1024, "100", n_frames=4, language_emb_model=None, language_emb_model_type=None)
    with pytest.raises(RuntimeError):
        head.sample(torch.zeros(4, 768))
    L = _lib.lib()
    d = _lib._Desc(16, 768, 1024, 6, 1000, 100, 1)
    h = ctypes.c_void_p()
    assert L.uva_head_create(ctypes.byref(d), ctypes.byref(h)) < 0
    assert len(L.uva_last_error()) > 0


def test_module_state_dict_keys_and_init():
    """Same state_dict keys / shapes as the reference heads and the same zero-init pattern (DL:250-259)."""
    from oracle import uva_oracle as O
    from unified_video_action_b200 import DiffActLoss, DiffLoss

    kw = dict(n_frames=4, language_emb_model=None, language_emb_model_type=None)
    head = DiffLoss(16, 768, 6, 1024, "100", **kw)
    sd = head.state_dict()
    want = {"net." + n: s for n, s in O.param_specs(16, 768, 1024, 6)}
    assert {k: tuple(v.shape) for k, v in sd.items()} == want
    assert sum(p.numel() for p in head.parameters()) == 35749920
    assert float(sd["net.final_layer.linear.weight"].abs().max()) == 0.0
    assert float(sd["net.res_blocks.3.adaLN_modulation.1.weight"].abs().max()) == 0.0
    assert float(sd["net.time_embed.mlp.0.weight"].std()) == pytest.approx(0.02, rel=0.05)
    act = DiffActLoss(2, 768, 6, 1024, "100", **kw)
    assert sum(p.numel() for p in act.parameters()) == 52225876
    assert sum(p.numel() for p in act.net.parameters()) == 35706884
    keys = set(act.state_dict().keys())
    for k in ("conv.0.weight", "fc.0.weight", "fc.2.bias", "interpolate.weight", "refine.2.weight"):
        assert k in keys


def test_shard_rows_partition():
    from unified_video_action_b200 import shard_rows

    for n, w in ((65536, 8), (1000, 3), (7, 8)):
        spans = [shard_rows(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
    assert shard_rows(131072, 1, 8, pair_halves=True) == (8192, 16384)
