"""CPU: the oracle restatement (oracle/depth_oracle.py) against the golden vectors produced by
the REAL reference classes (oracle/make_golden.py, run where /root/reference exists)."""
import numpy as np
import pytest
import torch

from helpers import build_model, load_golden, synth_images, synth_state_dict
from oracle import depth_oracle as O
from oracle.cases import CASES, FUSION_CASES, REFINE_CASES, key_name, stride_for

import monodepth_estimation_b200 as mde

FAST = ["std_224_b2_none", "std_256_b1_byte_other", "std_256_b1_signed_none", "std_160x200_b2_ing", "enh_224_b2_none",
        "enh_320x256_b1_sat", "std_256x320_b1_ingalt", "cfg4_256_b1_rover", "std_518_b1_rover", "enh_518_b1_roverh"]


def check_against_golden(out, gold, tol=2e-5):
    for k, v in out.items():
        n = key_name(k) if isinstance(k, tuple) else k
        ref = torch.from_numpy(gold[n])
        assert tuple(gold[n + "__shape"]) == tuple(v.shape), (k, v.shape)
        sy, sx = stride_for(v.shape[2]), stride_for(v.shape[3])
        got = v[:, :, ::sy, ::sx]
        err = float((got - ref).norm() / ref.norm())
        assert err < tol, f"{k}: rel l2 {err}"
        st = gold[n + "__stats"]
        assert abs(float(v.double().mean()) - st[0]) < 1e-4 * max(1.0, abs(st[0]))
    assert {key_name(k) if isinstance(k, tuple) else k for k in out} == {f[:-7] for f in gold.files if f.endswith("__stats")}


_SD = {}


def state(kind, seed):
    if (kind, seed) not in _SD:
        _SD[(kind, seed)] = build_model(kind, seed).state_dict()
    return _SD[(kind, seed)]


@pytest.mark.parametrize("name", FAST)
def test_oracle_matches_reference_golden(name):
    kind, B, H, W, ikind, md, wseed, iseed = CASES[name]
    x = synth_images(B, H, W, iseed, ikind)
    with torch.no_grad():
        if kind == "standard":
            out = O.standard_forward(state("standard", wseed), x, md)
        elif kind == "enhanced":
            out = O.enhanced_forward(state("enhanced", wseed), x, md)
        else:
            fu, rf = mde.MultiScaleDepthFusion(), mde.TextureAwareRefinementModule()
            out = O.config4_forward(state("enhanced", wseed), synth_state_dict(fu.state_dict(), wseed + 100),
                                    synth_state_dict(rf.state_dict(), wseed + 200), x, md)
    check_against_golden(out, load_golden(name))


@pytest.mark.parametrize("name", list(REFINE_CASES))
def test_refine_module_golden(name):
    H, W, rh, rw, seed = REFINE_CASES[name]
    sd = synth_state_dict(mde.TextureAwareRefinementModule().state_dict(), seed)
    depth = synth_images(1, H, W, seed, "unit")[:, :1].contiguous()
    rgb = synth_images(1, rh, rw, seed + 1, "unit")
    with torch.no_grad():
        check_against_golden({"refined": O.texture_refine_module(sd, depth, rgb)}, load_golden(name))


@pytest.mark.parametrize("name", list(FUSION_CASES))
def test_fusion_module_golden(name):
    H, W, seed = FUSION_CASES[name]
    sd = synth_state_dict(mde.MultiScaleDepthFusion().state_dict(), seed)
    maps = {("disp", s): synth_images(1, H >> s, W >> s, seed + s, "unit")[:, :1].contiguous() for s in range(4)}
    with torch.no_grad():
        check_against_golden({"fused": O.multiscale_depth_fusion(sd, maps)}, load_golden(name))


def test_source_scale_table():
    """model.py:63-87 == enhanced_model.py:218-242."""
    cases = [(None, None), ({}, None), ({"source": "rover"}, 1.0), ({"source": "rover", "height": 1.6}, 2.0 / 1.6),
             ({"source": "rover", "height": 0.01}, 20.0), ({"source": "ingenuity"}, 0.5), ({"source": "ingenuity", "altitude": 5.0}, 2.0),
             ({"source": "ingenuity", "altitude": 0.2}, 10.0), ({"source": "satellite"}, 0.01), ({"source": "other"}, 1.0)]
    for md, want in cases:
        assert O.source_scale(md) == want
        assert mde.source_scale(md) == want


def test_tokens_to_grid_quirks():
    """encoder.py:69-110: CLS stripped only for an exact 14x14 grid, else kept + 13 zero cells (14x15)."""
    tok = torch.arange(197 * 2, dtype=torch.float32).view(1, 197, 2)
    g = O.tokens_to_grid(tok, 14, 14)
    assert g.shape == (1, 2, 14, 14) and g[0, 0, 0, 0] == 2.0
    g = O.tokens_to_grid(tok, 64, 64)
    assert g.shape == (1, 2, 14, 15) and g[0, 0, 0, 0] == 0.0 and float(g[0, :, 13, 2:].abs().sum()) == 0.0
