"""CPU tests of the host side: C-ABI surface, placement planning, model containers, config handling."""
import ctypes
import os
import re

import pytest
import torch

from oracle import css_oracle as O
from tests.conftest import ROOT
from tests.helpers import CONF_B1, CONF_C7


def test_library_exports_every_declared_symbol():
    from css_b200 import _native as nv

    header = open(os.path.join(ROOT, "include", "css_b200.h")).read()
    names = sorted(set(re.findall(r"\b(css_[a-z0-9_]+)\s*\(", header)))
    assert len(names) >= 25
    lib = ctypes.CDLL(nv.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/css_b200.h but not exported: {missing}"
    assert lib.css_version() >= 100


def test_placement_plan_matches_oracle_and_layers_are_disjoint():
    from css_b200.executor.beamformer import layered_items, placement_plan

    for B, T in ((72, 960000), (747, 9600000), (10, 160000), (33, 38656 + 32 * 12800), (1, 38656), (4497, 57600000)):
        plan = placement_plan(B, T, 12800, 32000, 38144, 32)
        assert plan == O.placement_plan(B, T, 12800, 32000, 38144, 32)
        layers = layered_items(plan, T)
        total = 0
        for lay in layers:
            spans = sorted((d, d + (e - s)) for (_, s, e, d) in lay)
            assert all(a[1] <= b[0] for a, b in zip(spans, spans[1:])), "items of one launch overlap"
            for (c, s, e, d) in lay:
                assert 0 <= s < e <= 38144 and e - s <= 2048 and (e - 1) // 256 - s // 256 < 8 and d + (e - s) <= T
                total += e - s
        assert total == sum(p[2] - p[1] for p in plan)
        assert len(layers) <= 3


def test_model_containers_keep_reference_state_dict_names():
    import css_b200.models as M

    assert set(M.MODELS) >= {"Conformer", "BLSTM"}
    c = M.build_model("Conformer", CONF_C7)
    keys = list(c.state_dict().keys())
    for k in ("conformer.embed.0.weight", "conformer.pos_emb.pe_k.weight", "conformer.encoders.0.self_attn.linear_q.weight",
              "conformer.encoders.0.conv.pw_conv_1.weight", "conformer.encoders.1.conv.BN.running_mean", "linear.weight",
              "conformer.encoders.0.feed_forward_in.net.3.bias"):
        assert k in keys
    assert c.state_dict()["conformer.encoders.0.conv.pw_conv_1.weight"].shape == (2, 1, 1, 1)
    assert c.state_dict()["linear.weight"].shape == (771, 256)
    b = M.build_model("BLSTM", CONF_B1)
    assert "blstm.encoders.1.lstm.weight_hh_l0_reverse" in b.state_dict()
    full = M.Conformer(in_features=1799)
    assert sum(p.numel() for p in full.parameters()) == 22004835


@pytest.mark.skipif(not os.path.isdir("/root/reference/css"), reason="reference tree not present on this machine")
def test_seeded_init_equals_reference():
    import subprocess
    import sys

    code = (
        "import sys, torch; sys.dont_write_bytecode=True; sys.path.insert(0,'/root/reference'); sys.path.insert(0,%r)\n"
        "from css.models import build_model as rb; import css_b200.models as M; from tests.helpers import CONF_C7, CONF_B1\n"
        "for n,c in (('Conformer',CONF_C7),('BLSTM',CONF_B1)):\n"
        "    torch.manual_seed(3); a=rb(n,c).state_dict(); torch.manual_seed(3); b=M.build_model(n,c).state_dict()\n"
        "    assert list(a)==list(b) and all(torch.equal(a[k],b[k]) for k in a), n\n"
        "print('ok')\n" % ROOT
    )
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "ok" in out.stdout, out.stderr[-2000:]


def test_product_never_imports_the_oracle():
    for dp, _, files in os.walk(os.path.join(ROOT, "css_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f"{f} references the oracle"


def test_cpu_inputs_fail_loudly():
    from css_b200 import _native as nv
    import css_b200.models as M
    from css_b200.executor.separator import Separator
    from css_b200.pipeline import CONFIG_7CH

    with pytest.raises(nv.NativeError):
        M.build_model("BLSTM", CONF_B1)(torch.rand(1, 10, 257))
    with pytest.raises(nv.NativeError):
        Separator(CONFIG_7CH["separation"], device="cpu", model=M.build_model("Conformer", CONF_C7))
    with pytest.raises(ValueError):
        Separator(CONFIG_7CH["separation"], backend="tensorrt", model=None)


def test_steering_vectors_match_oracle():
    from css_b200.executor.separator import steervec_7ch

    sv, ang = steervec_7ch()
    sv_o, ang_o = O.steervec_7ch()
    assert torch.equal(ang, ang_o) and float((sv - sv_o).abs().max()) < 1e-6


def test_synth_is_deterministic_and_shaped():
    from css_b200.synth import make_recording, n_chunks

    a = make_recording(4, channels=7, seed=5)
    b = make_recording(4, channels=7, seed=5)
    assert a.shape == (7, 64000) and a.dtype == torch.float32 and torch.equal(a, b)
    assert float(a.abs().max()) == pytest.approx(0.5, rel=1e-5)
    assert n_chunks(960000) == 72 and n_chunks(9600000) == 747 and n_chunks(57600000) == 4497 and n_chunks(1000) == 0
