"""ONNX checkpoint importer (css/utils/model_util.py:12-25, conf/config_*.yaml:11) and the
``backend="onnx"`` Separator (css/executor/separator.py:26-34), plus the ``css`` alias package."""
import os

import pytest
import torch

from oracle import css_oracle as O
from tests.helpers import CONF_B1, CONF_C7, cfg1, cpu_state_dict, rel_inf
from tests.onnx_util import cpu_session_module, export_onnx


def _randomised_conformer(blocks=1, idim=257, units=64, seed=5):
    import css_b200.models as M

    torch.manual_seed(seed)
    m = M.build_model("Conformer", dict(CONF_C7, conformer_num_blocks=blocks, idim=idim, conformer_linear_units=units)).eval()
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for name, p in m.named_parameters():          # trained checkpoints have distinct LayerNorm / BatchNorm values
            if "layer_norm" in name or ".BN." in name or name.startswith("conformer.embed.1"):
                p.add_(0.1 * torch.randn(p.shape, generator=g))
        for blk in m.conformer.encoders:
            blk.conv.BN.running_mean.copy_(torch.randn(256, generator=g) * 0.1)
            blk.conv.BN.running_var.copy_(torch.rand(256, generator=g) + 0.5)
    return m


def _fold_bn(sd, p):
    scale = sd[p + "BN.weight"] / torch.sqrt(sd[p + "BN.running_var"] + 1e-5)
    w = sd[p + "dw_conv_1d.weight"] * scale.view(-1, 1, 1)
    b = (sd[p + "dw_conv_1d.bias"] - sd[p + "BN.running_mean"]) * scale + sd[p + "BN.bias"]
    return w, b


def test_conformer_onnx_roundtrip(tmp_path):
    from css_b200.utils.onnx_import import state_dict_from_onnx

    m = _randomised_conformer(blocks=2)
    sd = cpu_state_dict(m)
    mod = cpu_session_module(m)
    x = torch.randn(1, 257, 40)
    with torch.no_grad():                               # the exported module IS the reference network (oracle check)
        got = mod(x)
        ref = O.conformer_masks(x.transpose(1, 2), sd)
    for a, b in zip(got, ref):
        assert rel_inf(a, b[0].t()) < 1e-5
    path = export_onnx(mod, str(tmp_path / "c.onnx"), 257, 40)
    name, conf, sd2 = state_dict_from_onnx(path)
    assert name == "Conformer"
    assert conf == dict(CONF_C7, conformer_num_blocks=2, idim=257, conformer_linear_units=64)
    assert set(sd2) == set(sd)
    for k, v in sd.items():
        if ".BN." in k or "dw_conv_1d" in k:
            continue
        assert torch.equal(sd2[k], v.float()), k
    for i in range(2):                                  # eval-mode export folds BatchNorm into the depthwise conv
        p = f"conformer.encoders.{i}.conv."
        w1, b1 = _fold_bn(sd, p)
        w2, b2 = _fold_bn(sd2, p)
        assert rel_inf(w2, w1) < 1e-6 and rel_inf(b2, b1) < 1e-6


def test_blstm_onnx_roundtrip(tmp_path):
    import css_b200.models as M
    from css_b200.utils.onnx_import import load_onnx_model

    torch.manual_seed(9)
    m = M.build_model("BLSTM", dict(CONF_B1, blstm_hdim=64, blstm_num_layers=2)).eval()
    sd = cpu_state_dict(m)
    mod = cpu_session_module(m)
    x = torch.randn(1, 257, 30)
    with torch.no_grad():
        got = mod(x)
        ref = O.blstm_masks(x.transpose(1, 2), sd)
    for a, b in zip(got, ref):
        assert rel_inf(a, b[0].t()) < 1e-5
    path = export_onnx(mod, str(tmp_path / "b.onnx"), 257, 30)
    m2, name, conf = load_onnx_model(path)
    assert name == "BLSTM" and conf["blstm_hdim"] == 64 and conf["blstm_num_layers"] == 2 and conf["idim"] == 257
    sd2 = m2.state_dict()
    assert set(sd2) == set(sd)
    for k, v in sd.items():                             # incl. the LSTM gate re-ordering (ONNX i,o,f,c -> torch i,f,g,o)
        assert torch.equal(sd2[k], v), k


def test_reader_rejects_garbage(tmp_path):
    from css_b200.utils.onnx_reader import read_onnx

    p = tmp_path / "x.onnx"
    p.write_bytes(b"\x0a\x03abc")
    with pytest.raises(ValueError):
        read_onnx(str(p))


def test_css_alias_package():
    """`import css...` (the reference's package name, scripts/python/separate_libricss.py:14-19) resolves to css_b200."""
    import css.executor.separator as a
    import css_b200.executor.separator as b
    from css.executor.beamformer import Beamformer
    from css.executor.stitcher import Stitcher
    from css.models import MODELS, build_model
    from css.objectives import OBJECTIVES
    from css.utils.model_util import get_onnx_model

    assert a is b and a.Separator is b.Separator
    assert {"Conformer", "BLSTM"} <= set(MODELS) and "MSE" in OBJECTIVES
    assert callable(get_onnx_model) and callable(build_model) and Beamformer and Stitcher
    with pytest.raises(ImportError):
        import css.no_such_module  # noqa: F401


@pytest.mark.gpu
def test_separator_onnx_backend(cuda, tmp_path):
    """Separator(cfg, backend="onnx") with model_path: *.onnx (the reference's call, separate_libricss.py:55-61) gives the
    masks of the native model the file was exported from; the session protocol answers like onnxruntime's."""
    import numpy as np

    from css_b200.executor.separator import Separator
    from css_b200.synth import make_recording
    from css_b200.utils.model_util import get_onnx_model

    m = _randomised_conformer(blocks=2, units=256)
    path = export_onnx(cpu_session_module(m), str(tmp_path / "c1.onnx"), 257, 150)
    cfg = cfg1()["separation"]
    cfg["model_path"] = path
    wav = make_recording(8, channels=1, seed=3)
    sep = Separator(cfg, sess_options=None, device="cuda", backend="onnx")
    masks, mags = sep.separate(wav)
    ref = Separator(cfg, device="cuda", backend="b200", model=m.to(cuda))
    masks_ref, _ = ref.separate(wav)
    assert rel_inf(masks, masks_ref) < 1e-5
    with pytest.raises(AssertionError):
        Separator(dict(cfg, batch_size=2), device="cuda", backend="onnx")
    session, dev = get_onnx_model(path, None, "cuda")
    f = np.random.default_rng(0).standard_normal((1, 257, 150)).astype(np.float32)
    out = session.run(None, {session.get_inputs()[0].name: f})
    with torch.no_grad():
        want = cpu_session_module(_randomised_conformer(blocks=2, units=256))(torch.from_numpy(f))
    assert len(out) == 3 and out[0].shape == (257, 150)
    for a, b in zip(out, want):
        assert rel_inf(torch.from_numpy(a), b) < 1e-4


@pytest.mark.gpu
def test_beamformer_normalize_scale(cuda):
    from css_b200.executor.beamformer import Beamformer
    from tests.helpers import cfg7

    g = torch.Generator().manual_seed(2)
    res = torch.randn(3, 257, 150, generator=g) + 1j * torch.randn(3, 257, 150, generator=g)
    ref = torch.randn(3, 257, 150, generator=g) + 1j * torch.randn(3, 257, 150, generator=g)
    mask = torch.rand(3, 257, 150, generator=g)
    got = Beamformer(cfg7()["beamforming"]).normalize_scale(res.to(cuda), ref.to(cuda), mask.to(cuda))
    want = O.normalize_scale(res, ref, mask)
    assert rel_inf(torch.view_as_real(got.cpu()), torch.view_as_real(want)) < 1e-5
