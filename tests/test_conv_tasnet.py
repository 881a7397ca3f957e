"""Conv-TasNet (css/models/conv_tasnet.py) -- registry key, reference state_dict names, native forward
against the reference's own output (tests/golden/golden_r2.npz case I, written by make_golden_r2.py from the reference class)."""
import os

import numpy as np
import pytest
import torch

from oracle import css_oracle as O
from tests.helpers import CONF_CTN, conv_tasnet_input, cpu_state_dict, randomise_conv_tasnet, rel_inf

ROOT = os.path.dirname(os.path.abspath(__file__))


def _model():
    import css_b200.models as M

    assert list(M.MODELS) == ["Conformer", "ConvTasNet", "BLSTM"]        # css/models/__init__.py:19-27
    torch.manual_seed(2)
    return randomise_conv_tasnet(M.build_model("ConvTasNet", CONF_CTN)).eval()


def test_oracle_matches_reference_output():
    g = np.load(os.path.join(ROOT, "golden", "golden_r2.npz"))
    m = _model()
    with torch.no_grad():
        y = O.conv_tasnet_forward(conv_tasnet_input(), cpu_state_dict(m))
    assert y.shape == (2, 2, 4000) and rel_inf(y, g["I_out"]) < 1e-5
    with pytest.raises(ValueError):
        import css_b200.models as M

        M.build_model("ConvTasNet", dict(CONF_CTN, conv_tasnet_norm="bn"))


@pytest.mark.gpu
def test_native_forward_matches_reference(cuda):
    g = np.load(os.path.join(ROOT, "golden", "golden_r2.npz"))
    m = _model()
    sd = cpu_state_dict(m)
    x = conv_tasnet_input()
    y = m.to(cuda)(x.to(cuda))
    assert y.shape == (2, 2, 4000)
    assert rel_inf(y, g["I_out"]) < 1e-4
    # another length / batch of one / 1-D input, against the oracle
    x2 = conv_tasnet_input(seed=9, n=1, length=2408)[0]
    with torch.no_grad():
        ref = O.conv_tasnet_forward(x2, sd)
    assert rel_inf(m(x2.to(cuda)), ref) < 1e-4
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 2, 100, device=cuda))
