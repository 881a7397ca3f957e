"""Model registry (reference: css/models/__init__.py:19-27)."""
from .blstm import BLSTM
from .conformer import Conformer

# ConvTasNet is out of scope for the hot path: the STFT-mask executor
# cannot use it and BASELINE.json's north_star does not name it.
MODELS = {"Conformer": Conformer, "BLSTM": BLSTM}


def build_model(modelname, conf):
    return MODELS[modelname].build_model(conf)
