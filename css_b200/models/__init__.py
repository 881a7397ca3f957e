"""Model registry (reference: css/models/__init__.py:19-27)."""
from .blstm import BLSTM
from .conformer import Conformer
from .conv_tasnet import ConvTasNet

MODELS = {"Conformer": Conformer, "ConvTasNet": ConvTasNet, "BLSTM": BLSTM}


def build_model(modelname, conf):
    return MODELS[modelname].build_model(conf)
