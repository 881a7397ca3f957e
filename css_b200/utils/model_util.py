"""Mask-network session (reference: css/utils/model_util.py:12-31).

The reference's plug-in point is an ONNX Runtime ``InferenceSession``: an object with
``get_inputs()[0].name`` and ``run(None, {name: float32 (1, F, B*T)}) -> [spk1, spk2, noise]``
arrays of shape (257, B*T) (css/executor/separator.py:94-103).  ``B200Session`` implements that
protocol on top of a css_b200 model, so code written against the ORT session keeps working; the
executor itself bypasses the numpy round trip and calls the model on device tensors.
"""
import logging
from collections import namedtuple

import numpy as np
import torch

_Input = namedtuple("_Input", ["name", "shape", "type"])


class B200Session:
    """ORT-session-compatible wrapper of a css_b200 model (synchronous, single stream)."""

    def __init__(self, model, chunk_frames=150):
        self.model = model
        self.chunk_frames = chunk_frames

    def get_inputs(self):
        return [_Input("input", [1, self.model.in_features, "frames"], "tensor(float)")]

    def run(self, output_names, input_feed):
        (f,) = input_feed.values()
        f = np.asarray(f, dtype=np.float32)
        if f.ndim != 3 or f.shape[0] != 1 or f.shape[1] != self.model.in_features:
            raise RuntimeError(f"expected input of shape (1, {self.model.in_features}, frames), got {f.shape}")
        dev = self.model.linear.weight.device
        x = torch.from_numpy(f[0].T.copy()).to(dev)           # (frames, F)
        frames = x.shape[0]
        T = self.chunk_frames if frames % self.chunk_frames == 0 else frames
        _, masks = self.model(x.reshape(frames // T, T, -1))
        return [m.reshape(frames, -1).T.contiguous().cpu().numpy() for m in masks]


def save_model(path, model_name, conf, model):
    """Checkpoint in the reference's ``.mdl`` convention (scripts/python/train.py:304-317) plus the conf."""
    torch.save({"model_name": model_name, "conf": conf, "model": model.state_dict()}, path)


def get_b200_model(model_path, sess_options=None, device="cuda"):
    """Counterpart of ``get_onnx_model`` (model_util.py:12-25): -> (session, device)."""
    from ..models import build_model

    if str(model_path).endswith(".onnx"):
        return get_onnx_model(model_path, sess_options, device)
    ckpt = torch.load(model_path, map_location="cpu")
    model = build_model(ckpt["model_name"], ckpt["conf"])
    model.load_state_dict(ckpt["model"])
    model = model.to(device).eval()
    logging.info("loaded %s from %s", ckpt["model_name"], model_path)
    return B200Session(model), torch.device(device)


def get_onnx_model(model_path, sess_options=None, device="cuda"):
    """Reference signature (css/utils/model_util.py:12-25): -> (session, device).

    The reference builds an ``onnxruntime.InferenceSession`` here.  This build reads the weights out of the ONNX file
    (css_b200/utils/onnx_import.py; no onnxruntime involved) into the native Conformer / BLSTM and returns a session
    object with the same ``get_inputs`` / ``run`` protocol.  ``sess_options`` (ORT thread counts) has no meaning on the GPU
    and is ignored; a CPU ``device`` is rejected (no CPU path)."""
    from .onnx_import import load_onnx_model

    device = torch.device("cuda" if str(device) in ("gpu", "GPU") else device)
    if device.type != "cuda":
        raise RuntimeError("css_b200 runs only on CUDA devices (no CPU fallback): pass device='cuda'")
    model, name, _ = load_onnx_model(model_path)
    model = model.to(device).eval()
    logging.info("imported %s from ONNX file %s", name, model_path)
    return B200Session(model), device


def to_numpy(tensor):
    return tensor.detach().cpu().numpy() if tensor.requires_grad else tensor.cpu().numpy()
