"""Test helper: write ONNX files the way the reference's checkpoints were produced -- ``torch.onnx.export`` (TorchScript
exporter) of the mask network behind the session contract of css/executor/separator.py:94-103, input (1, F, T), outputs
3 x (257, T).  The css_b200 model classes are parameter containers with the reference's module tree, so a CPU forward that
calls their sub-modules exports with the same initialiser names / node scopes as the reference's own classes (checked
against an export of the reference classes in tests/golden/make_golden_r2.py).

The ``onnx`` package is not installed: the exporter only needs it for an onnxscript post-processing step that does not
apply here, so that step is bypassed.
"""
import math
import warnings

import torch

EPS = torch.finfo(torch.float32).eps


class _ConformerCpu(torch.nn.Module):
    def __init__(self, model):
        super().__init__()
        self.m = model

    def forward(self, x):                                    # (1, F, T)
        m = self.m
        enc = m.conformer
        f = x.transpose(1, 2)
        f = (f - f.mean(-2, keepdim=True)) / (f.std(-2, keepdim=True) + EPS)
        h = enc.embed(f)
        t = h.shape[1]
        heads = m.conf["attention_heads"]
        dk = h.shape[-1] // heads
        pos_k = None
        if enc.pos_emb is not None:
            idx = torch.arange(t)
            rel = (idx[:, None] - idx[None, :]).clamp(-1000, 999) + 1000
            pos_k = enc.pos_emb.pe_k(rel)
        for blk in enc.encoders:
            ff = blk.feed_forward_in
            h = h + 0.5 * ff.net(ff.layer_norm(h))
            at = blk.self_attn
            y = at.layer_norm(h)
            n = y.shape[0]
            q = at.linear_q(y).view(n, t, heads, dk).transpose(1, 2)
            k = at.linear_k(y).view(n, t, heads, dk).transpose(1, 2)
            v = at.linear_v(y).view(n, t, heads, dk).transpose(1, 2)
            a = torch.matmul(q, k.transpose(-2, -1))
            if pos_k is not None:
                qq = q.contiguous().view(n * heads, t, dk).transpose(0, 1)
                a = a + torch.matmul(qq, pos_k.transpose(-2, -1)).transpose(0, 1).view(n, heads, t, t)
            o = torch.matmul(torch.softmax(a / math.sqrt(dk), dim=-1), v).transpose(1, 2).contiguous().view(n, t, heads * dk)
            h = h + at.linear_out(o)
            cv = blk.conv
            y = cv.pw_conv_1(cv.layer_norm(h).unsqueeze(1))
            y = y[:, 0] * torch.sigmoid(y[:, 1])
            y = torch.relu(cv.BN(cv.dw_conv_1d(y.transpose(1, 2)))).transpose(1, 2)
            h = h + cv.pw_conv_2(y.unsqueeze(1)).squeeze(1)
            ff = blk.feed_forward_out
            h = h + 0.5 * ff.net(ff.layer_norm(h))
            h = blk.layer_norm(h)
        o = torch.relu(m.linear(h))
        return tuple(c[0].transpose(0, 1) for c in torch.chunk(o, m.num_spk + m.num_noise, -1))


class _BlstmCpu(torch.nn.Module):
    def __init__(self, model):
        super().__init__()
        self.m = model

    def forward(self, x):
        m = self.m
        f = x.transpose(1, 2)
        f = (f - f.mean(-2, keepdim=True)) / (f.std(-2, keepdim=True) + EPS)
        h = m.blstm.embed(f)
        for layer in m.blstm.encoders:
            h = layer.layer_norm(layer.lstm(h)[0])
        o = torch.relu(m.linear(h))
        return tuple(c[0].transpose(0, 1) for c in torch.chunk(o, m.num_spk + m.num_noise, -1))


def cpu_session_module(model):
    """CPU torch module with the session contract (1,F,T) -> 3 x (257,T) over a css_b200 (or reference) model."""
    return (_ConformerCpu if hasattr(model, "conformer") else _BlstmCpu)(model).eval()


def export_onnx(module, path, in_features, frames=150):
    from torch.onnx._internal.torchscript_exporter import onnx_proto_utils

    saved = onnx_proto_utils._add_onnxscript_fn
    onnx_proto_utils._add_onnxscript_fn = lambda proto, custom_opsets: proto
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            torch.onnx.export(module, (torch.randn(1, in_features, frames),), path, input_names=["input"],
                              output_names=["spk1", "spk2", "noise"], dynamo=False, opset_version=13,
                              dynamic_axes={"input": {2: "frames"}})
    finally:
        onnx_proto_utils._add_onnxscript_fn = saved
    return path
