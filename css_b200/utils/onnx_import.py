"""ONNX checkpoint importer: the reference's ``model_path: checkpoints/*.onnx`` (conf/config_7ch.yaml:11,
conf/config_1ch.yaml:11; loaded by css/utils/model_util.py:12-25, consumed at css/executor/separator.py:26-34,94-103)
-> a css_b200 Conformer / BLSTM with the same weights.

The graph is NOT interpreted: the mask networks are fixed architectures (css/models/conformer.py, css/models/blstm.py), so
the importer only has to find every ``state_dict`` entry among the graph's initialisers:

* tensors the exporter kept under their parameter name (biases, LayerNorm / scalar-conv / embedding weights), possibly
  aliased through ``Identity`` nodes (the exporter de-duplicates equal constants);
* ``nn.Linear`` weights, which ``torch.onnx.export`` stores transposed under anonymous names (``onnx::MatMul_123``) as
  the second input of a ``MatMul`` node: named from the node's module scope (``/conformer/encoders.0/.../net.0/MatMul``)
  or, for exporters without scope names, from the named bias that the following ``Add`` consumes;
* the depthwise ``Conv`` of the conv module with BatchNorm folded in by the eval-mode export (anonymous weight + bias):
  stored as ``dw_conv_1d`` with an identity BatchNorm;
* ``LSTM`` nodes (W, R, B with ONNX gate order i,o,f,c) -> ``nn.LSTM`` parameters (gate order i,f,g,o).

Parity status: the reference ships no checkpoint, so this path is pinned only against files written by
``torch.onnx.export`` of the reference's own model classes (tests/golden/make_golden_r2.py) -- "parity unpinned" for any
third-party export.
"""
import re

import numpy as np
import torch

from .onnx_reader import read_onnx

_ROOT = re.compile(r"(conformer|blstm)\.")
_HEAD = re.compile(r"(^|\.)linear\.(weight|bias)$")


def _canonical(name):
    """'m.conformer.embed.0.bias' -> 'conformer.embed.0.bias'; 'm.linear.bias' -> 'linear.bias'; else None."""
    m = _ROOT.search(name)
    if m:
        return name[m.start():]
    m = _HEAD.search(name)
    if m:
        return "linear." + m.group(2)
    return None


def _scope_path(value_name):
    """'/conformer/encoders.0/feed_forward_in/net/net.0/MatMul_output_0' -> 'conformer.encoders.0.feed_forward_in.net.0'."""
    parts = [p for p in value_name.split("/") if p][:-1]
    out = []
    for p in parts:
        if out and p.startswith(out[-1] + "."):
            out[-1] = p
        else:
            out.append(p)
    path = ".".join(out)
    m = _ROOT.search(path + ".")
    if m:
        return path[m.start():]
    return "linear" if path == "linear" or path.endswith(".linear") else None


def _lstm_reorder(x, H):
    """ONNX gate blocks (i, o, f, c) along dim 0 -> torch (i, f, g, o)."""
    i, o, f, c = x[0:H], x[H:2 * H], x[2 * H:3 * H], x[3 * H:4 * H]
    return np.concatenate([i, f, c, o], 0)


def state_dict_from_onnx(path):
    """-> (model_name, conf, state_dict of torch tensors keyed like the reference's ``model.state_dict()``)."""
    g = read_onnx(path)
    ini = dict(g["initializers"])
    for n in g["nodes"]:                       # Identity aliases of initialisers
        if n["op_type"] == "Identity" and n["input"] and n["input"][0] in ini:
            ini[n["output"][0]] = ini[n["input"][0]]
    sd = {}
    for name, arr in ini.items():
        key = _canonical(name)
        if key is not None and np.issubdtype(arr.dtype, np.floating):
            sd[key] = arr
    consumers = {}
    for n in g["nodes"]:
        for i in n["input"]:
            consumers.setdefault(i, []).append(n)

    def bias_follow(out_name):
        for c in consumers.get(out_name, []):
            if c["op_type"] == "Add":
                for i in c["input"]:
                    k = _canonical(i) if i in ini else None
                    if k and k.endswith(".bias"):
                        return k[: -len(".bias")]
        return None

    folded_bn = []
    last_conv_prefix, n_lstm = None, 0           # graph-order fallbacks for exporters that do not record module scopes
    for n in g["nodes"]:
        op = n["op_type"]
        if op == "Conv" and n["input"][1] in ini:
            k = _canonical(n["input"][1])
            if k and k.endswith("pw_conv_1.weight"):
                last_conv_prefix = k[: -len("pw_conv_1.weight")]
        if op in ("MatMul", "Gemm") and len(n["input"]) >= 2 and n["input"][1] in ini and _canonical(n["input"][1]) is None:
            mod = _scope_path(n["output"][0]) or _scope_path("/" + n["name"]) or bias_follow(n["output"][0])
            if mod is None:
                raise ValueError(f"{path}: cannot name the weight of node {n['name'] or n['output'][0]}")
            w = ini[n["input"][1]]
            transposed = op == "MatMul" or not n["attr"].get("transB", 0)
            sd[mod + ".weight"] = np.ascontiguousarray(w.T) if transposed else w
            if op == "Gemm" and len(n["input"]) > 2 and n["input"][2] in ini:
                sd.setdefault(mod + ".bias", ini[n["input"][2]])
        elif op == "Conv" and n["input"][1] in ini and _canonical(n["input"][1]) is None:
            mod = _scope_path(n["output"][0]) or _scope_path("/" + n["name"])
            if mod is None and last_conv_prefix and n["attr"].get("group", 1) > 1:
                mod = last_conv_prefix + "dw_conv_1d"        # the depthwise conv that follows this block's pw_conv_1
            if mod is None:
                raise ValueError(f"{path}: cannot name the weight of Conv node {n['name'] or n['output'][0]}")
            sd[mod + ".weight"] = ini[n["input"][1]]
            if len(n["input"]) > 2 and n["input"][2] in ini:
                sd[mod + ".bias"] = ini[n["input"][2]]
            if mod.endswith("dw_conv_1d"):
                folded_bn.append(mod[: -len("dw_conv_1d")] + "BN")
        elif op == "LSTM":
            mod = _scope_path(n["output"][0]) or _scope_path("/" + n["name"]) or f"blstm.encoders.{n_lstm}.lstm"
            n_lstm += 1
            W, R = ini[n["input"][1]], ini[n["input"][2]]
            Bv = ini[n["input"][3]] if len(n["input"]) > 3 and n["input"][3] in ini else np.zeros((W.shape[0], 2 * W.shape[1]), W.dtype)
            H = R.shape[-1]
            for d, suffix in enumerate(("", "_reverse")[: W.shape[0]]):
                sd[f"{mod}.weight_ih_l0{suffix}"] = _lstm_reorder(W[d], H)
                sd[f"{mod}.weight_hh_l0{suffix}"] = _lstm_reorder(R[d], H)
                sd[f"{mod}.bias_ih_l0{suffix}"] = _lstm_reorder(Bv[d, : 4 * H], H)
                sd[f"{mod}.bias_hh_l0{suffix}"] = _lstm_reorder(Bv[d, 4 * H:], H)
    for bn in folded_bn:
        if bn + ".weight" in sd:
            continue                            # the graph kept a BatchNormalization node with named statistics
        c = sd[bn[: -len("BN")] + "dw_conv_1d.weight"].shape[0]
        sd[bn + ".weight"], sd[bn + ".bias"] = np.ones(c, np.float32), np.zeros(c, np.float32)
        sd[bn + ".running_mean"] = np.zeros(c, np.float32)
        sd[bn + ".running_var"] = np.full(c, 1.0 - 1e-5, np.float32)     # sqrt(var + eps) == 1: identity
    out = {k: torch.from_numpy(np.array(v, dtype=np.float32, copy=True)) for k, v in sd.items()}
    for k in [k for k in out if k.endswith(".running_mean")]:
        out[k[: -len("running_mean")] + "num_batches_tracked"] = torch.zeros((), dtype=torch.long)
    return _infer_conf(out, path) + (out,)


def _infer_conf(sd, path):
    nout = sd["linear.weight"].shape[0]
    num_bins = 257 if nout % 257 == 0 else nout // 3
    base = dict(num_bins=num_bins, num_spk=nout // num_bins - 1, num_noise=1)
    if any(k.startswith("conformer.") for k in sd):
        d, idim = sd["conformer.embed.0.weight"].shape
        blocks = 1 + max(int(k.split(".")[2]) for k in sd if k.startswith("conformer.encoders."))
        rel = "conformer.pos_emb.pe_k.weight" in sd
        heads = d // sd["conformer.pos_emb.pe_k.weight"].shape[1] if rel else 4
        conf = dict(base, idim=idim, conformer_attention_dim=d, conformer_attention_heads=heads,
                    conformer_linear_units=sd["conformer.encoders.0.feed_forward_in.net.0.weight"].shape[0],
                    conformer_num_blocks=blocks, conformer_kernel_size=sd["conformer.encoders.0.conv.dw_conv_1d.weight"].shape[-1],
                    conformer_dropout_rate=0.1, conformer_relative_pos_emb=rel)
        return "Conformer", conf
    if any(k.startswith("blstm.") for k in sd):
        h, idim = sd["blstm.embed.0.weight"].shape
        layers = 1 + max(int(k.split(".")[2]) for k in sd if k.startswith("blstm.encoders."))
        return "BLSTM", dict(base, idim=idim, blstm_hdim=h, blstm_num_layers=layers, blstm_dropout_rate=0.1)
    raise ValueError(f"{path}: neither a css Conformer nor a css BLSTM graph (no conformer.* / blstm.* parameters found)")


def load_onnx_model(path):
    """-> (model on CPU in eval mode, model_name, conf).  ``load_state_dict`` is strict: a graph that does not carry every
    parameter of the architecture is rejected."""
    from ..models import build_model

    name, conf, sd = state_dict_from_onnx(path)
    model = build_model(name, conf)
    model.load_state_dict(sd, strict=True)
    return model.eval(), name, conf
