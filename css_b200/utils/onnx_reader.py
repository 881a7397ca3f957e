"""Minimal ONNX reader: protobuf wire format -> initialisers and graph nodes (no ``onnx`` / ``onnxruntime`` needed).

The reference's only inference backend is an ONNX session (css/utils/model_util.py:12-25, model named by
conf/config_*.yaml:11 ``model_path: checkpoints/*.onnx``).  This module reads such a file far enough to recover the
network's weights: ``ModelProto.graph`` (field 7) -> ``GraphProto.node`` (1), ``.initializer`` (5), ``.input`` (11),
``.output`` (12); ``TensorProto`` dims (1), data_type (2), float_data (4), int32_data (5), int64_data (7), name (8),
raw_data (9), double_data (10); ``NodeProto`` input (1), output (2), name (3), op_type (4), attribute (5).
Field numbers are those of onnx.proto3 (ONNX IR, stable since IR version 3).
"""
import struct

import numpy as np

_DTYPES = {1: np.float32, 2: np.uint8, 3: np.int8, 5: np.int16, 6: np.int32, 7: np.int64, 9: np.bool_, 10: np.float16,
           11: np.float64, 12: np.uint32, 13: np.uint64}


def _varint(buf, pos):
    out = shift = 0
    while True:
        b = buf[pos]
        pos += 1
        out |= (b & 0x7F) << shift
        if not b & 0x80:
            return out, pos
        shift += 7
        if shift > 70:
            raise ValueError("malformed varint")


def fields(buf):
    """Yield (field_number, wire_type, value) of one protobuf message; length-delimited values are memoryviews."""
    buf = memoryview(buf)
    pos, end = 0, len(buf)
    while pos < end:
        key, pos = _varint(buf, pos)
        num, wt = key >> 3, key & 7
        if wt == 0:
            val, pos = _varint(buf, pos)
        elif wt == 1:
            val, pos = bytes(buf[pos:pos + 8]), pos + 8
        elif wt == 2:
            n, pos = _varint(buf, pos)
            if pos + n > end:
                raise ValueError("truncated protobuf message")
            val, pos = buf[pos:pos + n], pos + n
        elif wt == 5:
            val, pos = bytes(buf[pos:pos + 4]), pos + 4
        else:
            raise ValueError(f"unsupported protobuf wire type {wt}")
        yield num, wt, val


def _packed_varints(val):
    out, pos = [], 0
    while pos < len(val):
        v, pos = _varint(val, pos)
        out.append(v)
    return out


def _signed(v):
    return v - (1 << 64) if v >= (1 << 63) else v


def parse_tensor(buf):
    """TensorProto -> (name, numpy array)."""
    dims, dtype, name, raw = [], 1, "", None
    floats, ints, doubles = [], [], []
    for num, wt, val in fields(buf):
        if num == 1:
            dims += [_signed(v) for v in (_packed_varints(val) if wt == 2 else [val])]
        elif num == 2:
            dtype = val
        elif num == 4:
            floats.append(np.frombuffer(val, dtype="<f4") if wt == 2 else np.frombuffer(val, dtype="<f4"))
        elif num in (5, 7):
            ints += [_signed(v) for v in (_packed_varints(val) if wt == 2 else [val])]
        elif num == 8:
            name = bytes(val).decode("utf-8")
        elif num == 9:
            raw = bytes(val)
        elif num == 10:
            doubles.append(np.frombuffer(val, dtype="<f8"))
        elif num == 13 or num == 14:
            if num == 14 and val == 1:
                raise ValueError(f"initializer '{name}' uses external data, which is not supported")
    if dtype not in _DTYPES:
        raise ValueError(f"initializer '{name}': unsupported ONNX data type {dtype}")
    np_dtype = np.dtype(_DTYPES[dtype])
    if raw is not None:
        arr = np.frombuffer(raw, dtype=np_dtype.newbyteorder("<"))
    elif floats:
        arr = np.concatenate(floats).astype(np_dtype)
    elif doubles:
        arr = np.concatenate(doubles).astype(np_dtype)
    else:
        arr = np.array(ints, dtype=np_dtype)
    return name, arr.reshape(dims) if dims else arr.reshape(())


def _parse_attribute(buf):
    name, out = "", {}
    for num, wt, val in fields(buf):
        if num == 1:
            name = bytes(val).decode()
        elif num == 2:
            out["f"] = struct.unpack("<f", val)[0]
        elif num == 3:
            out["i"] = _signed(val)
        elif num == 4:
            out["s"] = bytes(val)
        elif num == 5:
            out["t"] = parse_tensor(val)[1]
        elif num == 7:
            out.setdefault("floats", []).extend(np.frombuffer(val, dtype="<f4").tolist() if wt == 2 else [struct.unpack("<f", val)[0]])
        elif num == 8:
            out.setdefault("ints", []).extend([_signed(v) for v in (_packed_varints(val) if wt == 2 else [val])])
    for k in ("t", "f", "i", "s", "floats", "ints"):
        if k in out:
            return name, out[k]
    return name, None


def _parse_node(buf):
    node = {"input": [], "output": [], "name": "", "op_type": "", "attr": {}}
    for num, wt, val in fields(buf):
        if num == 1:
            node["input"].append(bytes(val).decode())
        elif num == 2:
            node["output"].append(bytes(val).decode())
        elif num == 3:
            node["name"] = bytes(val).decode()
        elif num == 4:
            node["op_type"] = bytes(val).decode()
        elif num == 5:
            k, v = _parse_attribute(val)
            node["attr"][k] = v
    return node


def _value_info_name(buf):
    for num, wt, val in fields(buf):
        if num == 1:
            return bytes(val).decode()
    return ""


def read_onnx(path):
    """-> {"initializers": {name: ndarray} (graph order), "nodes": [...], "inputs": [...], "outputs": [...]}."""
    with open(path, "rb") as f:
        data = f.read()
    graph = None
    for num, wt, val in fields(data):
        if num == 7 and wt == 2:
            graph = val
    if graph is None:
        raise ValueError(f"{path}: no GraphProto found (not an ONNX ModelProto?)")
    out = {"initializers": {}, "nodes": [], "inputs": [], "outputs": []}
    for num, wt, val in fields(graph):
        if num == 1:
            out["nodes"].append(_parse_node(val))
        elif num == 5:
            name, arr = parse_tensor(val)
            out["initializers"][name] = arr
        elif num == 11:
            out["inputs"].append(_value_info_name(val))
        elif num == 12:
            out["outputs"].append(_value_info_name(val))
    # Constant nodes carry tensors too (exporters fold some weights into them)
    for n in out["nodes"]:
        if n["op_type"] == "Constant" and isinstance(n["attr"].get("value"), np.ndarray):
            out["initializers"].setdefault(n["output"][0], n["attr"]["value"])
    out["inputs"] = [i for i in out["inputs"] if i not in out["initializers"]]
    return out
