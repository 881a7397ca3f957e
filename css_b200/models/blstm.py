"""BLSTM mask estimator, B200-native forward.

Drop-in for ``css.models.blstm.BLSTM`` (reference css/models/blstm.py:17-88): same constructor,
``add_args``/``build_model``, ``state_dict()`` names and default initialisation order.  The torch
modules are parameter containers; ``forward`` runs ``css_blstm_forward`` from libcss_b200.so
(input-projection GEMM + one fused recurrence kernel per time step).  Inference only.
"""
import ctypes

import torch
from torch import nn

from .. import _native as nv
from .conformer import pad_cols

DEFAULT_BLSTM_CONF = {"hidden_dim": 512, "num_layers": 4, "dropout_rate": 0.1}


def _layer_params(h_dim, p):
    m = nn.Module()
    m.lstm = nn.LSTM(input_size=h_dim, hidden_size=h_dim // 2, num_layers=1, batch_first=True, bidirectional=True)
    m.layer_norm = nn.LayerNorm(h_dim)
    m.dropout = nn.Dropout(p, inplace=True)
    return m


def _encoder_params(idim, hidden_dim, num_layers, dropout_rate):
    enc = nn.Module()
    enc.embed = nn.Sequential(nn.Linear(idim, hidden_dim), nn.LayerNorm(hidden_dim), nn.Dropout(dropout_rate), nn.ReLU())
    enc.encoders = nn.Sequential(*[_layer_params(hidden_dim, dropout_rate) for _ in range(num_layers)])
    return enc


class BLSTM(nn.Module):
    """BLSTM mask estimator (reference: css/models/blstm.py:17-88)."""

    @staticmethod
    def add_args(parser):
        parser.add_argument("--idim", type=int, default=257)
        parser.add_argument("--num-bins", type=int, default=257)
        parser.add_argument("--num-spk", type=int, default=2)
        parser.add_argument("--num-noise", type=int, default=1)
        parser.add_argument("--blstm-hdim", type=int, default=1024)
        parser.add_argument("--blstm-num-layers", type=int, default=3)
        parser.add_argument("--blstm-dropout-rate", type=float, default=0.1)

    @classmethod
    def build_model(cls, conf):
        bc = {"hidden_dim": int(conf["blstm_hdim"]), "num_layers": int(conf["blstm_num_layers"]),
              "dropout_rate": float(conf["blstm_dropout_rate"])}
        return cls(conf["idim"], conf["num_bins"], conf["num_spk"], conf["num_noise"], bc)

    def __init__(self, in_features=257, num_bins=257, num_spk=2, num_noise=1, blstm_conf=DEFAULT_BLSTM_CONF):
        super().__init__()
        self.blstm = _encoder_params(in_features, **blstm_conf)
        self.in_features = in_features
        self.num_bins = num_bins
        self.num_spk = num_spk
        self.num_noise = num_noise
        self.conf = dict(blstm_conf)
        self.linear = nn.Linear(blstm_conf["hidden_dim"], num_bins * (num_spk + num_noise))
        self._engine = None

    def invalidate(self):
        if self._engine is not None:
            nv.lib().css_blstm_destroy(ctypes.c_void_p(self._engine["handle"]))
        self._engine = None

    def load_state_dict(self, *a, **k):
        self.invalidate()
        return super().load_state_dict(*a, **k)

    def _apply(self, fn, *a, **k):
        self.invalidate()
        return super()._apply(fn, *a, **k)

    def __del__(self):
        try:
            self.invalidate()
        except Exception:
            pass

    def _build_engine(self):
        dev = self.linear.weight.device
        if dev.type != "cuda":
            raise nv.NativeError("BLSTM weights must be on a CUDA device (model.cuda()): no CPU fallback")
        keep = []

        def f32(t):
            t = t.detach().to(dev, torch.float32).contiguous()
            keep.append(t)
            return t.data_ptr()

        enc = self.blstm
        h = self.conf["hidden_dim"]
        ldin = (self.in_features + 15) // 16 * 16
        desc = nv.BlstmDesc()
        desc.idim, desc.ldin, desc.hdim, desc.num_layers = self.in_features, ldin, h, len(enc.encoders)
        desc.num_bins, desc.num_out = self.num_bins, self.num_spk + self.num_noise
        desc.embed_w = f32(pad_cols(enc.embed[0].weight.detach(), ldin))
        desc.embed_b, desc.embed_ln_g, desc.embed_ln_b = f32(enc.embed[0].bias), f32(enc.embed[1].weight), f32(enc.embed[1].bias)
        desc.head_w, desc.head_b = f32(self.linear.weight), f32(self.linear.bias)
        layers = (nv.BlstmLayer * len(enc.encoders))()
        for i, lay in enumerate(enc.encoders):
            r = lay.lstm
            layers[i].w_ih = f32(torch.cat([r.weight_ih_l0, r.weight_ih_l0_reverse], 0))
            layers[i].b_gates = f32(torch.cat([r.bias_ih_l0 + r.bias_hh_l0, r.bias_ih_l0_reverse + r.bias_hh_l0_reverse], 0))
            layers[i].w_hh_f, layers[i].w_hh_b = f32(r.weight_hh_l0), f32(r.weight_hh_l0_reverse)
            layers[i].ln_g, layers[i].ln_b = f32(lay.layer_norm.weight), f32(lay.layer_norm.bias)
        handle = nv.lib().css_blstm_create(ctypes.byref(desc), layers)
        if not handle:
            raise nv.NativeError(nv.lib().css_last_error().decode())
        self._engine = {"handle": handle, "keep": keep, "ldin": ldin, "ws": None, "device": dev}

    def engine(self):
        if self._engine is None:
            self._build_engine()
        return self._engine

    def run_padded(self, feat, N, T, apply_mvn=True, clamp_max=0.0, masks=None, head=None):
        eng = self.engine()
        need = nv.lib().css_blstm_workspace_bytes(ctypes.c_void_p(eng["handle"]), nv.c_long(N), nv.c_int(T))
        if eng["ws"] is None or eng["ws"].numel() < need:
            eng["ws"] = None
            eng["ws"] = torch.empty(need, dtype=torch.uint8, device=eng["device"])
        ws = eng["ws"]
        ld_head = 0 if head is None else head.stride(0)
        nv.call("css_blstm_forward", ctypes.c_void_p(eng["handle"]), nv.ptr(feat), nv.c_long(N), nv.c_int(T),
                nv.c_int(1 if apply_mvn else 0), nv.c_float(clamp_max), nv.ptr(masks), nv.ptr(head), nv.c_int(ld_head),
                nv.ptr(ws), nv.c_size_t(ws.numel()))

    def masks_layout(self, feat, N, T, clamp_max=1.0):
        out = torch.empty(N, self.num_bins, T, self.num_spk + self.num_noise, dtype=torch.float32, device=feat.device)
        self.run_padded(feat, N, T, True, clamp_max, masks=out)
        return out

    def forward(self, f):
        """css/models/blstm.py:68-88 with defect D3 patched (``f_orig[..., :num_bins]``)."""
        nv.require_cuda(f, "f")
        if f.ndim == 4:
            f = f.squeeze(0)
        N, T, F = f.shape
        if F != self.in_features:
            raise RuntimeError(f"expected {self.in_features} input features, got {F}")
        eng = self.engine()
        x = torch.zeros(N * T, eng["ldin"], dtype=torch.float32, device=f.device)
        x[:, :F] = f.reshape(N * T, F)
        nout = self.num_spk + self.num_noise
        ldh = (self.num_bins * nout + 7) // 8 * 8
        head = torch.empty(N * T, ldh, dtype=torch.float32, device=f.device)
        self.run_padded(x, N, T, True, 0.0, head=head)
        masks = tuple(head[:, c * self.num_bins : (c + 1) * self.num_bins].reshape(N, T, self.num_bins) for c in range(nout))
        f_orig = f[..., : self.num_bins].detach()
        y_pred = torch.stack([m * f_orig for m in masks[: self.num_spk]], dim=1)
        return y_pred, masks
