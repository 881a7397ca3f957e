"""Conformer mask estimator, B200-native forward.

Drop-in for ``css.models.conformer.Conformer`` (reference css/models/conformer.py:24-104): same
constructor, ``add_args``/``build_model`` and ``state_dict()`` key names (the ``.mdl`` interchange
contract), same default PyTorch initialisation order, so a given
``torch.manual_seed`` yields the reference's weights.  The torch modules below are parameter
containers only: ``forward`` hands the weight pointers to ``css_conformer_forward`` in
libcss_b200.so.  In eval mode dropout is the identity and BatchNorm uses running stats; in train mode ``forward`` runs the native
training step (css_b200/trainer), with batch-statistics BatchNorm, counter-based dropout masks at the reference's six
nn.Dropout sites (csrc/dropout.cu) and a native backward.
"""
import ctypes

import torch
from torch import nn

from .. import _native as nv

DEFAULT_CONFORMER_CONF = {
    "attention_dim": 256,
    "attention_heads": 4,
    "linear_units": 1024,
    "num_blocks": 16,
    "kernel_size": 33,
    "dropout_rate": 0.1,
    "relative_pos_emb": True,
}


def _ffn_params(d_model, d_inner, p):
    m = nn.Module()
    m.layer_norm = nn.LayerNorm(d_model)
    m.net = nn.Sequential(nn.Linear(d_model, d_inner), nn.ReLU(inplace=True), nn.Dropout(p), nn.Linear(d_inner, d_model), nn.Dropout(p))
    return m


def _attn_params(heads, d_model, p):
    m = nn.Module()
    m.layer_norm = nn.LayerNorm(d_model)
    for name in ("linear_q", "linear_k", "linear_v", "linear_out"):
        setattr(m, name, nn.Linear(d_model, d_model))
    m.dropout = nn.Dropout(p)
    return m


def _conv_params(d_model, ksize, p):
    m = nn.Module()
    m.layer_norm = nn.LayerNorm(d_model)
    m.pw_conv_1 = nn.Conv2d(1, 2, 1, 1, 0)
    m.dw_conv_1d = nn.Conv1d(d_model, d_model, ksize, 1, padding=(ksize - 1) // 2, groups=d_model)
    m.BN = nn.BatchNorm1d(d_model)
    m.pw_conv_2 = nn.Conv2d(1, 1, 1, 1, 0)
    m.dropout = nn.Dropout(p)
    return m


def _block_params(d_model, heads, d_ffn, ksize, p):
    m = nn.Module()
    m.feed_forward_in = _ffn_params(d_model, d_ffn, p)
    m.self_attn = _attn_params(heads, d_model, p)
    m.conv = _conv_params(d_model, ksize, p)
    m.feed_forward_out = _ffn_params(d_model, d_ffn, p)
    m.layer_norm = nn.LayerNorm(d_model)
    return m


def _encoder_params(idim, attention_dim, attention_heads, linear_units, num_blocks, kernel_size, dropout_rate, relative_pos_emb):
    enc = nn.Module()
    enc.embed = nn.Sequential(nn.Linear(idim, attention_dim), nn.LayerNorm(attention_dim), nn.Dropout(dropout_rate), nn.ReLU())
    if relative_pos_emb:
        enc.pos_emb = nn.Module()
        enc.pos_emb.pe_k = nn.Embedding(2 * 1000, attention_dim // attention_heads)
    else:
        enc.pos_emb = None
    enc.encoders = nn.Sequential(
        *[_block_params(attention_dim, attention_heads, linear_units, kernel_size, dropout_rate) for _ in range(num_blocks)]
    )
    return enc


def pad_cols(w, ld):
    """(rows, k) -> zero-padded contiguous (rows, ld)."""
    out = w.new_zeros(w.shape[0], ld)
    out[:, : w.shape[1]] = w
    return out


class Conformer(nn.Module):
    """Conformer mask estimator (reference: css/models/conformer.py:24-104)."""

    @staticmethod
    def add_args(parser):
        parser.add_argument("--idim", type=int, default=257)
        parser.add_argument("--num-bins", type=int, default=257)
        parser.add_argument("--num-spk", type=int, default=2)
        parser.add_argument("--num-noise", type=int, default=1)
        parser.add_argument("--conformer-attention-dim", type=str, default=256)
        parser.add_argument("--conformer-attention-heads", type=int, default=4)
        parser.add_argument("--conformer-linear-units", type=int, default=1024)
        parser.add_argument("--conformer-num-blocks", type=int, default=16)
        parser.add_argument("--conformer-kernel-size", type=int, default=33)
        parser.add_argument("--conformer-dropout-rate", type=float, default=0.1)
        parser.add_argument("--conformer-relative-pos-emb", type=bool, default=True)

    @classmethod
    def build_model(cls, conf):
        keys = ("attention_dim", "attention_heads", "linear_units", "num_blocks", "kernel_size")
        cc = {k: int(conf["conformer_" + k]) for k in keys}
        cc["dropout_rate"] = float(conf["conformer_dropout_rate"])
        cc["relative_pos_emb"] = bool(conf["conformer_relative_pos_emb"])
        return cls(conf["idim"], conf["num_bins"], conf["num_spk"], conf["num_noise"], cc)

    def __init__(self, in_features=257, num_bins=257, num_spk=2, num_noise=1, conformer_conf=DEFAULT_CONFORMER_CONF):
        super().__init__()
        self.conformer = _encoder_params(in_features, **conformer_conf)
        self.in_features = in_features
        self.num_bins = num_bins
        self.num_spk = num_spk
        self.num_noise = num_noise
        self.conf = dict(conformer_conf)
        self.linear = nn.Linear(conformer_conf["attention_dim"], num_bins * (num_spk + num_noise))
        self._engine = None
        self._train_engine = None

    # ------------------------------------------------------------------ native engine
    def train_engine(self):
        """Flat-buffer training engine (css_b200/trainer/engine.py), built on first use; re-homes the Parameters."""
        if self._train_engine is None:
            from ..trainer.engine import ConformerTrainEngine

            self._train_engine = ConformerTrainEngine(self)
        return self._train_engine

    def invalidate(self):
        """Drop the packed weights (call after loading a state_dict or moving devices)."""
        if self._engine is not None:
            nv.lib().css_conformer_destroy(ctypes.c_void_p(self._engine["handle"]))
        self._engine = None

    def load_state_dict(self, *a, **k):
        self.invalidate()
        return super().load_state_dict(*a, **k)

    def _apply(self, fn, *a, **k):
        self.invalidate()
        self._train_engine = None    # a device/dtype move re-allocates the parameters
        return super()._apply(fn, *a, **k)

    def __del__(self):
        try:
            self.invalidate()
        except Exception:
            pass

    def _build_engine(self):
        dev = self.linear.weight.device
        if dev.type != "cuda":
            raise nv.NativeError("Conformer weights must be on a CUDA device (model.cuda()): no CPU fallback")
        keep = []   # packed tensors referenced by raw pointers

        def f32(t):
            t = t.detach().to(dev, torch.float32).contiguous()
            keep.append(t)
            return t.data_ptr()

        enc = self.conformer
        d = self.conf["attention_dim"]
        ldin = (self.in_features + 63) // 64 * 64
        desc = nv.ConformerDesc()
        desc.idim, desc.ldin, desc.d_model = self.in_features, ldin, d
        desc.heads, desc.d_ffn, desc.num_blocks = self.conf["attention_heads"], self.conf["linear_units"], self.conf["num_blocks"]
        desc.num_bins, desc.num_out = self.num_bins, self.num_spk + self.num_noise
        desc.embed_w = f32(pad_cols(enc.embed[0].weight.detach(), ldin))
        desc.embed_b, desc.embed_ln_g, desc.embed_ln_b = f32(enc.embed[0].bias), f32(enc.embed[1].weight), f32(enc.embed[1].bias)
        if enc.pos_emb is not None:
            desc.pe, desc.pe_rows = f32(enc.pos_emb.pe_k.weight), enc.pos_emb.pe_k.weight.shape[0]
        else:
            desc.pe, desc.pe_rows = None, 0
        desc.head_w, desc.head_b = f32(self.linear.weight), f32(self.linear.bias)
        blocks = (nv.ConformerBlock * len(enc.encoders))()
        scal = []   # host scalars fetched with one sync
        for blk in enc.encoders:
            scal.append(torch.cat([blk.conv.pw_conv_1.weight.detach().reshape(2), blk.conv.pw_conv_1.bias.detach().reshape(2),
                                   blk.conv.pw_conv_2.weight.detach().reshape(1), blk.conv.pw_conv_2.bias.detach().reshape(1)]))
        scal = torch.stack(scal).float().cpu().tolist()
        for i, blk in enumerate(enc.encoders):
            b = blocks[i]
            for dst, src in ((b.ff_in, blk.feed_forward_in), (b.ff_out, blk.feed_forward_out)):
                dst.ln_g, dst.ln_b = f32(src.layer_norm.weight), f32(src.layer_norm.bias)
                dst.w1, dst.b1 = f32(src.net[0].weight), f32(src.net[0].bias)
                dst.w2, dst.b2 = f32(src.net[3].weight), f32(src.net[3].bias)
            at = blk.self_attn
            b.att_ln_g, b.att_ln_b = f32(at.layer_norm.weight), f32(at.layer_norm.bias)
            b.wqkv = f32(torch.cat([at.linear_q.weight, at.linear_k.weight, at.linear_v.weight], 0))
            b.bqkv = f32(torch.cat([at.linear_q.bias, at.linear_k.bias, at.linear_v.bias], 0))
            b.wo, b.bo = f32(at.linear_out.weight), f32(at.linear_out.bias)
            cv = blk.conv
            b.conv_ln_g, b.conv_ln_b = f32(cv.layer_norm.weight), f32(cv.layer_norm.bias)
            taps = cv.dw_conv_1d.weight.detach().reshape(d, -1)
            ks = taps.shape[1]
            if ks > 33 or ks % 2 == 0:
                raise nv.NativeError(f"conformer kernel_size {ks} unsupported (odd, <= 33)")
            padded = taps.new_zeros(d, 33)
            padded[:, (33 - ks) // 2 : (33 - ks) // 2 + ks] = taps
            b.dw_w, b.dw_b = f32(padded), f32(cv.dw_conv_1d.bias)
            scale = cv.BN.weight.detach() / torch.sqrt(cv.BN.running_var + cv.BN.eps)
            b.bn_scale, b.bn_shift = f32(scale), f32(cv.BN.bias.detach() - cv.BN.running_mean * scale)
            b.pw1_w0, b.pw1_w1, b.pw1_b0, b.pw1_b1, b.pw2_w, b.pw2_b = scal[i]
            b.out_ln_g, b.out_ln_b = f32(blk.layer_norm.weight), f32(blk.layer_norm.bias)
        handle = nv.lib().css_conformer_create(ctypes.byref(desc), blocks)
        if not handle:
            raise nv.NativeError(nv.lib().css_last_error().decode())
        self._engine = {"handle": handle, "keep": keep, "ldin": ldin, "ws": None, "device": dev}

    def engine(self):
        if self._engine is None:
            self._build_engine()
        return self._engine

    def _workspace(self, eng, N, T):
        need = nv.lib().css_conformer_workspace_bytes(ctypes.c_void_p(eng["handle"]), nv.c_long(N), nv.c_int(T))
        if eng["ws"] is None or eng["ws"].numel() < need:
            eng["ws"] = None
            eng["ws"] = torch.empty(need, dtype=torch.uint8, device=eng["device"])
        return eng["ws"]

    def run_padded(self, feat, N, T, apply_mvn=True, clamp_max=0.0, masks=None, head=None):
        """feat: (N*T, ldin) fp32 CUDA, zero padded, modified in place when apply_mvn."""
        eng = self.engine()
        ws = self._workspace(eng, N, T)
        ld_head = 0 if head is None else head.stride(0)
        nv.call("css_conformer_forward", ctypes.c_void_p(eng["handle"]), nv.ptr(feat), nv.c_long(N), nv.c_int(T),
                nv.c_int(1 if apply_mvn else 0), nv.c_float(clamp_max), nv.ptr(masks), nv.ptr(head), nv.c_int(ld_head),
                nv.ptr(ws), nv.c_size_t(ws.numel()))

    def masks_from_spec(self, spec, stride_d, stride_chunk, N, T, ipd_l, ipd_r, clamp_max=1.0, shared_right=False):
        """Executor fast path with the fused front end: chunk spectra -> (masks (N, num_bins, T, 3), mag (N, num_bins, T)),
        or None when the shape / GEMM mode is not covered (the caller then uses features() + masks_layout())."""
        n_pairs = int(ipd_l.numel())
        if nv.get_gemm_mode() == 0 or self.in_features != self.num_bins * (n_pairs + 1) or self.num_bins != 257 \
                or not nv.lib().css_features_planes_supported(nv.c_int(T), nv.c_int(n_pairs)):
            return None
        eng = self.engine()
        ws = self._workspace(eng, N, T)
        out = torch.empty(N, self.num_bins, T, self.num_spk + self.num_noise, dtype=torch.float32, device=spec.device)
        mag = torch.empty(N, self.num_bins, T, dtype=torch.float32, device=spec.device)
        nv.call("css_conformer_forward_spec", ctypes.c_void_p(eng["handle"]), nv.ptr(spec), nv.c_long(stride_d),
                nv.c_long(stride_chunk), nv.ptr(ipd_l), nv.ptr(ipd_r), nv.c_int(n_pairs), nv.c_int(1 if shared_right else 0), nv.c_long(N),
                nv.c_int(T), nv.c_float(clamp_max), nv.ptr(mag), nv.ptr(out), nv.ptr(None), nv.c_int(0), nv.ptr(ws), nv.c_size_t(ws.numel()))
        return out, mag

    def masks_layout(self, feat, N, T, clamp_max=1.0):
        """Executor fast path: padded features -> masks (N, num_bins, T, 3) (separator.py:100-104)."""
        out = torch.empty(N, self.num_bins, T, self.num_spk + self.num_noise, dtype=torch.float32, device=feat.device)
        self.run_padded(feat, N, T, True, clamp_max, masks=out)
        return out

    def forward(self, f):
        """f: (N, T, F) or (1, N, T, F) -> (y_pred (N, num_spk, T, num_bins), masks tuple of (N, T, num_bins)).

        css/models/conformer.py:83-104, with defect D3 patched (``f_orig[..., :num_bins]``).
        """
        nv.require_cuda(f, "f")
        if f.ndim == 4:
            f = f.squeeze(0)
        N, T, F = f.shape
        if F != self.in_features:
            raise RuntimeError(f"expected {self.in_features} input features, got {F}")
        if self.training and torch.is_grad_enabled():
            from ..trainer.autograd import train_forward

            return train_forward(self, f)
        eng = self.engine()
        x = torch.zeros(N * T, eng["ldin"], dtype=torch.float32, device=f.device)
        x[:, :F] = f.reshape(N * T, F)
        nout = self.num_spk + self.num_noise
        ldh = (self.num_bins * nout + 7) // 8 * 8
        head = torch.empty(N * T, ldh, dtype=torch.float32, device=f.device)
        self.run_padded(x, N, T, True, 0.0, head=head)
        masks = tuple(head[:, c * self.num_bins : (c + 1) * self.num_bins].reshape(N, T, self.num_bins) for c in range(nout))
        f_orig = f[..., : self.num_bins]
        y_pred = torch.stack([m * f_orig for m in masks[: self.num_spk]], dim=1)
        return y_pred, masks
