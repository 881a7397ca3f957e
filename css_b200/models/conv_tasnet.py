"""Conv-TasNet, B200-native forward (reference: css/models/conv_tasnet.py:17-330, itself taken from JusperLee/Conv-TasNet).

Drop-in for ``css.models.conv_tasnet.ConvTasNet``: same constructor / ``add_args`` / ``build_model`` and ``state_dict()`` names
(encoder, LayerN_S, BottleN_S, separation.R.X.{conv1x1, PReLU_1, norm_1, dwconv, PReLU_2, norm_2, Sc_conv}, gen_masks, decoder),
same default initialisation order.  The torch modules are parameter containers; ``forward`` keeps the activations time-major on
the GPU, runs every 1x1 convolution as ``css_gemm_f32`` and the rest through the ``css_ctn_*`` kernels of libcss_b200.so.
Inference only (the reference trains its mask models with the STFT-domain PIT-MSE objective; this waveform-domain model is
outside that training path).  Separation-block norm: "gln" (the argparse default); the first norm is always "cln" as in the
reference.
"""
import torch
from torch import nn

from .. import _native as nv

DEFAULT_CONV_TASNET_CONF = {
    "num_filters": 512,
    "filter_length": 16,
    "bottleneck_channels": 128,
    "conv_channels": 512,
    "kernel_size": 3,
    "num_blocks": 8,
    "num_layers": 3,
}


class _GlobalLayerNorm(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim, 1))
        self.bias = nn.Parameter(torch.zeros(dim, 1))


def _block_params(in_channels, out_channels, kernel_size, dilation):
    m = nn.Module()
    m.conv1x1 = nn.Conv1d(in_channels, out_channels, 1)
    m.PReLU_1 = nn.PReLU()
    m.norm_1 = _GlobalLayerNorm(out_channels)
    m.dwconv = nn.Conv1d(out_channels, out_channels, kernel_size, groups=out_channels, padding=dilation * (kernel_size - 1) // 2,
                         dilation=dilation)
    m.PReLU_2 = nn.PReLU()
    m.norm_2 = _GlobalLayerNorm(out_channels)
    m.Sc_conv = nn.Conv1d(out_channels, in_channels, 1, bias=True)
    m.dilation = dilation
    return m


class ConvTasNet(nn.Module):
    @staticmethod
    def add_args(parser):
        parser.add_argument("--num-spk", type=int, default=2)
        parser.add_argument("--num-noise", type=int, default=1)
        parser.add_argument("--conv-tasnet-num-filters", type=int, default=256)
        parser.add_argument("--conv-tasnet-filter-length", type=int, default=16)
        parser.add_argument("--conv-tasnet-bottleneck-channels", type=int, default=128)
        parser.add_argument("--conv-tasnet-conv-channels", type=int, default=256)
        parser.add_argument("--conv-tasnet-kernel-size", type=int, default=3)
        parser.add_argument("--conv-tasnet-num-blocks", type=int, default=8)
        parser.add_argument("--conv-tasnet-num-layers", type=int, default=3)
        parser.add_argument("--conv-tasnet-norm", type=str, default="gln", choices=["gln", "cln", "bn"])

    @classmethod
    def build_model(cls, conf):
        keys = ("num_filters", "filter_length", "bottleneck_channels", "conv_channels", "kernel_size", "num_blocks", "num_layers")
        cc = {k: int(conf["conv_tasnet_" + k]) for k in keys}
        cc["norm"] = conf["conv_tasnet_norm"]
        return cls(num_spk=conf["num_spk"], num_noise=conf["num_noise"], conv_tasnet_conf=cc)

    def __init__(self, num_spk=2, num_noise=1, conv_tasnet_conf=DEFAULT_CONV_TASNET_CONF, activate="relu", causal=False):
        super().__init__()
        c = conv_tasnet_conf
        N, L, B, H = c["num_filters"], c["filter_length"], c["bottleneck_channels"], c["conv_channels"]
        P, X, R = c["kernel_size"], c["num_blocks"], c["num_layers"]
        norm = c["norm"]          # KeyError for the defaults dict, as in the reference
        if norm != "gln":
            raise ValueError(f"Conv-TasNet separation-block norm '{norm}' is not built (the sm_100a kernels implement 'gln', the default)")
        if causal or activate != "relu" or P % 2 == 0:
            raise ValueError("the sm_100a Conv-TasNet kernels are non-causal, ReLU masks, odd kernel size")
        if N % 16 or B % 16 or H % 16:
            raise ValueError("num_filters / bottleneck_channels / conv_channels must be multiples of 16 (GEMM K granularity)")
        self.encoder = nn.Conv1d(1, N, L, stride=L // 2, padding=0)
        self.LayerN_S = nn.LayerNorm(N)
        self.BottleN_S = nn.Conv1d(N, B, 1)
        self.separation = nn.Sequential(*[nn.Sequential(*[_block_params(B, H, P, 2 ** i) for i in range(X)]) for _ in range(R)])
        self.gen_masks = nn.Conv1d(B, (num_spk + num_noise) * N, 1)
        self.decoder = nn.ConvTranspose1d(N, 1, L, stride=L // 2)
        self.num_spk, self.num_noise = num_spk, num_noise
        self.dims = (N, L, B, H, P)

    def forward(self, x):
        """x: (L,) or (n, L) waveform on the GPU -> (n, num_spk, L') separated waveforms (conv_tasnet.py:127-156)."""
        if x.dim() >= 3:
            raise RuntimeError("ConvTasNet accept 1/2D tensor as input, but got {:d}".format(x.dim()))
        nv.require_cuda(x, "x")
        if x.dim() == 1:
            x = x.unsqueeze(0)
        x = x.float().contiguous()
        n, length = x.shape
        N, L, B, H, P = self.dims
        stride = L // 2
        if length < L:
            raise RuntimeError(f"waveform of {length} samples is shorter than the {L}-sample encoder filter")
        T = (length - L) // stride + 1
        M = n * T
        dev = x.device

        def f32(t):
            return t.detach().to(dev, torch.float32).contiguous()

        def gemm(a, conv, out_ch, act=0, residual=None):
            w = f32(conv.weight).reshape(out_ch, -1)
            out = torch.empty(M, out_ch, dtype=torch.float32, device=dev)
            nv.call("css_gemm_f32", nv.ptr(a), nv.c_int(a.shape[1]), nv.ptr(w), nv.c_int(w.shape[1]), nv.ptr(f32(conv.bias)), nv.ptr(out),
                    nv.c_int(out_ch), nv.c_long(M), nv.c_int(out_ch), nv.c_int(a.shape[1]), nv.c_int(act), nv.c_float(1.0),
                    nv.ptr(residual), nv.c_int(0 if residual is None else residual.shape[1]), nv.c_float(0.0))
            return out

        w = torch.empty(M, N, dtype=torch.float32, device=dev)
        nv.call("css_ctn_encoder", nv.ptr(x), nv.c_long(n), nv.c_long(length), nv.ptr(f32(self.encoder.weight).reshape(N, L)),
                nv.ptr(f32(self.encoder.bias)), nv.c_int(N), nv.c_int(L), nv.c_int(stride), nv.ptr(w))
        e = torch.empty_like(w)
        nv.call("css_row_layernorm", nv.ptr(w), nv.ptr(f32(self.LayerN_S.weight)), nv.ptr(f32(self.LayerN_S.bias)), nv.c_long(M), nv.c_int(N),
                nv.c_float(self.LayerN_S.eps), nv.ptr(e))
        e = gemm(e, self.BottleN_S, B)
        stats = torch.empty(n, 2, dtype=torch.float64, device=dev)
        c2 = torch.empty(M, H, dtype=torch.float32, device=dev)
        c3 = torch.empty(M, H, dtype=torch.float32, device=dev)
        for rep in self.separation:
            for blk in rep:
                c1 = gemm(e, blk.conv1x1, H)
                a1, g1, b1 = f32(blk.PReLU_1.weight), f32(blk.norm_1.weight).reshape(-1), f32(blk.norm_1.bias).reshape(-1)
                nv.call("css_ctn_prelu_stats", nv.ptr(c1), nv.c_long(n), nv.c_long(T * H), nv.ptr(a1), nv.ptr(stats))
                nv.call("css_ctn_dwconv", nv.ptr(c1), nv.c_long(n), nv.c_long(T), nv.c_int(H), nv.c_int(P), nv.c_int(blk.dilation), nv.ptr(a1),
                        nv.ptr(stats), nv.ptr(g1), nv.ptr(b1), nv.c_float(1e-5), nv.ptr(f32(blk.dwconv.weight).reshape(H, P)),
                        nv.ptr(f32(blk.dwconv.bias)), nv.ptr(c2))
                a2, g2, b2 = f32(blk.PReLU_2.weight), f32(blk.norm_2.weight).reshape(-1), f32(blk.norm_2.bias).reshape(-1)
                nv.call("css_ctn_prelu_stats", nv.ptr(c2), nv.c_long(n), nv.c_long(T * H), nv.ptr(a2), nv.ptr(stats))
                nv.call("css_ctn_norm_apply", nv.ptr(c2), nv.c_long(n), nv.c_long(T), nv.c_int(H), nv.ptr(a2), nv.ptr(stats), nv.ptr(g2),
                        nv.ptr(b2), nv.c_float(1e-5), nv.ptr(c3))
                e = gemm(c3, blk.Sc_conv, B, residual=e)
        S = self.num_spk + self.num_noise
        masks = gemm(e, self.gen_masks, S * N, act=1)
        out_len = (T - 1) * stride + L
        frames = torch.empty(self.num_spk * M * L, dtype=torch.float32, device=dev)
        y = torch.empty(self.num_spk, n, out_len, dtype=torch.float32, device=dev)
        nv.call("css_ctn_decode", nv.ptr(w), nv.ptr(masks), nv.c_int(S * N), nv.c_long(n), nv.c_long(T), nv.c_int(N), nv.c_int(L),
                nv.c_int(stride), nv.c_int(self.num_spk), nv.ptr(f32(self.decoder.weight).reshape(N, L)), nv.ptr(f32(self.decoder.bias)),
                nv.ptr(frames), nv.ptr(y))
        return y.permute(1, 0, 2).contiguous()
