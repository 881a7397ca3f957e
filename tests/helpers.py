"""Shared builders for the parity tests (models/configs identical to tests/golden/make_golden.py)."""
import copy

import torch

from css_b200.pipeline import CONFIG_1CH, CONFIG_7CH

CONF_C7 = dict(idim=1799, num_bins=257, num_spk=2, num_noise=1, conformer_attention_dim=256, conformer_attention_heads=4,
               conformer_linear_units=1024, conformer_num_blocks=2, conformer_kernel_size=33, conformer_dropout_rate=0.1,
               conformer_relative_pos_emb=True)
CONF_B1 = dict(idim=257, num_bins=257, num_spk=2, num_noise=1, blstm_hdim=256, blstm_num_layers=2, blstm_dropout_rate=0.1)


def golden_conformer(num_blocks=2, idim=1799, seed=0, bn_seed=7):
    """The case-A model: seeded reference init + non-trivial BatchNorm running stats."""
    import css_b200.models as M

    conf = dict(CONF_C7, conformer_num_blocks=num_blocks, idim=idim)
    torch.manual_seed(seed)
    m = M.build_model("Conformer", conf).eval()
    g = torch.Generator().manual_seed(bn_seed)
    for blk in m.conformer.encoders:
        blk.conv.BN.running_mean.copy_(torch.randn(256, generator=g) * 0.1)
        blk.conv.BN.running_var.copy_(torch.rand(256, generator=g) + 0.5)
    return m


def golden_blstm(seed=1, conf=None):
    import css_b200.models as M

    torch.manual_seed(seed)
    return M.build_model("BLSTM", conf or CONF_B1).eval()


def cpu_state_dict(model):
    return {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}


def cfg7():
    return copy.deepcopy(CONFIG_7CH)


def cfg1():
    c = copy.deepcopy(CONFIG_1CH)
    c["separation"].update(merge=False, merge_threshold=16)
    return c


def rel_inf(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def rel_l2(a, b):
    a, b = torch.as_tensor(a).double().cpu(), torch.as_tensor(b).double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def train_samples(steps, n, t, seed=21, idim=1799):
    """Seeded synthetic training batches shaped like the reference's (mix feats, 2 source magnitudes)."""
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(steps):
        s1 = torch.rand(n, t, 257, generator=g) ** 2
        s2 = torch.rand(n, t, 257, generator=g) ** 2
        mix = torch.empty(n, t, idim)
        mix[..., :257] = s1 + s2 + 0.05 * torch.rand(n, t, 257, generator=g)
        if idim > 257:
            mix[..., 257:] = (torch.rand(n, t, idim - 257, generator=g) * 2 - 1) * 3.14159
        out.append({"mix": mix, "source1": s1, "source2": s2, "ovl": 0.5})
    return out


def train_conformer(num_blocks=2, idim=1799, seed=3, dropout=0.0):
    """Case-E model: reference init under `seed`, train mode (dropout 0 unless asked: the reference-made fixtures use 0)."""
    import css_b200.models as M

    conf = dict(CONF_C7, conformer_num_blocks=num_blocks, idim=idim, conformer_dropout_rate=dropout)
    torch.manual_seed(seed)
    return M.build_model("Conformer", conf).train()


F_FLIP_SEED = 31


def flip_set(n_chunks, seed=F_FLIP_SEED):
    g = torch.Generator().manual_seed(seed)
    return (torch.rand(n_chunks, generator=g) < 0.4).tolist()


def doa_case(seed=77):
    """Constructed spectra X (4,7,257,150) c64 and masks (4,257,150,2) for case G (shared with the GPU test)."""
    from oracle import css_oracle as O

    sv, angles = O.steervec_7ch(257, 30)
    g = torch.Generator().manual_seed(seed)
    B, F, T = 4, 257, 150
    pairs = [(3, 8), (10, 11), (20, 22), (5, 5)]            # steering-grid indices: 60, 12, 24 degrees apart; single
    X = torch.zeros(B, 7, F, T, dtype=torch.complex64)
    mask = torch.zeros(B, F, T, 2)
    for b, (a0, a1) in enumerate(pairs):
        own = torch.rand(F, T, generator=g) < 0.5            # TF bins of source 0; the rest belong to source 1
        s0 = torch.randn(F, T, generator=g) + 1j * torch.randn(F, T, generator=g)
        s1 = torch.randn(F, T, generator=g) + 1j * torch.randn(F, T, generator=g)
        if b == 3:
            # one source, no noise: X is exactly proportional to a steering vector -> radicand ~ 0 +- rounding
            X[b] = (sv[:, a0, :].t().unsqueeze(-1) * s0.unsqueeze(0)).to(torch.complex64)
            mask[b, :, :, 0] = own.float() * 0.9
            mask[b, :, :, 1] = (~own).float() * 0.8
            continue
        d0 = sv[:, a0, :].t().unsqueeze(-1) * (s0 * own).unsqueeze(0)
        d1 = sv[:, a1, :].t().unsqueeze(-1) * (s1 * ~own).unsqueeze(0)
        noise = 0.05 * (torch.randn(7, F, T, generator=g) + 1j * torch.randn(7, F, T, generator=g))
        X[b] = (d0 + d1 + noise).to(torch.complex64)
        mask[b, :, :, 0] = torch.where(own, 0.9, 0.1) + 0.05 * torch.rand(F, T, generator=g)
        mask[b, :, :, 1] = torch.where(own, 0.1, 0.9) - 0.05 * torch.rand(F, T, generator=g)
    return X, mask, pairs


# ---------------------------------------------------------------------------------------------------------------
# numpy mirror of the counter-based dropout masks of css_b200/csrc/dropout.cu (csrc/dropout.h: css_dropout_mix / keep)
# ---------------------------------------------------------------------------------------------------------------
def _mix32(x):
    import numpy as np

    x = x.astype(np.uint32)
    x ^= x >> np.uint32(16)
    x *= np.uint32(0x7FEB352D)
    x ^= x >> np.uint32(15)
    x *= np.uint32(0x846CA68B)
    x ^= x >> np.uint32(16)
    return x


def dropout_mask(n, p, seed, site):
    """keep-mask (bool, length n) of (p, seed, site): keep = (mix(index ^ mix(seed * 0x9E3779B1 + site)) >> 8) >= floor(p 2^24)."""
    import numpy as np

    with np.errstate(over="ignore"):
        key = _mix32(np.array([(seed * 0x9E3779B1 + site) & 0xFFFFFFFF], dtype=np.uint64).astype(np.uint32))[0]
        h = _mix32(np.arange(n, dtype=np.uint32) ^ key)
    return torch.from_numpy((h >> np.uint32(8)) >= np.uint32(int(p * 16777216.0)))


class DropHook:
    """The dropout hook of oracle.css_oracle.conformer_masks with the masks the CUDA kernels use for (p, seed)."""

    def __init__(self, p, seed):
        self.p, self.seed = p, seed

    def __call__(self, x, block, k):
        if self.p <= 0:
            return x
        site = 0 if block < 0 else 1 + 8 * block + k
        scale = 1.0 / (1.0 - self.p)
        if k == 2 and block >= 0:                   # probabilities (n, H, t, t): kernel buffer is [H][n][t][t rounded up to 32]
            n, h, t, _ = x.shape
            tp = (t + 31) // 32 * 32
            m = dropout_mask(h * n * t * tp, self.p, self.seed, site).view(h, n, t, tp)[..., :t].permute(1, 0, 2, 3)
        else:
            m = dropout_mask(x.numel(), self.p, self.seed, site).view(x.shape)
        return x * m.to(x.dtype) * scale


CONF_CTN = dict(num_spk=2, num_noise=1, conv_tasnet_num_filters=64, conv_tasnet_filter_length=16, conv_tasnet_bottleneck_channels=32,
                conv_tasnet_conv_channels=64, conv_tasnet_kernel_size=3, conv_tasnet_num_blocks=3, conv_tasnet_num_layers=2,
                conv_tasnet_norm="gln")


def randomise_conv_tasnet(model, seed=6):
    """Non-trivial PReLU slopes / norm affines (their defaults 0.25 / 1 / 0 would hide indexing mistakes)."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if "PReLU" in name or "norm_" in name or name.startswith("LayerN_S"):
                p.add_(0.2 * torch.randn(p.shape, generator=g))
    return model


def conv_tasnet_input(seed=7, n=2, length=4000):
    return torch.randn(n, length, generator=torch.Generator().manual_seed(seed)) * 0.3
