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


def train_conformer(num_blocks=2, idim=1799, seed=3):
    """Case-E model: reference init under `seed`, dropout 0 (the native step's supported setting), train mode."""
    import css_b200.models as M

    conf = dict(CONF_C7, conformer_num_blocks=num_blocks, idim=idim, conformer_dropout_rate=0.0)
    torch.manual_seed(seed)
    return M.build_model("Conformer", conf).train()
