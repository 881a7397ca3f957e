"""CPU restatement (torch-CPU, fp32) of the desh2608/css hot path.

TEST INFRASTRUCTURE ONLY -- never imported by the product (``css_b200``).

Parity status: the reference ships NO tests, golden vectors or fixtures
(SURVEY.md section 4 / 8c).  This restatement is therefore pinned against outputs of the
reference itself, executed in the build container by ``tests/golden/make_golden.py``
(which imports ``/root/reference`` with ``onnxruntime``/``asteroid`` stand-ins) and
committed under ``tests/golden/``.  ``tests/test_oracle_golden.py`` replays them.
The asteroid beamformer arithmetic (``asteroid.dsp.beamforming``, version unpinned, not
vendored by the reference) is restated from its published algorithm: that sub-boundary is
"parity unpinned" and cross-checked against ``torchaudio.functional`` instead.

Every function cites the reference file:line it follows (paths relative to the reference
root).  All tensors are torch CPU tensors; arithmetic is fp32 (complex64) except the MVDR
solve (complex128), as in the reference.
"""
import itertools
import math

import torch

EPS = torch.finfo(torch.float32).eps  # css/executor/feature.py:9, css/models/conformer.py:21


# --------------------------------------------------------------------------------------
# STFT / features  (css/executor/feature.py)
# --------------------------------------------------------------------------------------
def dft_kernel(frame_len=512):
    """Windowed DFT basis, (2*(N/2+1), 1, frame_len).  css/executor/feature.py:13-30."""
    n_fft = 2 ** math.ceil(math.log2(frame_len))
    basis = torch.view_as_real(torch.fft.rfft(torch.eye(n_fft), dim=1))[:frame_len]
    basis = basis.permute(2, 1, 0) * torch.hann_window(frame_len)
    return basis.reshape(n_fft + 2, 1, frame_len)


def stft_conv(x, frame_len=512, frame_hop=256, kernel=None):
    """Uncentred conv1d STFT.  css/executor/feature.py:45-82.

    x: (B, N) or (B, D, N) -> mag, phase, re, im of shape (B, [D,] F, T).
    """
    if x.dim() not in (2, 3):
        raise RuntimeError("stft_conv expects a 2D/3D tensor")
    K = dft_kernel(frame_len) if kernel is None else kernel
    if x.dim() == 2:
        c = torch.nn.functional.conv1d(x.unsqueeze(1), K, stride=frame_hop)
        re, im = torch.chunk(c, 2, dim=1)
    else:
        B, D, N = x.shape
        c = torch.nn.functional.conv1d(x.reshape(B * D, 1, N), K, stride=frame_hop)
        c = c.view(B, D, -1, c.shape[-1])
        re, im = torch.chunk(c, 2, dim=2)
    mag = (re ** 2 + im ** 2) ** 0.5
    pha = torch.atan2(im, re)
    return mag, pha, re, im


def parse_ipd(ipd_index):
    """'1,0;2,0' -> ([1,2],[0,0]).  css/executor/feature.py:95-103."""
    pairs = [tuple(int(v) for v in p.split(",")) for p in ipd_index.split(";")]
    return [p[0] for p in pairs], [p[1] for p in pairs]


def ipd_feature(pha, ipd_index):
    """Mean-normalised inter-channel phase difference.  css/executor/feature.py:105-134."""
    if pha.dim() == 3:
        pha = pha.unsqueeze(0)
    left, right = parse_ipd(ipd_index)
    N, _, _, T = pha.shape
    delta = pha[:, left] - pha[:, right]
    c, s = torch.cos(delta), torch.sin(delta)
    out = torch.atan2(s - s.mean(-1, keepdim=True), c - c.mean(-1, keepdim=True))
    return out.view(N, -1, T)


def feature_extract(x, frame_len=512, frame_hop=256, ipd_index=None, kernel=None):
    """mag(ch0), features, re, im.  css/executor/feature.py:156-175."""
    mag, pha, re, im = stft_conv(x, frame_len, frame_hop, kernel)
    if mag.dim() == 4:
        mag = mag[:, 0]
    f = torch.clamp(mag, min=EPS)
    f = (f - f.mean(-1, keepdim=True)) / (f.std(-1, keepdim=True) + EPS)
    if ipd_index:
        f = torch.cat([f, ipd_feature(pha, ipd_index)], dim=1)
    return mag, f, re, im


# --------------------------------------------------------------------------------------
# Mask networks, functional over a reference-named state_dict  (css/models/*.py)
# --------------------------------------------------------------------------------------
def _linear(x, sd, name):
    return torch.nn.functional.linear(x, sd[name + ".weight"], sd[name + ".bias"])


def _layer_norm(x, sd, name):
    w = sd[name + ".weight"]
    return torch.nn.functional.layer_norm(x, (w.shape[0],), w, sd[name + ".bias"], 1e-5)


def mvn_time(f):
    """(f - mean_T) / (std_T(unbiased) + eps) over dim -2.  css/models/conformer.py:95."""
    return (f - f.mean(-2, keepdim=True)) / (f.std(-2, keepdim=True) + EPS)


def _no_drop(x, block, k):
    return x


def _conformer_ffn(x, sd, p, drop=_no_drop, block=0, k=0):
    """css/models/conformer.py:306-326 (Linear, ReLU, Dropout, Linear, Dropout)."""
    y = _layer_norm(x, sd, p + ".layer_norm")
    y = drop(torch.relu(_linear(y, sd, p + ".net.0")), block, k)
    return drop(_linear(y, sd, p + ".net.3"), block, k + 1)


def _conformer_mhsa(x, pos_k, sd, p, heads, drop=_no_drop, block=0):
    """css/models/conformer.py:213-254 (mask=None); dropout on the probabilities (:249) and on the output (:254)."""
    n, t, d = x.shape
    dk = d // heads
    y = _layer_norm(x, sd, p + ".layer_norm")
    q = _linear(y, sd, p + ".linear_q").view(n, t, heads, dk).transpose(1, 2)
    k = _linear(y, sd, p + ".linear_k").view(n, t, heads, dk).transpose(1, 2)
    v = _linear(y, sd, p + ".linear_v").view(n, t, heads, dk).transpose(1, 2)
    a = torch.matmul(q, k.transpose(-2, -1))
    if pos_k is not None:
        # b[n,h,i,j] = q[n,h,i,:] . pos_k[i,j,:]
        qq = q.contiguous().view(n * heads, t, dk).transpose(0, 1)
        b = torch.matmul(qq, pos_k.transpose(-2, -1)).transpose(0, 1).view(n, heads, t, t)
        a = a + b
    attn = drop(torch.softmax(a / math.sqrt(dk), dim=-1), block, 2)
    o = torch.matmul(attn, v).transpose(1, 2).contiguous().view(n, t, d)
    return drop(_linear(o, sd, p + ".linear_out"), block, 3)


def _conformer_conv(x, sd, p, training_stats=False, drop=_no_drop, block=0):
    """css/models/conformer.py:289-303.  The two 'pointwise convs' are scalar affines; dropout at the end (:302)."""
    y = _layer_norm(x, sd, p + ".layer_norm")
    w1, b1 = sd[p + ".pw_conv_1.weight"].view(2), sd[p + ".pw_conv_1.bias"]
    y = (y * w1[0] + b1[0]) * torch.sigmoid(y * w1[1] + b1[1])
    wd = sd[p + ".dw_conv_1d.weight"]
    y = torch.nn.functional.conv1d(
        y.transpose(1, 2), wd, sd[p + ".dw_conv_1d.bias"], padding=(wd.shape[-1] - 1) // 2, groups=wd.shape[0]
    )
    y = torch.nn.functional.batch_norm(
        y, sd[p + ".BN.running_mean"], sd[p + ".BN.running_var"], sd[p + ".BN.weight"], sd[p + ".BN.bias"],
        training_stats, 0.0 if training_stats else 0.1, 1e-5,
    )
    y = torch.relu(y).transpose(1, 2)
    return drop(y * sd[p + ".pw_conv_2.weight"].view(()) + sd[p + ".pw_conv_2.bias"].view(()), block, 4)


def conformer_masks(f, sd, heads=4, num_out=3, training=False, drop=None):
    """Reference Conformer forward up to the masks.  css/models/conformer.py:83-101,152-165.

    f: (N, T, F) -> list of num_out tensors (N, T, num_bins).  ``training``: BatchNorm uses batch statistics
    (model.train()).  ``drop``: callable (x, block, k) -> x applied at the reference's six nn.Dropout sites (block -1 /
    k 0 = embedding; per block k = 0 FFN-in hidden, 1 FFN-in output, 2 attention probabilities (n, heads, t, t),
    3 attention output, 4 conv module, 5 FFN-out hidden, 6 FFN-out output); None = dropout probability 0.
    """
    if f.ndim == 4:
        f = f.squeeze(0)
    drop = drop or _no_drop
    x = mvn_time(f)
    x = torch.relu(drop(_layer_norm(_linear(x, sd, "conformer.embed.0"), sd, "conformer.embed.1"), -1, 0))
    t = x.shape[1]
    pos_k = None
    if "conformer.pos_emb.pe_k.weight" in sd:
        table = sd["conformer.pos_emb.pe_k.weight"]
        maxlen = table.shape[0] // 2
        idx = torch.arange(t)
        rel = (idx[:, None] - idx[None, :]).clamp(-maxlen, maxlen - 1) + maxlen
        pos_k = table[rel]
    nblk = 1 + max(int(k.split(".")[2]) for k in sd if k.startswith("conformer.encoders."))
    for i in range(nblk):
        p = "conformer.encoders.%d" % i
        x = x + 0.5 * _conformer_ffn(x, sd, p + ".feed_forward_in", drop, i, 0)
        x = x + _conformer_mhsa(x, pos_k, sd, p + ".self_attn", heads, drop, i)
        x = x + _conformer_conv(x, sd, p + ".conv", training, drop, i)
        x = x + 0.5 * _conformer_ffn(x, sd, p + ".feed_forward_out", drop, i, 5)
        x = _layer_norm(x, sd, p + ".layer_norm")
    m = torch.relu(_linear(x, sd, "linear"))
    return list(torch.chunk(m, num_out, 2))


def blstm_masks(f, sd, num_out=3):
    """Reference BLSTM forward up to the masks.  css/models/blstm.py:68-86,122-128,156-160."""
    if f.ndim == 4:
        f = f.squeeze(0)
    x = mvn_time(f)
    x = torch.relu(_layer_norm(_linear(x, sd, "blstm.embed.0"), sd, "blstm.embed.1"))
    nlay = 1 + max(int(k.split(".")[2]) for k in sd if k.startswith("blstm.encoders."))
    for i in range(nlay):
        p = "blstm.encoders.%d" % i
        x = _bilstm(x, sd, p + ".lstm")
        x = _layer_norm(x, sd, p + ".layer_norm")
    m = torch.relu(_linear(x, sd, "linear"))
    return list(torch.chunk(m, num_out, -1))


def conv_tasnet_forward(x, sd, num_spk=2, num_noise=1):
    """Reference Conv-TasNet forward (css/models/conv_tasnet.py:127-156, blocks :296-320, gLN :168-202), functional over a
    reference-named state_dict; separation-block norm "gln", non-causal, ReLU masks.  x (n, L) -> (n, num_spk, L')."""
    F = torch.nn.functional
    if x.dim() == 1:
        x = x.unsqueeze(0)
    L = sd["encoder.weight"].shape[-1]
    w = F.conv1d(x.unsqueeze(1), sd["encoder.weight"], sd["encoder.bias"], stride=L // 2)
    e = F.layer_norm(w.transpose(1, 2), (w.shape[1],), sd["LayerN_S.weight"], sd["LayerN_S.bias"], 1e-5).transpose(1, 2)
    e = F.conv1d(e, sd["BottleN_S.weight"], sd["BottleN_S.bias"])

    def gln(v, p):
        mean = v.mean((1, 2), keepdim=True)
        var = ((v - mean) ** 2).mean((1, 2), keepdim=True)
        return sd[p + ".weight"] * (v - mean) / torch.sqrt(var + 1e-5) + sd[p + ".bias"]

    r = 0
    while f"separation.{r}.0.conv1x1.weight" in sd:
        i = 0
        while f"separation.{r}.{i}.conv1x1.weight" in sd:
            p = f"separation.{r}.{i}"
            c = F.conv1d(e, sd[p + ".conv1x1.weight"], sd[p + ".conv1x1.bias"])
            c = gln(F.prelu(c, sd[p + ".PReLU_1.weight"]), p + ".norm_1")
            wd = sd[p + ".dwconv.weight"]
            dil = 2 ** i
            c = F.conv1d(c, wd, sd[p + ".dwconv.bias"], padding=dil * (wd.shape[-1] - 1) // 2, dilation=dil, groups=wd.shape[0])
            c = gln(F.prelu(c, sd[p + ".PReLU_2.weight"]), p + ".norm_2")
            e = e + F.conv1d(c, sd[p + ".Sc_conv.weight"], sd[p + ".Sc_conv.bias"])
            i += 1
        r += 1
    m = torch.relu(F.conv1d(e, sd["gen_masks.weight"], sd["gen_masks.bias"]))
    outs = [F.conv_transpose1d(w * mk, sd["decoder.weight"], sd["decoder.bias"], stride=L // 2).squeeze(1)
            for mk in torch.chunk(m, num_spk + num_noise, dim=1)]
    return torch.stack(outs[:-1], dim=1)


def _lstm_dir(x, w_ih, w_hh, b_ih, b_hh, reverse):
    """One LSTM direction, gate order i,f,g,o (torch.nn.LSTM semantics, css/models/blstm.py:139-145)."""
    n, t, _ = x.shape
    hdim = w_hh.shape[1]
    xp = torch.nn.functional.linear(x, w_ih, b_ih)
    h = x.new_zeros(n, hdim)
    c = x.new_zeros(n, hdim)
    out = x.new_empty(n, t, hdim)
    steps = range(t - 1, -1, -1) if reverse else range(t)
    for s in steps:
        g = xp[:, s] + torch.nn.functional.linear(h, w_hh, b_hh)
        gi, gf, gg, go = torch.chunk(g, 4, dim=1)
        c = torch.sigmoid(gf) * c + torch.sigmoid(gi) * torch.tanh(gg)
        h = torch.sigmoid(go) * torch.tanh(c)
        out[:, s] = h
    return out


def _bilstm(x, sd, p):
    fw = _lstm_dir(x, sd[p + ".weight_ih_l0"], sd[p + ".weight_hh_l0"], sd[p + ".bias_ih_l0"], sd[p + ".bias_hh_l0"], False)
    bw = _lstm_dir(
        x, sd[p + ".weight_ih_l0_reverse"], sd[p + ".weight_hh_l0_reverse"],
        sd[p + ".bias_ih_l0_reverse"], sd[p + ".bias_hh_l0_reverse"], True,
    )
    return torch.cat([fw, bw], dim=-1)


def model_masks(f, sd, heads=4, training=False, drop=None):
    """Dispatch on the state_dict's own key names (the reference's .mdl interchange contract)."""
    if any(k.startswith("conformer.") for k in sd):
        return conformer_masks(f, sd, heads, training=training, drop=drop)
    return blstm_masks(f, sd)


def model_forward(f, sd, num_bins=257, heads=4, training=False, drop=None):
    """(y_pred, masks) with the documented patch D3 (``f_orig[..., :num_bins]``).

    css/models/conformer.py:103-104, css/models/blstm.py:87-88; SURVEY.md A.9 D3.
    """
    if f.ndim == 4:
        f = f.squeeze(0)
    masks = model_masks(f, sd, heads, training, drop)
    y_pred = torch.stack([m * f[..., :num_bins] for m in masks[:-1]], dim=1)
    return y_pred, masks


# --------------------------------------------------------------------------------------
# Separator  (css/executor/separator.py)
# --------------------------------------------------------------------------------------
CHUNK_EXTRA = 256  # css/executor/separator.py:67


def chunk_signal(s, eval_win, eval_hop):
    """(D, T) -> (B, D, win+256).  css/executor/separator.py:67,76.  Tail is dropped."""
    return s.unfold(-1, eval_win + CHUNK_EXTRA, eval_hop).permute(1, 0, 2)


def steervec_7ch(nfreqs=257, nvecs=30, radius=0.0425, sndvelocity=340, samplerate=16000):
    """Far-field steering vectors of the 7-mic circular array.  css/executor/separator.py:113-163."""
    ang = torch.tensor([2 * math.pi * k / nvecs for k in range(nvecs)])
    offs = [None, math.pi / 6, -math.pi / 6, -math.pi / 2, -5 * math.pi / 6, 5 * math.pi / 6, math.pi / 2]
    cols = [torch.zeros(nvecs) if o is None else torch.cos(ang + o) for o in offs]
    deltas = radius * torch.stack(cols, dim=1) / sndvelocity * samplerate
    sv = torch.stack([torch.exp(1j * deltas * math.pi * f / (nfreqs - 1)) for f in range(nfreqs)])
    sv /= math.sqrt(7)
    return sv, ang * 360 / (2 * math.pi)


def doa_likelihood(X, sv, mask, lo, hi, compression=0.5, epsilon=1e-12):
    """css/executor/separator.py:202-250 (softmax=False branch).  -> (B, C, A)."""
    Xs = X[:, :, lo:hi]
    xpow = torch.abs(torch.einsum("bmft,bmft->bft", Xs, Xs.conj()))
    xh = torch.abs(torch.einsum("bmft,fam->bfta", Xs.conj(), sv[lo:hi])) ** 2
    tf = -torch.pow(xpow[..., None] - xh / (1 + epsilon), compression)
    return torch.einsum("bftc,bfta->bca", mask[:, lo:hi], tf)


def angle_merge(X, mask, sv, angles, lo, hi, thresh=16):
    """css/executor/separator.py:165-200.  Victim channel is always 0."""
    binmask = (mask > 0.5).float()
    like = doa_likelihood(X, sv, binmask, lo, hi)
    idx = torch.argmax(like, dim=-1)
    out = mask.clone()
    merged = []
    for b in range(mask.shape[0]):
        d0, d1 = angles[idx[b, 0]], angles[idx[b, 1]]
        hit = bool(min((d0 - d1) % 360, (d1 - d0) % 360) <= thresh)
        merged.append(hit)
        if hit:
            out[b, ..., 0] = 1e-12
    return out, idx, merged


def separate(s, cfg, mask_fn, sr=16000, lowcut=80, highcut=2000, return_debug=False):
    """css/executor/separator.py:63-111 with the session replaced by ``mask_fn``.

    mask_fn: f (1, T, F) -> list of 3 tensors (1, T, 257)  (the (1,F,T)->3x(257,T) session
    contract of separator.py:94-103 in model layout).  One chunk per call (batch_size 1).
    Returns masks (B,257,150,3), mags (B,257,150).
    """
    win, hop = int(cfg["eval_win"] * sr), int(cfg["eval_hop"] * sr)
    ipd = cfg.get("ipd", None)
    merge = cfg.get("merge", False)
    thresh = cfg.get("merge_threshold", 16)
    K = dft_kernel(cfg["frame_length"])
    if merge:
        sv, angles = steervec_7ch(257, 30, samplerate=sr)
        step = (sr // 2) / 256
        lo, hi = int(math.floor(lowcut / step)), int(math.ceil(highcut / step))
    chunks = chunk_signal(s, win, hop)
    masks, mags, dbg = [], [], []
    for b in range(chunks.shape[0]):
        mag, f, re, im = feature_extract(chunks[b : b + 1], cfg["frame_length"], cfg["frame_shift"], ipd, K)
        m = mask_fn(f.transpose(1, 2))
        m = torch.stack([mm[0].transpose(0, 1) for mm in m[:3]], dim=-1).unsqueeze(0)  # (1,F,T,3)
        m = torch.clamp(m, max=1.0)
        if merge:
            new, idx, hit = angle_merge(re + 1j * im, m[..., :2], sv, angles, lo, hi, thresh)
            m[..., :2] = new
            dbg.append((idx[0].tolist(), hit[0]))
        masks.append(m)
        mags.append(mag)
    out = torch.cat(masks, 0), torch.cat(mags, 0)
    return out + (dbg,) if return_debug else out


# --------------------------------------------------------------------------------------
# Stitcher  (css/executor/stitcher.py)
# --------------------------------------------------------------------------------------
def stitch_margin(cfg, sr=16000):
    """css/executor/stitcher.py:10-12 -> 100 frames for the shipped configs."""
    return int((cfg["eval_win"] * 10 - cfg["eval_hop"] * 10) / 10 * sr / cfg["hop_size"])


def get_stitch(mags, masks, margin=100, return_sims=False):
    """Adjacent-chunk permutation by overlap similarity.  css/executor/stitcher.py:14-52.

    mags (B,F,T), masks (B,F,T,3) -> list of B-1 permutations ([0,1] or [1,0]).
    Sums are fp32 ``torch.sum`` over the same (F, margin) slabs as the reference so the
    decision is reproduced bit for bit on CPU.
    """
    perms, sims = [], []
    for n in range(len(masks) - 1):
        e_prev = masks[n][:, :, :2].permute(2, 0, 1) * torch.abs(mags[n])
        e_now = masks[n + 1][:, :, :2].permute(2, 0, 1) * torch.abs(mags[n + 1])
        sim = torch.zeros((2, 2))
        for i in range(2):
            for j in range(2):
                d = e_prev[j, :, -margin:] - e_now[i, :, :margin]
                sim[i, j] = -torch.sum(torch.pow(torch.abs(d), 0.5))
        keep, swap = sim[0, 0] + sim[1, 1], sim[0, 1] + sim[1, 0]
        perms.append([0, 1] if keep >= swap else [1, 0])
        sims.append((float(keep), float(swap)))
    return (perms, sims) if return_sims else perms


def perm_states(perms):
    """Prefix-XOR of the swap bits.  css/executor/stitcher.py:55-60."""
    state, out = 0, [0]
    for p in perms:
        if p[0] == 1:
            state = 1 - state
        out.append(state)
    return out


def get_connect(perms, masks, hop=50):
    """Reorder by state, winner-take-all (floor 1e-4), overlap-average.  css/executor/stitcher.py:54-110."""
    states = perm_states(perms)
    B, F, win, _ = masks.shape
    L = hop * (B - 1) + win
    acc = torch.zeros((3, F, L))
    cnt = torch.zeros((1, L))
    floor = torch.tensor(1e-4, dtype=torch.float32)
    for i in range(B):
        m = masks[i]
        if states[i] == 1:
            m = torch.stack((m[:, :, 1], m[:, :, 0], m[:, :, 2]), dim=2)
        m = torch.where(m == torch.amax(m, dim=2, keepdim=True), m, floor)
        acc[:, :, hop * i : hop * i + win] += m.permute(2, 0, 1)
        cnt[:, hop * i : hop * i + win] += 1
    cnt[cnt == 0] = 1
    return acc[0] / cnt, acc[1] / cnt, acc[2] / cnt


# --------------------------------------------------------------------------------------
# asteroid.dsp.beamforming restatement (third party, unvendored, version unpinned;
# call sites css/executor/beamformer.py:4,11-14,151-158).  PARITY UNPINNED at this seam:
# cross-checked against torchaudio.functional.{psd,mvdr_weights_souden,apply_beamforming}.
# --------------------------------------------------------------------------------------
def compute_scm(x, mask, normalize=True):
    """scm[b,m,n,f] = sum_t mask x x^H (/ sum_t mask).  x (B,M,F,T) c64, mask (B,1,F,T)."""
    scm = torch.einsum("bmft,bnft->bmnf", mask * x, x.conj())
    if normalize:
        scm = scm / mask.sum(-1, keepdim=True).transpose(-1, -2)
    return scm


def souden_mvdr(mix, target_scm, noise_scm, ref_mic=0, eps=1e-8):
    """w = (Phi_n^-1 Phi_s / (tr(.)+eps))[:, ref]; out = w^H x.  Solve in complex128."""
    ps = target_scm.permute(0, 3, 1, 2)
    pn = noise_scm.permute(0, 3, 1, 2)
    num = torch.linalg.solve(pn.to(torch.complex128), ps.to(torch.complex128)).to(ps.dtype)
    tr = torch.diagonal(num, dim1=-2, dim2=-1).sum(-1)
    w = (num / (tr[..., None, None] + eps))[..., ref_mic]  # (B,F,M)
    return torch.einsum("bfm,bmft->bft", w.conj(), mix)


# --------------------------------------------------------------------------------------
# Beamformer  (css/executor/beamformer.py)
# --------------------------------------------------------------------------------------
def centered_stft(x, n_fft=512, hop=256):
    """css/executor/beamformer.py:137-145.  (dtype follows x: float64 input gives the complex128 bracket chain.)"""
    return torch.stft(x, n_fft=n_fft, hop_length=hop, center=True, onesided=True,
                      window=torch.hann_window(n_fft, dtype=x.dtype), return_complex=True)


def normalize_scale(result, ch0, mask):
    """css/executor/beamformer.py:165-183."""
    e_mask = torch.sqrt(torch.mean(torch.abs(mask * ch0) ** 2, dim=(1, 2), keepdim=True))
    e_out = torch.sqrt(torch.mean(torch.abs(result) ** 2, dim=(1, 2), keepdim=True))
    return result / e_out * e_mask


def process_one_batch(x, speech_mask, noise_mask, rescale=True):
    """css/executor/beamformer.py:126-163.  x (B,D,N), masks (B,F,T) -> (B,F,T) complex64."""
    B, D, N = x.shape
    S = centered_stft(x.reshape(B * D, N))
    S = S.reshape(B, D, S.shape[-2], S.shape[-1])
    L = min(S.shape[-1], speech_mask.shape[-1])
    S = S[..., :L]
    tgt = compute_scm(S, speech_mask[..., :L].unsqueeze(1))
    dis = compute_scm(S, noise_mask[..., :L].unsqueeze(1))
    out = souden_mvdr(S, tgt, dis)
    if rescale:
        out = normalize_scale(out, S[:, 0], speech_mask[..., :L])
    return out


def dedup(S):
    """In-place low-volume-competitor suppression.  css/executor/beamformer.py:68-80.  S (B,2,F,T)."""
    power = 10 * torch.log10(torch.sum(torch.abs(S) ** 2, dim=(2, 3)))
    mag = torch.abs(S)
    gain = mag / torch.amax(mag, dim=1, keepdim=True)
    for b in range(S.shape[0]):
        if power[b, 0] - power[b, 1] > 15:
            S[b, 1] = torch.clamp(gain[b, 1], min=10 ** (-40 / 20)) * S[b, 1]
        elif power[b, 1] - power[b, 0] > 15:
            S[b, 0] = torch.clamp(gain[b, 0], min=10 ** (-40 / 20)) * S[b, 0]
    return S


def placement_plan(n_chunks, T, hop, margin, wav_len, batch_size):
    """Enumerate the slice additions of css/executor/beamformer.py:54-57,104-118.

    Returns a list of (chunk, src_start, src_end, dst_start) in the reference's order
    (chunk 32k appears twice; the ``en == T`` branch is reproduced as written).
    """
    plan = []
    for i in range(math.ceil(n_chunks / batch_size)):
        bs = i * batch_size
        be = min((i + 1) * batch_size, n_chunks) + 1
        st, en = hop * bs, hop * be
        for c in range(bs, min(be, n_chunks)):
            if st == 0:
                plan.append((c, 0, margin, 0))
            elif en == T:
                plan.append((c, margin - hop, wav_len, st + margin - hop))
            else:
                plan.append((c, margin - hop, margin, st + margin - hop))
            st += hop
            en += hop
    return plan


def continuous_process(X, masks, cfg, sr=16000, rescale=True, dtype=None):
    """css/executor/beamformer.py:32-124.  X (D,T); masks 3x(F,L) -> two (1,T) waveforms.

    ``dtype=torch.float64`` runs the same chain (STFT, covariances, solve, apply, rescale, dedup, iSTFT, placement) in
    float64 / complex128 on the given fp32 inputs: the "exact" answer the fp32 implementations are bracketed against.
    """
    if dtype is not None:
        X = X.to(dtype)
        masks = tuple(m.to(dtype) for m in masks)
    n_fft, hop_len = cfg["n_fft"], cfg["hop_size"]
    win, hop = int(cfg["eval_win"] * sr), int(cfg["eval_hop"] * sr)
    margin = cfg["proceed_margin"] * sr
    bsz = cfg["batch_size"]
    mhop, mwin = int(hop / hop_len), int(win / hop_len)
    D, T = X.shape
    Xc = X.unfold(-1, win + 256, hop).permute(1, 0, 2)
    m0, m1, mn = (m.unfold(1, mwin, mhop).permute(1, 0, 2) for m in masks)
    res = torch.zeros(2, T, dtype=X.dtype)
    window = torch.hann_window(n_fft, dtype=X.dtype)
    for i in range(math.ceil(Xc.shape[0] / bsz)):
        bs, be = i * bsz, min((i + 1) * bsz, Xc.shape[0]) + 1
        xb = Xc[bs:be]
        o0 = process_one_batch(xb, m0[bs:be], mn[bs:be], rescale)
        o1 = process_one_batch(xb, m1[bs:be], mn[bs:be], rescale)
        S = dedup(torch.stack((o0, o1), dim=1))
        wavs = []
        for k in range(2):
            w = torch.istft(S[:, k], n_fft=n_fft, hop_length=hop_len, center=True, onesided=True,
                            window=window, return_complex=False)
            wavs.append(w[:, : xb.shape[-1]])
        st, en = hop * bs, hop * be
        for b in range(len(wavs[0])):
            for k in range(2):
                w = wavs[k]
                if st == 0:
                    res[k, st : st + margin] += w[b, :margin]
                elif en == T:
                    res[k, st + margin - hop : st + w.shape[1]] += w[b, margin - hop :]
                else:
                    res[k, st + margin - hop : st + margin] += w[b, margin - hop : margin]
            st += hop
            en += hop
    out0 = res[0] * 0.9 / torch.max(torch.abs(res[0]))
    out1 = res[1] * 0.9 / torch.max(torch.abs(res[1]))
    return out0.unsqueeze(0), out1.unsqueeze(0)


# --------------------------------------------------------------------------------------
# Objectives  (css/objectives/pit.py, css/objectives/mse.py)
# --------------------------------------------------------------------------------------
def pit_loss(estimate, target, axis, loss_fn=torch.nn.functional.mse_loss, return_permutation=False, **kw):
    """Exhaustive-permutation loss, first minimum wins.  css/objectives/pit.py:5-62."""
    k = estimate.size()[axis]
    assert k < 30, "too many sources"
    assert estimate.size() == target.size(), f"{estimate.size()} != {target.size()}"
    perms = list(itertools.permutations(range(k)))
    cands = [loss_fn(estimate.index_select(axis, torch.tensor(p)), target, **kw) for p in perms]
    best, idx = torch.min(torch.stack(cands), dim=0)
    return (best, perms[int(idx)]) if return_permutation else best


def pit_mse(y_pred, y_true, return_permutation=False):
    """Batch mean of per-utterance PIT-MSE.  css/objectives/mse.py:29-37.  (B,K,T,F) x2."""
    assert y_pred.shape == y_true.shape, f"{y_pred.shape} != {y_true.shape}"
    total, perms = 0, []
    for b in range(y_pred.shape[0]):
        l, p = pit_loss(y_pred[b], y_true[b], axis=0, return_permutation=True)
        total = total + l
        perms.append(p)
    loss = total / y_pred.shape[0]
    return (loss, perms) if return_permutation else loss


def mse_objective(sd, sample, heads=4, training=False, drop=None):
    """MeanSquaredError.forward with patches D3/D4.  css/objectives/mse.py:22-37."""
    y_pred, _ = model_forward(sample["mix"], sd, heads=heads, training=training, drop=drop)
    y_true = torch.stack((sample["source1"], sample["source2"]), dim=1)
    return pit_mse(y_pred, y_true)


def train_steps(sd, samples, lr=1e-4, weight_decay=1e-2, grad_thresh=5.0, heads=4, drops=None):
    """The body of train_one_epoch (css/trainer/iteration_types.py:34-63) with the optimiser of
    scripts/python/train.py:209-211 (torch.optim.Adam(lr, weight_decay)) and train.sh:16 (clip 5.0).  ``drops``: one
    dropout hook per step (see conformer_masks) -- the masks are an INPUT here; None = dropout probability 0.

    sd: state_dict (modified in place: parameters updated, BatchNorm running stats are NOT tracked here).
    Returns per step: loss, gradient norm (before clipping), and the gradients of the first step.
    """
    names = [k for k, v in sd.items() if v.dtype.is_floating_point and "running_" not in k]
    params = [sd[k].detach().clone().requires_grad_(True) for k in names]
    opt = torch.optim.Adam(params, lr=lr, weight_decay=weight_decay)
    out = {"loss": [], "norm": [], "grads": None}
    for step, sample in enumerate(samples):
        live = dict(sd)
        live.update(zip(names, params))
        loss = mse_objective(live, sample, heads=heads, training=True, drop=drops[step] if drops else None)
        opt.zero_grad()
        loss.backward()
        if step == 0:
            out["grads"] = {k: (p.grad.detach().clone() if p.grad is not None else torch.zeros_like(p)) for k, p in zip(names, params)}
        norm = torch.nn.utils.clip_grad_norm_(params, grad_thresh)
        opt.step()
        out["loss"].append(float(loss.detach()))
        out["norm"].append(float(norm.detach()))
    for k, p in zip(names, params):
        sd[k] = p.detach()
    return out


# --------------------------------------------------------------------------------------
# End-to-end + quality metric
# --------------------------------------------------------------------------------------
def css_pipeline(wav, cfg, mask_fn, sr=16000, stages=None):
    """scripts/python/separate_libricss.py:71-78: separate -> get_stitch -> get_connect -> beamform."""
    import time

    t0 = time.perf_counter()
    masks, mags = separate(wav, cfg["separation"], mask_fn, sr)
    t1 = time.perf_counter()
    perms = get_stitch(mags, masks, stitch_margin(cfg["stitching"], sr))
    t2 = time.perf_counter()
    hop = int(cfg["stitching"]["eval_hop"] * sr / cfg["stitching"]["hop_size"])
    stitched = get_connect(perms, masks, hop)
    t3 = time.perf_counter()
    if wav.shape[0] > 1:
        out = continuous_process(wav, stitched, cfg["beamforming"], sr)
    else:
        out = None
    t4 = time.perf_counter()
    if stages is not None:
        stages.update(separate=t1 - t0, stitch=t2 - t1, connect=t3 - t2, beamform=t4 - t3)
    return {"masks": masks, "mags": mags, "perms": perms, "stitched": stitched, "wavs": out}


def si_sdr(est, ref):
    """Scale-invariant SDR in dB (standard definition; the reference implements no metric)."""
    est = est.double() - est.double().mean()
    ref = ref.double() - ref.double().mean()
    a = torch.dot(est, ref) / torch.dot(ref, ref).clamp_min(1e-30)
    proj = a * ref
    return float(10 * torch.log10(torch.dot(proj, proj) / torch.dot(est - proj, est - proj).clamp_min(1e-30)))
