"""Inter-chunk permutation + overlap averaging (reference: css/executor/stitcher.py).

``get_stitch`` runs one CTA per adjacent chunk pair (all pairs in one launch) and ``get_connect`` one
gather-style kernel over the output frames; the permutation state is a device-side prefix XOR.
"""
import torch

from .. import _native as nv


class Stitcher:
    def __init__(self, stitching_config, sr=16000):
        self.eval_win = stitching_config["eval_win"]
        self.eval_hop = stitching_config["eval_hop"]
        self.fft_hop = stitching_config["hop_size"]
        self.sr = sr
        self.stitch_margin = int((self.eval_win * 10 - self.eval_hop * 10) / 10 * self.sr / self.fft_hop)
        self.last_sims = None

    def stitch_device(self, x, masks):
        """-> swap (B-1,) int32 CUDA (1 where the reference emits [1,0]); keeps sims for margin reports."""
        nv.require_cuda(masks, "masks")
        x = nv.require_cuda(x, "x").float().contiguous()
        masks = masks.float().contiguous()
        B, F, T, C = masks.shape
        if F != nv.NBINS or x.shape != (B, F, T):
            raise RuntimeError(f"get_stitch: bad shapes {tuple(x.shape)} / {tuple(masks.shape)}")
        swap = torch.zeros(max(B - 1, 1), dtype=torch.int32, device=masks.device)
        sims = torch.zeros(max(B - 1, 1), 4, dtype=torch.float32, device=masks.device)
        nv.call("css_stitch_perm", nv.ptr(x), nv.ptr(masks), nv.c_long(B), nv.c_int(T), nv.c_int(C),
                nv.c_int(self.stitch_margin), nv.ptr(swap), nv.ptr(sims))
        self.last_sims = sims[: B - 1]
        return swap[: B - 1]

    def get_stitch(self, x, masks):
        """reference: stitcher.py:14-52 -> list of [0,1] / [1,0] (one device->host read)."""
        swap = self.stitch_device(x, masks)
        return [[1, 0] if v else [0, 1] for v in swap.tolist()]

    def connect_device(self, swap, masks, state0=None):
        """-> stitched [3][L][KP] fp32 CUDA (bin fastest; the beamformer's native layout).

        ``state0`` (1-element int32 CUDA tensor, optional) is the permutation state of the first chunk -- non-zero when
        ``masks`` is a chunk range in the middle of a recording that is sharded across GPUs (css_b200/dist.py)."""
        masks = nv.require_cuda(masks, "masks").float().contiguous()
        B, F, T, C = masks.shape
        if F != nv.NBINS or C != 3:
            raise RuntimeError(f"get_connect: masks must be (B,257,T,3), got {tuple(masks.shape)}")
        if swap.numel() != B - 1:
            raise RuntimeError(f"get_connect: {swap.numel()} permutations for {B} chunks")
        hop = int(self.eval_hop * self.sr / self.fft_hop)
        state = torch.empty(B, dtype=torch.int32, device=masks.device)
        swap = swap.to(masks.device, torch.int32).contiguous() if swap.numel() else torch.zeros(1, dtype=torch.int32, device=masks.device)
        nv.call("css_perm_states", nv.ptr(swap), nv.c_long(B), nv.ptr(state))
        if state0 is not None:
            state ^= state0.to(state.device, torch.int32).reshape(1)
        L = hop * (B - 1) + T
        out = torch.empty(3, L, nv.KP, dtype=torch.float32, device=masks.device)
        nv.call("css_stitch_connect", nv.ptr(masks), nv.ptr(state), nv.c_long(B), nv.c_int(T), nv.c_int(C), nv.c_int(hop), nv.ptr(out))
        self.last_state = state
        return out

    def get_connect(self, PERM, mask):
        """reference: stitcher.py:54-110 -> (res_1, res_2, res_noise), each (257, L)."""
        if isinstance(PERM, torch.Tensor):
            swap = PERM
        else:
            swap = torch.tensor([1 if p[0] == 1 else 0 for p in PERM], dtype=torch.int32)
        out = self.connect_device(swap, mask)
        res = tuple(out[c, :, : nv.NBINS].t() for c in range(3))
        for r in res:
            r._css_stitched = out   # lets Beamformer reuse the native layout without a copy
        return res
