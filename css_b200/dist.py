"""Multi-GPU plumbing: one process per GPU, recordings sharded across ranks (run.sh:21-36 of the reference
starts one process per LibriCSS session; here the sessions of a job are dealt round-robin to the ranks).

The data path needs no collective -- recordings are independent.  The only exchange is the final gather of
the separated waveforms to rank 0 (NCCL over NVLink on GPUs; gloo in the CPU tests).
"""
import os

import torch
import torch.distributed as dist

from . import _native as nv


def init_from_env(backend=None):
    """Initialise torch.distributed from torchrun's environment; returns (rank, world, local_rank)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world)
    return rank, world, local


def shard_recordings(lengths, rank, world):
    """Longest-first round-robin assignment of recordings to ranks -> indices owned by `rank`."""
    order = sorted(range(len(lengths)), key=lambda i: (-lengths[i], i))
    return sorted(order[rank::world])


def gather_waveforms(local, lengths, rank, world, dst=0):
    """Gather per-recording (2, T_i) outputs to `dst`.

    local: {recording index: (2, T_i) tensor} for the recordings this rank owns.  Every rank passes the full
    `lengths` list, so the owner of each recording is known everywhere.  Returns the complete list on `dst`
    (None elsewhere).  Uses one padded gather (NCCL gather / gloo gather) -- no data-path collective.
    """
    if world == 1:
        return [local[i] for i in range(len(lengths))]
    owners = {}
    for r in range(world):
        for i in shard_recordings(lengths, r, world):
            owners[i] = r
    mine = shard_recordings(lengths, rank, world)
    slots = max(len(shard_recordings(lengths, r, world)) for r in range(world))
    tmax = max(lengths)
    ref = next(iter(local.values())) if local else None
    device = ref.device if ref is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")
    buf = torch.zeros(slots, 2, tmax, dtype=torch.float32, device=device)
    for s, i in enumerate(mine):
        buf[s, :, : lengths[i]] = local[i]
    out = [torch.empty_like(buf) for _ in range(world)] if rank == dst else None
    dist.gather(buf, out, dst=dst)
    if rank != dst:
        return None
    res = [None] * len(lengths)
    for r in range(world):
        for s, i in enumerate(shard_recordings(lengths, r, world)):
            res[i] = out[r][s, :, : lengths[i]]
    return res


# ---------------------------------------------------------------------------------------------------------------
# One long recording sharded across the GPUs of a box (BASELINE.json configs[4]: 1-hour 7-ch meeting)
# ---------------------------------------------------------------------------------------------------------------
HALO = 2   # a stitched frame averages up to 3 consecutive chunks (stitcher.py:96-110): 2 neighbours on either side


def chunk_ranges(n_chunks, world):
    """Contiguous, near-equal chunk ranges [c0, c1) per rank (ranks beyond the chunk count get empty ranges)."""
    base, rem = divmod(n_chunks, world)
    out, c = [], 0
    for r in range(world):
        n = base + (1 if r < rem else 0)
        out.append((c, c + n))
        c += n
    return out


def extended_range(c0, c1, n_chunks, halo=HALO):
    """Chunks a rank evaluates: its own plus `halo` recomputed neighbours on each side (no mask exchange needed)."""
    return max(0, c0 - halo), min(n_chunks, c1 + halo)


def owned_parity(state_local, c0, c1, e0):
    """XOR of the swap bits of the pairs (k, k+1) with k+1 in [c0, c1), from the rank-local prefix states
    (state_local[j] = XOR of the local swap bits before local chunk j).  Works on tensors or lists."""
    lo = max(c0 - 1, 0) - e0
    return state_local[c1 - 1 - e0] ^ state_local[lo]


def first_state(prefix_parity, state_local, c0, e0):
    """Global permutation state of the rank's first evaluated chunk e0, given the XOR of the owned parities of all
    earlier ranks (= state of chunk c0-1)."""
    if c0 == 0:
        return prefix_parity * 0
    return prefix_parity ^ state_local[c0 - 1 - e0]


class ShardedCssPipeline:
    """``CssPipeline`` for ONE recording split into contiguous chunk ranges, one per rank.

    Every rank evaluates the masks of its own chunks plus a 2-chunk halo on each side (recomputed rather than exchanged:
    4 extra chunks against hundreds), decides the chunk-pair permutations locally (the kernels are deterministic, so a
    halo pair gets the same bit on both ranks), and the only data-path exchange is an all-gather of one parity bit per
    rank that turns the rank-local permutation states into global ones (stitcher.py:55-60 is a prefix XOR).  MVDR and
    the iSTFT placement run on the owned chunks with the GLOBAL placement plan (its 32-chunk batch seams are defined by
    global chunk indices, beamformer.py:54-57) into a buffer that covers only the rank's OWN sample range; a rank's last
    chunk can spill <= 6,144 samples into its right neighbour's range (the `en == T` branch), exchanged by one small
    all-gather; the global peak (beamformer.py:122-123) is an all-reduce(MAX) of two values and every rank scales its own
    range.  No rank ever holds or moves the whole (2, T) output unless a single writer asks for it (``gather=True``).
    Results equal the single-GPU pipeline's bit for bit.
    """

    def __init__(self, cfg, model, device, rank, world, group=None):
        from .pipeline import CssPipeline

        self.pipe = CssPipeline(cfg, model, device=device)
        self.rank, self.world, self.group = rank, world, group
        self.device = torch.device(device)
        self._layers = {}

    def _phase1(self, wav, c0, c1, B):
        """Masks, local pair decisions and local prefix states of the chunk range [c0, c1) plus its halo."""
        sep, st = self.pipe.separator, self.pipe.stitcher
        dev = self.device
        e0, e1 = extended_range(c0, c1, B)
        lo, hi = e0 * sep.eval_hop, (e1 - 1) * sep.eval_hop + sep.chunk_len
        if wav.is_cuda:
            local = wav[:, lo:hi].to(dev, torch.float32)
            masks, mags = sep.separate(local)
        elif wav.is_pinned() and wav.dtype == torch.float32 and e1 - e0 >= 64:
            # the Separator streams a pinned host slice in pieces, the mask network of the first chunks running underneath the copy
            masks, mags = sep.separate(wav[:, lo:hi])
            local = sep.last["wav"]
        else:   # row-wise copies: a strided host slice would be staged through pageable memory
            local = torch.empty(wav.shape[0], hi - lo, dtype=torch.float32, device=dev)
            for d in range(wav.shape[0]):
                local[d].copy_(wav[d, lo:hi], non_blocking=True)
            masks, mags = sep.separate(local)
        swap = st.stitch_device(mags, masks)
        state_local = torch.zeros(e1 - e0, dtype=torch.int32, device=dev)
        if e1 - e0 > 1:
            state_local[1:] = (torch.cumsum(swap, 0) & 1).to(torch.int32)
        return {"c0": c0, "c1": c1, "e0": e0, "e1": e1, "local": local, "masks": masks, "swap": swap, "state_local": state_local,
                "spec": sep.last["spec"], "parity": owned_parity(state_local, c0, c1, e0).reshape(1)}

    # ------------------------------------------------------------------ sample ownership
    def sample_start(self, c, T):
        """First output sample placed by chunk c (beamformer.py:104-118): 0 for chunk 0, else hop*c + margin - hop."""
        bf = self.pipe.beamformer
        return 0 if c == 0 else min(T, bf.eval_hop * c + bf.proceed_margin - bf.eval_hop)

    def owned_samples(self, c0, c1, B, T):
        """[s0, s1): the output samples whose LAST-rank-independent owner is the rank with chunks [c0, c1): from the first
        sample its first chunk places to the first sample the next rank's first chunk places (the last rank owns the tail)."""
        if c1 <= c0:
            return 0, 0
        return self.sample_start(c0, T), (T if c1 >= B else self.sample_start(c1, T))

    @property
    def spill(self):
        """Samples a rank's last chunk can place beyond its own range: only the `en == T` branch (A.6 Q4) writes
        wav_len - margin samples into the next chunk's region (6,144 for the shipped configs)."""
        bf = self.pipe.beamformer
        return max(0, bf.hop_length * (bf.mask_win - 1) - bf.proceed_margin)

    def _phase2(self, ctx, prefix, B, T, out, s0=0):
        """Global states -> stitched masks -> MVDR of the range -> placement of the OWNED chunks into out (2, n), whose
        sample 0 is recording sample s0."""
        st, bf = self.pipe.stitcher, self.pipe.beamformer
        c0, c1, e0, e1 = ctx["c0"], ctx["c1"], ctx["e0"], ctx["e1"]
        st0 = first_state(prefix, ctx["state_local"], c0, e0).reshape(1)
        stitched = st.connect_device(ctx["swap"], ctx["masks"], state0=st0)
        Y = bf.beamform_chunks(ctx["local"], stitched, e1 - e0, spec=ctx["spec"])
        key = (B, T, c0, c1, e0, s0, str(out.device))
        layers = self._layers.get(key)
        if layers is None:      # built once per shape: the boolean selection synchronises, the steady state must not
            layers = []
            for items in bf._plan(B, T, out.device):
                keep = (items[:, 0] >= c0) & (items[:, 0] < c1)
                sub = items[keep].clone()
                sub[:, 0] -= e0
                sub[:, 3] -= s0
                layers.append(sub.contiguous())
            self._layers[key] = layers
        bf.place(Y, layers, out)
        return st.last_state[c0 - e0 : c1 - e0]

    def _finish_local(self, loc, own, spill_prev, bits):
        """Add the previous rank's spill, take the maximum of the owned range into bits (2 x uint32)."""
        if spill_prev is not None and own > 0:
            n = min(spill_prev.shape[1], own)
            loc[:, :n] += spill_prev[:, :n]
        nv.call("css_absmax", nv.ptr(loc), nv.c_int(2), nv.c_long(own), nv.c_long(loc.shape[1]), nv.ptr(bits))

    def __call__(self, wav, return_intermediates=False, gather=True):
        """wav: (7, T) fp32 host (pinned) or device, the same on every rank.

        gather=True : -> (2, T) on rank 0 (NCCL gather of the ranks' normalised sample ranges), None elsewhere.
        gather=False: -> (slice (2, s1-s0), (s0, s1)) on every rank: each rank keeps (and can write / copy to the host) its
                      own sample range; no waveform bytes cross the GPUs.

        Exchange steps: all-gather of one permutation-parity bit per rank; all-gather of each rank's
        <= 6,144-sample spill into its right neighbour's range (A.6 Q4); all-reduce(MAX) of the two peak values."""
        sep, bf = self.pipe.separator, self.pipe.beamformer
        D, T = wav.shape
        B = sep.n_chunks(T)
        if B < 1:
            raise RuntimeError(f"waveform of {T} samples is shorter than one window of {sep.chunk_len}")
        ranges = chunk_ranges(B, self.world)
        c0, c1 = ranges[self.rank]
        dev = self.device
        s0, s1 = self.owned_samples(c0, c1, B, T)
        own, sp = s1 - s0, self.spill
        loc = torch.zeros(2, own + sp, dtype=torch.float32, device=dev)
        ctx = self._phase1(wav, c0, c1, B) if c1 > c0 else None
        parity = ctx["parity"] if ctx else torch.zeros(1, dtype=torch.int32, device=dev)
        if self.world > 1:
            allp = torch.zeros(self.world, dtype=torch.int32, device=dev)
            dist.all_gather_into_tensor(allp, parity.contiguous(), group=self.group)
        else:
            allp = parity
        states = None
        if ctx:
            prefix = (allp[: self.rank].sum() & 1).to(torch.int32).reshape(1)
            states = self._phase2(ctx, prefix, B, T, loc, s0)
        bits = torch.zeros(2, dtype=torch.int32, device=dev)
        spill_prev = None
        if self.world > 1 and sp > 0:
            spills = torch.empty(self.world, 2, sp, dtype=torch.float32, device=dev)
            dist.all_gather_into_tensor(spills, loc[:, own:].contiguous(), group=self.group)
            if self.rank > 0:
                spill_prev = spills[self.rank - 1]
        self._finish_local(loc, own, spill_prev, bits)
        if self.world > 1:
            dist.all_reduce(bits, op=dist.ReduceOp.MAX, group=self.group)    # float bits of non-negative values order as ints
        nv.call("css_peak_scale", nv.ptr(loc), nv.c_int(2), nv.c_long(own), nv.c_long(loc.shape[1]), nv.c_float(0.9), nv.ptr(bits))
        mine = loc[:, :own]
        self.last_collective_bytes = {"parity": 4 * self.world, "spill": 8 * sp * self.world if self.world > 1 else 0, "peak": 8,
                                      "gather": 8 * T if (gather and self.world > 1) else 0}
        if not gather:
            res = (mine, (s0, s1))
        elif self.world == 1:
            res = mine
        else:
            res = self._gather(mine, ranges, B, T)
        return (res, {"state": states, "range": (c0, c1), "samples": (s0, s1)}) if return_intermediates else res

    def _gather(self, mine, ranges, B, T):
        """NCCL gather of the ranks' sample ranges onto rank 0 -> (2, T)."""
        spans = [self.owned_samples(a, b, B, T) for a, b in ranges]
        width = max(b - a for a, b in spans)
        buf = torch.zeros(2, width, dtype=torch.float32, device=self.device)
        buf[:, : mine.shape[1]] = mine
        outs = [torch.empty_like(buf) for _ in range(self.world)] if self.rank == 0 else None
        dist.gather(buf, outs, dst=0, group=self.group)
        if self.rank != 0:
            return None
        full = torch.empty(2, T, dtype=torch.float32, device=self.device)
        for (a, b), o in zip(spans, outs):
            full[:, a:b] = o[:, : b - a]
        return full

    def emulate(self, wav, world):
        """Single-process emulation of a `world`-rank run (the ranks' work executed one after the other on this GPU, the
        collectives done in place).  Used by the parity tests on one-GPU boxes."""
        sep, bf = self.pipe.separator, self.pipe.beamformer
        D, T = wav.shape
        B = sep.n_chunks(T)
        dev = self.device
        ranges = chunk_ranges(B, world)
        ctxs = [self._phase1(wav, c0, c1, B) if c1 > c0 else None for c0, c1 in ranges]
        zero = torch.zeros(1, dtype=torch.int32, device=dev)
        allp = torch.cat([c["parity"] if c else zero for c in ctxs])
        sp = self.spill
        locs, states, spans = [], [], []
        for r, (ctx, (c0, c1)) in enumerate(zip(ctxs, ranges)):
            s0, s1 = self.owned_samples(c0, c1, B, T)
            loc = torch.zeros(2, s1 - s0 + sp, dtype=torch.float32, device=dev)
            if ctx:
                states.append(self._phase2(ctx, (allp[:r].sum() & 1).to(torch.int32).reshape(1), B, T, loc, s0))
            locs.append(loc)
            spans.append((s0, s1))
        bits = torch.zeros(2, dtype=torch.int32, device=dev)
        for r, (loc, (s0, s1)) in enumerate(zip(locs, spans)):
            prev = locs[r - 1][:, spans[r - 1][1] - spans[r - 1][0]:] if r > 0 and sp > 0 else None
            self._finish_local(loc, s1 - s0, prev.clone() if prev is not None else None, bits)
        out = torch.zeros(2, T, dtype=torch.float32, device=dev)
        for loc, (s0, s1) in zip(locs, spans):
            nv.call("css_peak_scale", nv.ptr(loc), nv.c_int(2), nv.c_long(s1 - s0), nv.c_long(loc.shape[1]), nv.c_float(0.9), nv.ptr(bits))
            out[:, s0:s1] = loc[:, : s1 - s0]
        return out, torch.cat(states)
