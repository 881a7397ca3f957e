"""Multi-GPU plumbing: one process per GPU, recordings sharded across ranks (run.sh:21-36 of the reference
starts one process per LibriCSS session; here the sessions of a job are dealt round-robin to the ranks).

The data path needs no collective -- recordings are independent.  The only exchange is the final gather of
the separated waveforms to rank 0 (NCCL over NVLink on GPUs; gloo in the CPU tests).
"""
import os

import torch
import torch.distributed as dist


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
