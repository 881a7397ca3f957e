"""world_size-2 gloo test of the multi-GPU host logic (sharding + final gather), on CPU."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, port, lengths, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    from css_b200.dist import gather_waveforms, init_from_env, shard_recordings

    r, w, _ = init_from_env(backend="gloo")
    mine = shard_recordings(lengths, r, w)
    local = {i: torch.full((2, lengths[i]), float(i + 1)) * torch.tensor([[1.0], [-1.0]]) for i in mine}
    res = gather_waveforms(local, lengths, r, w, dst=0)
    if r == 0:
        ok = all(res[i].shape == (2, lengths[i]) and float(res[i][0, 0]) == i + 1 and float(res[i][1, -1]) == -(i + 1)
                 for i in range(len(lengths)))
        q.put(ok)
    else:
        assert res is None
    dist.barrier()
    dist.destroy_process_group()


def test_shard_is_a_partition():
    from css_b200.dist import shard_recordings

    lengths = [100, 900, 300, 300, 50, 700, 20]
    for world in (1, 2, 4, 8):
        got = sorted(i for r in range(world) for i in shard_recordings(lengths, r, world))
        assert got == list(range(len(lengths)))
    # longest-first dealing balances the load
    loads = [sum(lengths[i] for i in shard_recordings(lengths, r, 2)) for r in range(2)]
    assert abs(loads[0] - loads[1]) <= max(lengths)


def test_gather_world2_gloo():
    lengths = [1000, 640, 2000, 333, 1200]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, lengths, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=10) is True
