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


def test_long_recording_shard_plan():
    """Host logic of the one-recording sharding: ranges partition the chunks, halos stay in bounds, and the parity
    exchange reproduces the global prefix-XOR permutation states for random swap bits."""
    import random

    from css_b200.dist import chunk_ranges, extended_range, first_state, owned_parity

    rng = random.Random(0)
    for B in (1, 2, 3, 7, 72, 747, 4497):
        swaps = [rng.randint(0, 1) for _ in range(B - 1)]
        state = [0]
        for s in swaps:
            state.append(state[-1] ^ s)                      # stitcher.py:55-60
        for world in (1, 2, 3, 8):
            ranges = chunk_ranges(B, world)
            assert ranges[0][0] == 0 and ranges[-1][1] == B and all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
            assert max(c1 - c0 for c0, c1 in ranges) - min(c1 - c0 for c0, c1 in ranges) <= 1
            parities, locals_ = [], []
            for c0, c1 in ranges:
                if c1 == c0:
                    parities.append(0); locals_.append(None)
                    continue
                e0, e1 = extended_range(c0, c1, B)
                assert 0 <= e0 <= c0 and c1 <= e1 <= B
                sl = [0]
                for s in swaps[e0:e1 - 1]:
                    sl.append(sl[-1] ^ s)
                locals_.append((sl, e0, e1))
                parities.append(owned_parity(sl, c0, c1, e0))
            for r, (c0, c1) in enumerate(ranges):
                if c1 == c0:
                    continue
                sl, e0, e1 = locals_[r]
                prefix = 0
                for p in parities[:r]:
                    prefix ^= p
                s0 = first_state(prefix, sl, c0, e0)
                assert [s0 ^ v for v in sl] == state[e0:e1]


def test_long_recording_sample_ownership():
    """Host logic of the sharded output: the ranks' sample ranges partition [0, T), every placement of a rank's
    own chunks lands inside its range plus the spill, and only the `en == T` items (A.6 Q4) reach into the neighbour's range."""
    import types

    from css_b200.dist import ShardedCssPipeline, chunk_ranges
    from css_b200.executor.beamformer import Beamformer, placement_plan
    from css_b200.pipeline import CONFIG_7CH

    sh = ShardedCssPipeline.__new__(ShardedCssPipeline)
    sh.pipe = types.SimpleNamespace(beamformer=Beamformer(CONFIG_7CH["beamforming"]))
    assert sh.spill == 6144
    for T, B in ((960000, 72), (1120000, 85), (9600000, 747)):
        plan = placement_plan(B, T, 12800, 32000, 38144, 32)
        for world in (1, 2, 3, 8, 13):
            ranges = chunk_ranges(B, world)
            spans = [sh.owned_samples(c0, c1, B, T) for c0, c1 in ranges]
            live = [s for s in spans if s[1] > s[0]]
            assert live[0][0] == 0 and live[-1][1] == T and all(a[1] == b[0] for a, b in zip(live, live[1:]))
            crossings = 0
            for (c0, c1), (s0, s1) in zip(ranges, spans):
                for c, a, b, dst in plan:
                    if c0 <= c < c1:
                        end = min(T, dst + (b - a))
                        assert dst >= s0 and end <= min(T, s1 + sh.spill)
                        crossings += end > s1
            if T % 12800:
                assert crossings == 0          # no `en == T` branch without T a multiple of the hop
