"""CPU tests of the training path's host logic: the oracle's training step against the reference-generated golden
vectors, the bucket plan of the overlapped gradient exchange, the LR schedule, and a world-size-2 gloo run of the
bucketed all-reduce."""
import argparse
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_oracle_train_step_matches_reference_golden():
    from oracle import css_oracle as O
    from tests.helpers import train_conformer, train_samples

    G = np.load(os.path.join(ROOT, "tests", "golden", "golden_train.npz"))
    torch.set_num_threads(8)
    model = train_conformer()          # css_b200 parameter container: reference init under the same seed
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    res = O.train_steps(sd, train_samples(2, 4, 50))
    assert np.allclose(res["loss"], G["E_loss"], rtol=1e-5)
    assert np.allclose(res["norm"], G["E_norm"], rtol=1e-4)
    for key in G.files:
        if key.startswith("E_grad/"):
            g = res["grads"][key[len("E_grad/"):]].reshape(-1)
            got = g[:: max(1, g.numel() // 4096)].numpy()
            assert np.abs(got - G[key]).max() <= 2e-3 * np.abs(G[key]).max(), key
        if key.startswith("E_param2/"):
            p = sd[key[len("E_param2/"):]].reshape(-1)
            got = p[:: max(1, p.numel() // 4096)].numpy()
            assert np.abs(got - G[key]).max() <= 1e-6 + 1e-5 * np.abs(G[key]).max(), key


def test_bucket_plan_partitions_the_flat_gradient():
    from css_b200.trainer.optim import bucket_ranges, slice_bounds

    for L in (0, 1, 2, 16):
        bounds = [1000 + 700 * l for l in range(L + 1)]     # block offsets, then the head
        size = bounds[-1] + 300
        for buckets in (1, 2, 4, 7, 64):
            ranges = bucket_ranges(L, buckets)
            assert len(ranges) <= buckets
            parts = [p for hi, lo in ranges for p in range(hi, lo - 1, -1)]
            assert parts == list(range(L, -2, -1))          # every backward part once, in execution order
            cover = sorted(slice_bounds(bounds, size, hi, lo) for hi, lo in ranges)
            assert cover[0][0] == 0 and cover[-1][1] == size
            assert all(a[1] == b[0] for a, b in zip(cover, cover[1:]))   # disjoint and gap-free


def test_lr_scheduler_matches_reference_rule():
    from css_b200.trainer.lr_scheduler import LRScheduler

    opt = argparse.Namespace(param_groups=[{"lr": 1e-3}])
    sched = LRScheduler(opt, argparse.Namespace(warmup=4, fixed=2, decay=0.1, min_lr=1e-9))
    assert sched.curr_lr == 1e-9
    seen = []
    for _ in range(9):
        sched.step(1.0)
        seen.append(opt.param_groups[0]["lr"])
    assert abs(seen[3] - 1e-3) < 1e-12 and seen[4] == 1e-3 and seen[5] == 1e-3
    assert abs(seen[6] - 1e-3 * np.exp(-0.1)) < 1e-12 and abs(seen[8] - 1e-3 * np.exp(-0.3)) < 1e-12
    state = sched.state_dict()
    other = LRScheduler(argparse.Namespace(param_groups=[{"lr": 1e-3}]), argparse.Namespace(warmup=0, fixed=0, decay=0.0, min_lr=0.0))
    other.load_state_dict(state)
    assert other.state_dict() == state


class _FakeEngine:
    """Stands in for ConformerTrainEngine in the exchange's host logic (flat buffer + part bounds)."""

    def __init__(self, num_blocks, rank):
        self.num_blocks = num_blocks
        self.part_bounds = [64 + 128 * l for l in range(num_blocks + 1)]
        self.size = self.part_bounds[-1] + 64
        self.flat_g = torch.arange(self.size, dtype=torch.float32) * (rank + 1)
        self.device = torch.device("cpu")

    def grad_slice(self, hi, lo):
        from css_b200.trainer.optim import slice_bounds

        a, b = slice_bounds(self.part_bounds, self.size, hi, lo)
        return self.flat_g[a:b]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from css_b200.trainer.optim import GradientExchange

    eng = _FakeEngine(5, rank)
    ex = GradientExchange(eng, buckets=3)
    for hi, lo in ex.ranges:
        ex.after_parts(hi, lo)
    ex.finish()
    want = torch.arange(eng.size, dtype=torch.float32) * sum(r + 1 for r in range(world))
    q.put(bool(torch.equal(eng.flat_g, want)))
    dist.barrier()
    dist.destroy_process_group()


def test_bucketed_allreduce_world2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=10) and q.get(timeout=10)
