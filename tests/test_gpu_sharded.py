"""One recording sharded over ranks (BASELINE configs[4]): the sharded result must equal the
single-GPU pipeline bit for bit -- permutation states exactly, waveforms exactly."""
import os

import pytest
import torch

from tests.helpers import cfg7, golden_conformer

pytestmark = pytest.mark.gpu


def _flip_some_chunks(separator):
    """A random-init network never swaps its outputs, so the test swaps the two speaker masks of pseudo-randomly chosen
    chunks (chosen from the chunk's own magnitudes, hence identically in the single and the sharded run)."""
    orig = separator.separate

    def separate(s):
        masks, mags = orig(s)
        key = ((mags.mean(dim=(1, 2)) * 1e4).floor().long() % 2) == 1
        masks[key] = masks[key][..., [1, 0, 2]]
        return masks, mags

    separator.separate = separate


def _single(wav, dev):
    from css_b200.pipeline import CssPipeline

    pipe = CssPipeline(cfg7(), golden_conformer(num_blocks=1).to(dev), device=dev)
    _flip_some_chunks(pipe.separator)
    got = pipe(wav, return_intermediates=True)
    return got["wavs"].clone(), pipe.stitcher.last_state.clone()


@pytest.mark.parametrize("world", [2, 3, 8])
def test_sharded_equals_single_emulated(cuda, world):
    """70 s, 85 chunks (crosses two 32-chunk MVDR batch seams): ranks emulated one after the other on one GPU."""
    from css_b200.dist import ShardedCssPipeline
    from css_b200.synth import make_recording

    wav = make_recording(70, channels=7, seed=17)
    ref_wav, ref_state = _single(wav, cuda)
    sh = ShardedCssPipeline(cfg7(), golden_conformer(num_blocks=1).to(cuda), cuda, 0, 1)
    _flip_some_chunks(sh.pipe.separator)
    out, state = sh.emulate(wav, world)
    assert torch.equal(state.cpu(), ref_state.cpu())
    assert ref_state.sum() > 0, "test recording has no permutation flips"
    assert torch.equal(out, ref_wav)


def test_short_recording_more_ranks_than_chunks(cuda):
    from css_b200.dist import ShardedCssPipeline
    from css_b200.synth import make_recording

    wav = make_recording(4.1, channels=7, seed=3)          # 3 chunks on 8 ranks: 5 ranks idle
    ref_wav, _ = _single(wav, cuda)
    sh = ShardedCssPipeline(cfg7(), golden_conformer(num_blocks=1).to(cuda), cuda, 0, 1)
    _flip_some_chunks(sh.pipe.separator)
    out, _ = sh.emulate(wav, 8)
    assert torch.equal(out, ref_wav)


def _nccl_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import torch.distributed as dist

    from css_b200.dist import ShardedCssPipeline, init_from_env
    from css_b200.synth import make_recording

    init_from_env("nccl")
    dev = torch.device("cuda", rank)
    wav = make_recording(70, channels=7, seed=17)
    sh = ShardedCssPipeline(cfg7(), golden_conformer(num_blocks=1).to(dev), dev, rank, world)
    _flip_some_chunks(sh.pipe.separator)
    out = sh(wav)
    if rank == 0:
        ref_wav, _ = _single(wav, dev)
        q.put(bool(torch.equal(out, ref_wav)))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_equals_single_nccl(cuda):
    """Real ranks over NCCL when the box has at least two GPUs."""
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 30500 + os.getpid() % 2000
    procs = [ctx.Process(target=_nccl_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    assert q.get(timeout=10) is True
