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


@pytest.mark.parametrize("world", [2, 13])
def test_sharded_60s_q4_spill_crosses_a_rank_boundary(cuda, world):
    """60 s = 72 chunks with T a multiple of the hop: chunks 42 and 66 take the `en == T` branch (A.6 Q4) and place 6,144
    samples into the next chunk's region.  With 13 ranks chunk 66 is the last chunk of a rank's range, so that spill lands in
    the right neighbour's sample range and has to be exchanged; the result still equals the single-GPU pipeline bit for bit."""
    from css_b200.dist import ShardedCssPipeline, chunk_ranges
    from css_b200.synth import make_recording

    wav = make_recording(60, channels=7, seed=23)
    if world == 13:
        assert 67 in [c1 for _, c1 in chunk_ranges(72, world)]
    ref_wav, ref_state = _single(wav, cuda)
    sh = ShardedCssPipeline(cfg7(), golden_conformer(num_blocks=1).to(cuda), cuda, 0, 1)
    _flip_some_chunks(sh.pipe.separator)
    assert sh.spill == 6144
    out, state = sh.emulate(wav, world)
    assert torch.equal(state.cpu(), ref_state.cpu())
    assert torch.equal(out, ref_wav)
    # world == 1 through the real entry point: one rank owns every sample
    res, info = sh(wav, return_intermediates=True)
    assert torch.equal(res, ref_wav) and info["samples"] == (0, wav.shape[1])
    piece, (s0, s1) = sh(wav, gather=False)
    assert (s0, s1) == (0, wav.shape[1]) and torch.equal(piece, ref_wav)


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
    piece, (s0, s1) = sh(wav, gather=False)
    if rank == 0:
        ref_wav, _ = _single(wav, dev)
        q.put(bool(torch.equal(out, ref_wav)) and bool(torch.equal(piece, ref_wav[:, s0:s1])) and s0 == 0)
    else:
        assert out is None and s1 == wav.shape[1]
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
