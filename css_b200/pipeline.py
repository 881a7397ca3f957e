"""End-to-end CSS of one recording: the loop body of scripts/python/separate_libricss.py:65-82.

    separator.separate -> stitcher.get_stitch -> stitcher.get_connect -> beamformer.continuous_process

``CssPipeline`` wires the three executor objects together without leaving the device: the global
STFT is computed once and shared by the feature kernel and the MVDR kernel, the permutation bits and
the stitched masks stay in HBM, and only the two separated waveforms are read back.
"""
import os

import torch
import yaml

from .executor.beamformer import Beamformer
from .executor.separator import Separator
from .executor.stitcher import Stitcher

# conf/config_7ch.yaml and conf/config_1ch.yaml of the reference, as dicts (model_path is unused when a
# model object is handed in).
CONFIG_7CH = {
    "sampling_rate": 16000,
    "separation": {"batch_size": 1, "eval_hop": 0.8, "eval_win": 2.4, "frame_length": 512, "frame_shift": 256,
                   "model_path": "checkpoints/conformer_large.onnx", "ipd": "1,0;2,0;3,0;4,0;5,0;6,0",
                   "merge": True, "merge_threshold": 16},
    "stitching": {"eval_hop": 0.8, "eval_win": 2.4, "hop_size": 256, "n_fft": 512, "connector": "mask", "stitcher": "mask"},
    "beamforming": {"batch_size": 32, "type": "SoudenMVDRBeamformer", "hop_size": 256, "n_fft": 512, "eval_hop": 0.8,
                    "proceed_margin": 2, "eval_win": 2.4, "wta_thresh": 0.0001},
}
CONFIG_1CH = {
    "sampling_rate": 16000,
    "separation": {"batch_size": 1, "eval_hop": 0.8, "eval_win": 2.4, "frame_length": 512, "frame_shift": 256,
                   "model_path": "checkpoints/1ch_conformer_large.onnx"},
    "stitching": CONFIG_7CH["stitching"],
    "beamforming": CONFIG_7CH["beamforming"],
}


def load_config(path):
    with open(path) as f:
        return yaml.load(f, Loader=yaml.FullLoader)


class GraphedPipeline:
    """One captured recording shape of a CssPipeline (``CssPipeline.capture``)."""

    def __init__(self, graph, inp, out):
        self.graph, self.input, self.output = graph, inp, out

    def replay(self):
        self.graph.replay()
        return self.output


class CssPipeline:
    def __init__(self, cfg, model, device="cuda"):
        sr = cfg.get("sampling_rate", 16000)
        self.cfg = cfg
        self.sr = sr
        self.device = torch.device(device)
        self.separator = Separator(cfg["separation"], sr=sr, device=device, backend="b200", model=model)
        self.stitcher = Stitcher(cfg["stitching"], sr=sr)
        self.beamformer = Beamformer(cfg["beamforming"], sr=sr)

    def __call__(self, wav, beamform=None, return_intermediates=False):
        """wav: (D, T) fp32, host (pinned for an async copy) or device.  -> (2, T) separated waveforms (CUDA).

        With a single channel (no MVDR possible, as in the reference's 1-ch flow) the stitched masks
        [3][L][KP] are returned instead.
        """
        masks, mags = self.separator.separate(wav)     # a pinned host recording is streamed to the device in pieces
        wav = self.separator.last["wav"]
        swap = self.stitcher.stitch_device(mags, masks)
        stitched = self.stitcher.connect_device(swap, masks)
        if beamform is None:
            beamform = wav.shape[0] > 1
        out = None
        if beamform:
            views = tuple(stitched[c, :, :257].t() for c in range(3))
            for v in views:
                v._css_stitched = stitched
            o0, o1 = self.beamformer.continuous_process(wav, views, spec=self.separator.last["spec"])
            out = torch.cat([o0, o1], 0)
        if return_intermediates:
            return {"masks": masks, "mags": mags, "swap": swap, "stitched": stitched, "wavs": out,
                    "sims": self.stitcher.last_sims, "doa": self.separator.angle_merge_debug()}
        return out if beamform else stitched

    def capture(self, wav, pool=None):
        """Capture the whole path for recordings of ``wav``'s shape into ONE CUDA graph (about 800 kernel launches for a
        10-minute 7-channel session).  ``wav`` (device, (D, T) fp32) becomes the graph's static input buffer: write the next
        recording into ``g.input`` and call ``g.replay()``; the result is ``g.output`` (static as well: consume it, or copy
        it out, before the next replay).  One replay costs the host a few microseconds instead of the ~18 ms it takes to
        issue the launches one by one, which is within 1.5x of the GPU time: a host that is descheduled for a moment no
        longer starves the GPU.  Same kernels, same bits as the eager call (tests/test_gpu_pipeline.py)."""
        if not (isinstance(wav, torch.Tensor) and wav.is_cuda and wav.dtype == torch.float32 and wav.dim() == 2):
            raise RuntimeError("capture needs a (D, T) fp32 CUDA tensor as the static input buffer")
        self(wav)                                   # eager once: lazy set-up (weight planes, workspaces, placement plans)
        torch.cuda.synchronize(wav.device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, pool=pool):
            out = self(wav)
        return GraphedPipeline(g, wav, out)

    def _graph_slots(self, shape):
        """Two captured graphs (one per staging buffer) for recordings of this shape.  Each has its own memory pool: with a
        shared pool one graph's intermediates may sit where the other's OUTPUT sits, and a replay could then overwrite a result
        whose device-to-host copy is still in flight."""
        if getattr(self, "_graphs", None) is None:
            self._graphs = {}
        if shape not in self._graphs:
            a = self.capture(torch.zeros(shape, dtype=torch.float32, device=self.device))
            b = self.capture(torch.zeros(shape, dtype=torch.float32, device=self.device))
            self._graphs[shape] = (a, b)
        return self._graphs[shape]

    def run_many(self, recordings, outputs=None, graphs=None):
        """Separate a list of pinned host recordings (D, T_i) back to back, as the session loop of
        scripts/python/separate_libricss.py:65-82 does, with the transfers double-buffered around the compute: the H2D
        copy of recording i+1 and the D2H copy of result i-1 run on their own streams while recording i is separated.

        outputs: optional list of pinned host tensors (2, T_i) to receive the results; returned (allocated if None).
        graphs: replay captured CUDA graphs (``capture``) instead of issuing the launches when all recordings have one
        shape and are multi-channel; default on (``CSS_B200_GRAPHS=0`` or ``graphs=False``: eager).
        The call returns when every result has landed on the host.
        """
        if graphs is None:
            graphs = os.environ.get("CSS_B200_GRAPHS", "1") != "0"
        # a captured graph keeps its intermediates (about 6.6 GB per 10-minute 7-channel session, two graphs): recordings beyond
        # GRAPH_MAX_SECONDS are issued eagerly, and so is everything if a capture fails (out of memory, ...)
        if graphs and len(recordings) > 1 and recordings[0].shape[0] > 1 and all(r.shape == recordings[0].shape for r in recordings) \
                and recordings[0].shape[1] <= self.GRAPH_MAX_SECONDS * self.sr:
            try:
                self._graph_slots(tuple(recordings[0].shape))
            except Exception:
                self.release_graphs()
                torch.cuda.synchronize(self.device)
                return self._run_many_eager(recordings, outputs)
            return self._run_many_graphed(recordings, outputs)
        return self._run_many_eager(recordings, outputs)

    GRAPH_MAX_SECONDS = 1500

    def release_graphs(self):
        """Drop the captured graphs (and the memory their pools hold)."""
        self._graphs = None
        torch.cuda.empty_cache()

    def _run_many_graphed(self, recordings, outputs):
        dev = self.device
        main = torch.cuda.current_stream(dev)
        if getattr(self, "_h2d", None) is None:
            self._h2d, self._d2h = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        n = len(recordings)
        if outputs is None:
            outputs = [torch.empty(2, r.shape[1], dtype=torch.float32).pin_memory() for r in recordings]
        slots = self._graph_slots(tuple(recordings[0].shape))
        staged, freed, copied = [None] * n, [None, None], [None, None]

        def stage(i):
            slot = i % 2
            with torch.cuda.stream(self._h2d):
                if freed[slot] is not None:
                    self._h2d.wait_event(freed[slot])          # the previous recording in this input buffer has been separated
                slots[slot].input.copy_(recordings[i], non_blocking=True)
                staged[i] = torch.cuda.Event()
                staged[i].record(self._h2d)

        self._h2d.wait_stream(main)
        stage(0)
        done = []
        for i in range(n):
            slot = i % 2
            if i + 1 < n:
                stage(i + 1)
            main.wait_event(staged[i])
            if copied[slot] is not None:
                main.wait_event(copied[slot])                  # the previous result in this output buffer has left
            slots[slot].replay()
            ev = torch.cuda.Event()
            ev.record(main)
            freed[slot] = ev
            with torch.cuda.stream(self._d2h):
                self._d2h.wait_event(ev)
                outputs[i].copy_(slots[slot].output, non_blocking=True)
                fin = torch.cuda.Event()
                fin.record(self._d2h)
                copied[slot] = fin
                done.append(fin)
        for fin in done:
            fin.synchronize()
        return outputs

    def _run_many_eager(self, recordings, outputs=None):
        dev = self.device
        main = torch.cuda.current_stream(dev)
        if getattr(self, "_h2d", None) is None:
            self._h2d, self._d2h = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
        n = len(recordings)
        if outputs is None:
            outputs = [torch.empty(2, r.shape[1], dtype=torch.float32).pin_memory() for r in recordings]
        # the two device staging buffers live with the pipeline: allocating them per call (on the copy stream) showed up as
        # occasional ~0.5 s stalls on the GPU boxes
        if getattr(self, "_stage_bufs", None) is None:
            self._stage_bufs = [None, None]
        bufs, staged, freed = self._stage_bufs, [None] * n, [None, None]

        def stage(i):
            r = recordings[i]
            slot = i % 2
            with torch.cuda.stream(self._h2d):
                if freed[slot] is not None:
                    self._h2d.wait_event(freed[slot])          # the previous user of this buffer has been separated
                if bufs[slot] is None or bufs[slot].shape != r.shape:
                    bufs[slot] = torch.empty(r.shape, dtype=torch.float32, device=dev)
                bufs[slot].copy_(r, non_blocking=True)
                staged[i] = torch.cuda.Event()
                staged[i].record(self._h2d)

        if n:
            stage(0)
        done = []
        for i in range(n):
            if i + 1 < n:
                stage(i + 1)
            main.wait_event(staged[i])
            out = self(bufs[i % 2])
            ev = torch.cuda.Event()
            ev.record(main)
            freed[i % 2] = ev
            with torch.cuda.stream(self._d2h):
                self._d2h.wait_event(ev)
                outputs[i].copy_(out, non_blocking=True)
                out.record_stream(self._d2h)
                fin = torch.cuda.Event()
                fin.record(self._d2h)
                done.append(fin)
        for fin in done:
            fin.synchronize()
        return outputs
