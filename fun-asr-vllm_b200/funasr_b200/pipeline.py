"""End-to-end audio path: host PCM -> pinned staging -> H2D -> fbank/LFR/CMVN -> SAN-M encoder -> adaptor -> bf16
``prompt_embeds`` rows -> D2H.  This is what ``bench.py`` times as ``e2e`` and what a batched driver
(``/root/reference/infer.py:444-475``) reduces to once the three seams are patched: the three calls
``extract_fbank`` / ``audio_encoder`` / ``audio_adaptor`` plus the per-utterance ``cat`` become one packed pass with
no padded intermediates.  vLLM wants CPU tensors for ``prompt_embeds`` (VLLM/renderers/base.py:755-758), hence the D2H."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib
from .encoder import AudioTower
from .frontend import WavFrontend
from .ingest import PCM_S16


class AudioPipeline:
    def __init__(self, frontend: WavFrontend, tower: AudioTower, device=None, private_plan: bool = False):
        """``private_plan``: run the single-pass schedule on a plan of this pipeline's own (own workspace, the tower's weights)
        instead of the tower's: needed when several pipelines over one tower have work in flight at once (StreamingPipeline)."""
        self.frontend, self.tower = frontend, tower
        self.device = torch.device(device) if device is not None else tower.encoder._device()
        self._pin_in = self._pin_out = self._dev_in = None
        self._private, self._own, self._own_of = bool(private_plan), None, None

    def _own_plan(self):
        plan = self.tower.plan()
        if self._own_of is not plan:
            self._own, self._own_of = plan.twin(), plan
        return self._own

    def _staging(self, n_in: int, n_out_rows: int, dim: int):
        if self._pin_in is None or self._pin_in.numel() < n_in:
            self._pin_in = torch.empty(n_in, dtype=torch.float32).pin_memory()
            self._dev_in = torch.empty(n_in, dtype=torch.float32, device=self.device)
        if self._pin_out is None or self._pin_out.shape[0] < n_out_rows or self._pin_out.shape[1] != dim:
            self._pin_out = torch.empty((n_out_rows, dim), dtype=torch.bfloat16).pin_memory()

    @torch.no_grad()
    def features(self, wav_dev: torch.Tensor, starts: np.ndarray, n_samples: np.ndarray):
        """Device-resident stage 1: packed PCM -> packed [sum T, 560] fp32 features."""
        return self.frontend.forward_packed(wav_dev, starts, n_samples)

    @torch.no_grad()
    def embed_device(self, wav_dev: torch.Tensor, starts: np.ndarray, n_samples: np.ndarray, out: torch.Tensor | None = None):
        """Device-resident whole path: packed PCM on the GPU -> packed bf16 ``[sum T, llm_dim]`` on the GPU."""
        feats, cu = self.frontend.forward_packed(wav_dev, starts, n_samples)
        if self._private:
            if out is None:
                out = torch.empty((int(cu[-1]), self.tower.adaptor.llm_dim), dtype=torch.bfloat16, device=feats.device)
            self._own_plan().encode_adapt(feats, 0, np.diff(cu), out, 0)
        else:
            out, _ = self.tower.forward_packed(feats, np.diff(cu), out=out)
        return out, cu

    def _twins(self, sizes):
        """One plan per piece (own workspaces, the tower's weights) + auxiliary streams for the overlapped schedule.  Each piece
        runs on ONE stream: the pieces already are the streams (measured: letting each half split again doubles the
        launches and loses 5 ms at cfg5, 5 ms at cfg3)."""
        plan = self.tower.plan()
        if getattr(self, "_twin_of", None) is not plan or len(self._twin_plans) != len(sizes):
            self._twin_plans = tuple(plan.twin(max(int(n), 128)) for n in sizes)
            for t in self._twin_plans:
                t.set_streams(1)
            self._twin_of = plan
            self._aux = torch.cuda.Stream(self.device)
            self._d2h = torch.cuda.Stream(self.device)
            self._ev_fork, self._ev_join = torch.cuda.Event(), torch.cuda.Event()
            self._ev_done = [torch.cuda.Event() for _ in sizes]
        return self._twin_plans

    def _pcm16_meta(self, starts: np.ndarray, n: np.ndarray) -> int:
        """Utterance starts (int64) and lengths (int32) of an int16 batch on the device, for ``fa_pcm_to_mono_f32``: one small
        copy from a pinned buffer kept across calls (an event guards its re-use).  Returns the device address."""
        B = len(n)
        if getattr(self, "_meta_pin", None) is None or self._meta_pin.numel() < 3 * B:
            self._meta_pin = torch.empty(3 * max(B, 64), dtype=torch.int32).pin_memory()
            self._meta_dev = torch.empty(self._meta_pin.numel(), dtype=torch.int32, device=self.device)
            self._meta_ev = torch.cuda.Event()
        else:
            self._meta_ev.synchronize()  # the previous call's copy has read the buffer
        h = self._meta_pin.numpy()
        h[: 2 * B] = np.ascontiguousarray(starts, np.int64).view(np.int32)
        h[2 * B: 3 * B] = n.astype(np.int32)
        self._meta_dev[: 3 * B].copy_(self._meta_pin[: 3 * B], non_blocking=True)
        self._meta_ev.record(torch.cuda.current_stream(self.device))
        return self._meta_dev.data_ptr()

    def _pcm16_to_f32(self, raw: torch.Tensor, dev: torch.Tensor, meta: int, B: int, b0: int, b1: int, max_len: int):
        """int16 samples of utterances [b0, b1) in ``raw`` -> f32 / 32768 at the same offsets of ``dev`` (what
        ``torchaudio.load`` / ``.float()`` hand the reference, /root/reference/infer.py:320-325), on the current stream."""
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().fa_pcm_to_mono_f32(raw.data_ptr(), PCM_S16, 1, 1.0 / 32768.0, meta + 8 * b0, meta + 8 * B + 4 * b0,
                                                      dev.data_ptr(), meta + 8 * b0, b1 - b0, int(max_len),
                                                      _lib.stream_ptr(self.device)), "fa_pcm_to_mono_f32")

    @staticmethod
    def _cut(T: np.ndarray) -> int:
        """Utterance index that halves the cost 0.4657 T + 1.5155e-4 T^2, 0 < cut < B."""
        c = np.cumsum(0.4657152 * T + 1.51552e-4 * T.astype(np.float64) ** 2)
        return int(min(max(np.searchsorted(c, 0.5 * c[-1]) + 1, 1), len(T) - 1))

    @staticmethod
    def _cuts(T: np.ndarray, pieces: int) -> list[int]:
        """Utterance boundaries 0 = c_0 < c_1 < ... < c_k = B of k <= pieces contiguous pieces of equal cost."""
        if pieces == 2:
            return [0, AudioPipeline._cut(T), len(T)]
        c = np.cumsum(0.4657152 * T + 1.51552e-4 * T.astype(np.float64) ** 2)
        cuts = [0]
        for i in range(1, pieces):
            b = int(min(max(np.searchsorted(c, i * c[-1] / pieces) + 1, cuts[-1] + 1), len(T) - 1))
            if b > cuts[-1]:
                cuts.append(b)
        if cuts[-1] != len(T):
            cuts.append(len(T))
        return cuts

    @torch.no_grad()
    def run_pinned(self, host_pcm: torch.Tensor, starts: np.ndarray, n_samples: np.ndarray, out_host: torch.Tensor,
                   overlap: bool | None = None, pieces: int = 2):
        """Pinned packed PCM on the host -> pinned bf16 ``[sum T, llm_dim]`` on the host: H2D, the whole path, D2H, all
        enqueued on (or joined back into) the current stream (no sync here; the caller synchronises once).

        ``overlap`` (default: on for batches of >= 8 utterances and >= 8192 rows): the batch is cut on an utterance boundary
        into two halves of equal cost, each with its own stream and workspace.  The copy engine then moves the second half's
        PCM while the first half computes, and the first half's embeddings go back to the host while the second half
        finishes; utterances are independent, so the rows are bit-identical to the single-pass schedule.

        ``pieces`` > 2 cuts the batch further (alternating between the two streams, copies back on a third): measured slower
        at cfg5 (45.7-46.5 ms with 2, 50.1 / 48.6 / 51.6 ms with 3 / 4 / 6: profiles/r02/e2e_pieces_r2bm.md), the smaller GEMMs
        cost more than the shorter first copy in and last copy out save; kept for batches far above an hour.

        ``host_pcm`` may be int16 (mono, what a 16 kHz wav file holds; same ``starts``): half the bytes cross PCIe and the
        conversion to the f32 samples ``torchaudio.load`` would have produced (x / 32768, exact) runs on the GPU."""
        total = host_pcm.numel()
        pcm16 = host_pcm.dtype == torch.int16
        if not pcm16 and host_pcm.dtype != torch.float32:
            raise TypeError(f"host PCM must be float32 or int16, not {host_pcm.dtype}")
        if pcm16 and (getattr(self, "_dev_raw", None) is None or self._dev_raw.numel() < total):
            self._dev_raw = torch.empty(total, dtype=torch.int16, device=self.device)
        if self._dev_in is None or self._dev_in.numel() < total:
            self._dev_in = torch.empty(total, dtype=torch.float32, device=self.device)
        dev = self._dev_in[:total]
        starts, n_samples = np.asarray(starts, np.int64), np.asarray(n_samples, np.int64)
        _, T = self.frontend.lengths(n_samples)
        rows = int(T.sum())
        if overlap is None:
            overlap = len(T) >= 8 and rows >= 8192
        B = len(T)
        meta = self._pcm16_meta(starts, n_samples) if pcm16 else 0
        if not overlap or len(T) < 2:
            if pcm16:
                self._dev_raw[:total].copy_(host_pcm, non_blocking=True)
                self._pcm16_to_f32(self._dev_raw, dev, meta, B, 0, B, n_samples.max())
            else:
                dev.copy_(host_pcm, non_blocking=True)
            out, cu = self.embed_device(dev, starts, n_samples)
            out_host[: out.shape[0]].copy_(out, non_blocking=True)
            return cu
        cuts = self._cuts(T, max(2, int(pieces)))
        rcum = np.concatenate([[0], np.cumsum(T)]).astype(np.int64)
        scut = [int(starts[c]) if c < len(T) else total for c in cuts]
        plans = self._twins([rcum[c1] - rcum[c0] for c0, c1 in zip(cuts[:-1], cuts[1:])])
        main = torch.cuda.current_stream(self.device)
        dim = self.tower.adaptor.llm_dim
        if getattr(self, "_dev_out", None) is None or self._dev_out.shape[0] < rows or self._dev_out.shape[1] != dim:
            self._dev_out = torch.empty((rows, dim), dtype=torch.bfloat16, device=self.device)
        out = self._dev_out[:rows]
        self._ev_fork.record(main)
        self._aux.wait_event(self._ev_fork)
        self._d2h.wait_event(self._ev_fork)
        # piece k on stream k % 2; (stream, utterances, first sample, end sample, first row, end row)
        parts = [((main, self._aux)[k % 2], slice(c0, c1), scut[k], scut[k + 1], int(rcum[c0]), int(rcum[c1]))
                 for k, (c0, c1) in enumerate(zip(cuts[:-1], cuts[1:]))]
        for st, _, s0, s1, _, _ in parts:  # all copies queued first: the DMA engine runs them back to back
            with torch.cuda.stream(st):
                (self._dev_raw if pcm16 else dev)[s0:s1].copy_(host_pcm[s0:s1], non_blocking=True)
        two = len(parts) == 2
        for k, (st, sl, s0, s1, r0, r1) in enumerate(parts):
            with torch.cuda.stream(st):
                if pcm16:
                    self._pcm16_to_f32(self._dev_raw, dev, meta, B, sl.start, sl.stop, n_samples[sl].max())
                feats, cu_h = self.frontend.forward_packed(dev[s0:s1], starts[sl] - s0, n_samples[sl])
                plans[k].encode_adapt(feats, 0, np.diff(cu_h), out[r0:r1], 0)
                if two:  # each stream's only piece: its copy back follows it on the same stream
                    out_host[r0:r1].copy_(out[r0:r1], non_blocking=True)
                else:  # a later piece follows on this stream: the copy back goes to a stream of its own
                    self._ev_done[k].record(st)
            if not two:
                self._d2h.wait_event(self._ev_done[k])
                with torch.cuda.stream(self._d2h):
                    out_host[r0:r1].copy_(out[r0:r1], non_blocking=True)
        self._ev_join.record(self._aux)
        main.wait_event(self._ev_join)
        if not two:
            self._ev_join.record(self._d2h)
            main.wait_event(self._ev_join)
        cu = np.zeros(len(T) + 1, np.int32)
        cu[1:] = np.cumsum(T)
        return cu

    @torch.no_grad()
    def __call__(self, wavs: list) -> tuple[list[torch.Tensor], np.ndarray]:
        """list of 1-D fp32 CPU waveforms -> list of CPU bf16 ``[T_b, llm_dim]`` tensors (views of one pinned buffer,
        valid until the next call) + lengths.  One H2D, one D2H, one stream sync."""
        lens = np.asarray([int(w.numel()) for w in wavs], dtype=np.int64)
        starts = np.zeros(len(wavs), np.int64)
        starts[1:] = np.cumsum((lens[:-1] + 3) // 4 * 4)
        total = int(starts[-1] + (lens[-1] + 3) // 4 * 4)
        _, T = self.frontend.lengths(lens)
        rows = int(T.sum())
        dim = self.tower.adaptor.llm_dim
        self._staging(total, rows, dim)
        host = self._pin_in[:total]
        for w, s, n in zip(wavs, starts, lens):
            host[s:s + n] = w.reshape(-1)
        dev = self._dev_in[:total]
        dev.copy_(host, non_blocking=True)
        out, cu = self.embed_device(dev, starts, lens)
        res = self._pin_out[:rows]
        res.copy_(out, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return [res[cu[b]:cu[b + 1]] for b in range(len(wavs))], np.diff(cu)


class StreamingPipeline:
    """A serving loop over ``AudioPipeline.run_pinned``: ``lanes`` pipelines (own streams, workspaces and device buffers, the
    tower's weights shared) take the submitted batches in turn, so that the copy in of batch n + 1 and the copy back of batch
    n - 1 run under the compute of batch n; every batch still makes its own H2D and D2H.  Results are bit-identical to
    ``run_pinned`` (utterances and batches are independent).

        sp = StreamingPipeline(frontend, tower, "cuda:0")
        for host_pcm, starts, n, out_host in batches:      # pinned buffers; a buffer pair may be re-used after its ticket is done
            ticket = sp.submit(host_pcm, starts, n, out_host)
        sp.drain()                                          # or ticket.synchronize() per batch
    """

    def __init__(self, frontend: WavFrontend, tower: AudioTower, device=None, lanes: int = 2, overlap: bool | None = None):
        self.overlap = overlap  # run_pinned's schedule inside a lane (None: its default, two halves from 8192 rows on)
        self.pipes = [AudioPipeline(frontend, tower, device, private_plan=True) for _ in range(max(1, int(lanes)))]
        self.device = self.pipes[0].device
        self.streams = [torch.cuda.Stream(self.device) for _ in self.pipes]
        self._n = 0

    @torch.no_grad()
    def submit(self, host_pcm: torch.Tensor, starts, n_samples, out_host: torch.Tensor, **kw) -> torch.cuda.Event:
        """Enqueue one batch (arguments of ``AudioPipeline.run_pinned``); returns the event that marks ``out_host`` complete."""
        kw.setdefault("overlap", self.overlap)
        k = self._n % len(self.pipes)
        self._n += 1
        st = self.streams[k]
        st.wait_stream(torch.cuda.current_stream(self.device))  # whatever produced the inputs on the caller's stream
        with torch.cuda.stream(st):
            self.pipes[k].run_pinned(host_pcm, starts, n_samples, out_host, **kw)
            ev = torch.cuda.Event()
            ev.record(st)
        return ev

    def drain(self) -> None:
        for st in self.streams:
            st.synchronize()
