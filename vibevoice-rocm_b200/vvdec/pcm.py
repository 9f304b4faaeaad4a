"""PCM16 output step and the audio streamer — what happens to decode()'s waveform next.

Reference behaviour this mirrors:
  * ``convert_to_16_bit_wav`` (main.py:2956-2979, demo/gradio_demo.py:1205): peak-normalise a chunk if max|x| > 1, scale
    by 32767, truncate to int16 — numpy, on the host, once per streamed chunk (main.py:937-951).
  * ``AudioStreamer.put`` (vibevoice/modular/streamer.py:41-55): for every sample of the step a ``sample_idx.item()`` and
    an ``audio_chunks[i].detach().cpu()`` — 2·B host synchronisations and B pageable copies per frame step.

Here: one ``vv_pcm16`` launch converts the whole step's chunks on the device (bit-exact with the numpy function), ONE
asynchronous copy moves the int16 block into a pinned ring slot (half the bytes of fp32), and one event wait hands the
per-sample int16 arrays to the same per-sample queues.  No CPU / numpy fallback: tensors must be on the CUDA device.
"""
from __future__ import annotations

import ctypes as C
from queue import Queue
from typing import Optional

import numpy as np
import torch

from . import _capi

PCM_ROW_MAX = 8192          # rows up to this many samples convert in one launch (csrc/pcm16.cuh)


def _rows(wave: torch.Tensor):
    """View a waveform tensor as [B, L] rows without copying; returns (B, L, row_stride)."""
    if wave.dim() == 3 and wave.shape[1] == 1:          # decode() output [B, 1, L]
        B, L, rs = wave.shape[0], wave.shape[2], wave.stride(0)
        inner = wave.stride(2)
    elif wave.dim() == 2:
        B, L, rs = wave.shape[0], wave.shape[1], wave.stride(0)
        inner = wave.stride(1)
    elif wave.dim() == 1:
        B, L, rs = 1, wave.shape[0], wave.shape[0]
        inner = wave.stride(0)
    else:
        raise ValueError(f"waveform must be [L], [B, L] or [B, 1, L]; got {tuple(wave.shape)}")
    if B == 1:
        rs = L                                          # the stride of a size-1 axis carries no information
    return B, L, rs, inner


def pcm16(wave: torch.Tensor, return_peak: bool = False):
    """int16 PCM of a waveform, row by row, exactly as ``convert_to_16_bit_wav`` would produce for each row.

    wave: CUDA tensor [L], [B, L] or [B, 1, L], float32 or bfloat16 (bf16 is widened exactly, as main.py:939-942 does).
    Returns an int16 tensor of the same shape (and the per-row peaks max|x| as fp32 [B] when ``return_peak``).
    """
    if not isinstance(wave, torch.Tensor):
        raise ValueError("wave must be a torch.Tensor")
    if not wave.is_cuda:
        raise RuntimeError("wave must live on the CUDA device; vvdec has no CPU path")
    if wave.dtype not in (torch.float32, torch.bfloat16):
        raise ValueError(f"wave must be float32 or bfloat16, got {wave.dtype}")
    B, L, rs, inner = _rows(wave)
    out = torch.empty(wave.shape, dtype=torch.int16, device=wave.device)
    peak = torch.empty((max(B, 1),), dtype=torch.float32, device=wave.device)
    if B == 0 or L == 0:
        return (out, peak[:B].zero_()) if return_peak else out
    if inner != 1 or (B > 1 and rs < L):
        wave = wave.contiguous()
        B, L, rs, inner = _rows(wave)
    lib = _capi.load()
    with torch.cuda.device(wave.device):
        stream = torch.cuda.current_stream(wave.device).cuda_stream
        rc = lib.vv_pcm16(C.c_void_p(wave.data_ptr()), _capi.VV_BF16 if wave.dtype == torch.bfloat16 else _capi.VV_F32,
                          int(B), int(L), int(rs), C.c_void_p(out.data_ptr()), C.c_void_p(peak.data_ptr()),
                          C.c_void_p(stream))
    _capi.check(rc, "vv_pcm16")
    return (out, peak) if return_peak else out


class B200PcmStreamer:
    """``AudioStreamer`` (vibevoice/modular/streamer.py:13-83) with the PCM16 step moved onto the device.

    Same constructor, ``put(audio_chunks, sample_indices)``, ``end(sample_indices=None)``, ``get_stream(i)`` and batch
    iteration; queue items are **int16 numpy arrays** (what main.py:951 computes from each float chunk) instead of float
    CPU tensors.  Per ``put``: one kernel launch, one pinned asynchronous copy, one event wait — against 2·B syncs.
    """

    def __init__(self, batch_size: int, stop_signal=None, timeout: Optional[float] = None, ring_slots: int = 4):
        self.batch_size = batch_size
        self.stop_signal = stop_signal
        self.timeout = timeout
        self.audio_queues = [Queue() for _ in range(batch_size)]
        self.finished_flags = [False for _ in range(batch_size)]
        self.sample_indices_map = {}
        self._ring = [None] * max(1, ring_slots)        # pinned int16 staging, grown on demand
        self._turn = 0
        self.puts = 0                                   # bookkeeping the tests / bench read
        self.d2h_bytes = 0

    def _slot(self, n: int) -> torch.Tensor:
        i = self._turn % len(self._ring)
        self._turn += 1
        buf = self._ring[i]
        if buf is None or buf.numel() < n:
            buf = torch.empty((max(n, 1),), dtype=torch.int16, pin_memory=True)
            self._ring[i] = buf
        return buf[:n]

    def put(self, audio_chunks: torch.Tensor, sample_indices):
        """audio_chunks: CUDA tensor [n, ...] (one chunk per row); sample_indices: the batch positions they belong to
        (tensor or list; a CUDA tensor costs one ``tolist()`` instead of the reference's ``item()`` per sample)."""
        if isinstance(sample_indices, torch.Tensor):
            idx = [int(i) for i in sample_indices.reshape(-1).tolist()]
        else:
            idx = [int(i) for i in sample_indices]
        n = audio_chunks.shape[0]
        if len(idx) != n:
            raise AssertionError("len(sample_indices) must match the number of chunks")
        if n == 0:
            return
        rows = audio_chunks.reshape(n, -1)
        pcm = pcm16(rows)                               # [n, L] int16 on the device
        host = self._slot(pcm.numel()).view(n, -1)
        host.copy_(pcm, non_blocking=True)
        done = torch.cuda.Event()
        done.record(torch.cuda.current_stream(pcm.device))
        done.synchronize()
        self.puts += 1
        self.d2h_bytes += pcm.numel() * 2
        arr = host.numpy()
        for i, q in enumerate(idx):
            if q < self.batch_size and not self.finished_flags[q]:
                self.audio_queues[q].put(arr[i].copy(), timeout=self.timeout)    # the ring slot is reused later

    def end(self, sample_indices=None):
        if sample_indices is None:
            targets = range(self.batch_size)
        elif isinstance(sample_indices, torch.Tensor):
            targets = [int(i) for i in sample_indices.reshape(-1).tolist()]
        else:
            targets = [int(i) for i in sample_indices]
        for q in targets:
            if q < self.batch_size and not self.finished_flags[q]:
                self.audio_queues[q].put(self.stop_signal, timeout=self.timeout)
                self.finished_flags[q] = True

    def get_stream(self, sample_idx: int):
        if sample_idx >= self.batch_size:
            raise ValueError(f"Sample index {sample_idx} exceeds batch size {self.batch_size}")
        return _SampleIterator(self, sample_idx)

    def __iter__(self):
        return _BatchIterator(self)


def _is_stop(value, stop_signal) -> bool:
    return value is stop_signal or (not isinstance(value, np.ndarray) and value == stop_signal)


class _SampleIterator:
    def __init__(self, streamer: B200PcmStreamer, sample_idx: int):
        self.streamer, self.sample_idx = streamer, sample_idx

    def __iter__(self):
        return self

    def __next__(self):
        value = self.streamer.audio_queues[self.sample_idx].get(timeout=self.streamer.timeout)
        if _is_stop(value, self.streamer.stop_signal):
            raise StopIteration()
        return value


class _BatchIterator:
    """Yields {sample_idx: int16 chunk} for whatever is ready, until every sample has ended (streamer.py:102-150)."""

    def __init__(self, streamer: B200PcmStreamer):
        self.streamer = streamer
        self.active = set(range(streamer.batch_size))

    def __iter__(self):
        return self

    def __next__(self):
        import queue
        import time
        while self.active:
            chunks, ended = {}, set()
            for i in self.active:
                try:
                    value = self.streamer.audio_queues[i].get(block=False)
                except queue.Empty:
                    continue
                if _is_stop(value, self.streamer.stop_signal):
                    ended.add(i)
                else:
                    chunks[i] = value
            self.active -= ended
            if chunks:
                return chunks
            if self.active:
                time.sleep(0.01)
        raise StopIteration()
