"""One generation-loop frame step on the device: acoustic streaming decode -> semantic-tokenizer
streaming encode of the chunk just decoded -> PCM16 of the chunk, captured as ONE CUDA graph per (stream-count bucket,
frames per step), with ``sample_indices`` consumed on the device.

What it replaces in the reference's ``generate()`` (vibevoice/modular/modeling_vibevoice_inference.py), per step:
  :635-641  ``acoustic_tokenizer.decode(scaled_latent, cache=acoustic_cache, sample_indices=diffusion_indices, use_cache=True)``
  :656-662  ``semantic_tokenizer.encode(audio_chunk, cache=semantic_cache, sample_indices=diffusion_indices, use_cache=True).mean``
  :651-653 + main.py:937-951  the streamer hand-off and ``convert_to_16_bit_wav`` of every chunk
  :540-544  ``acoustic_cache.set_to_zero(idx); semantic_cache.set_to_zero(idx)`` at a speech end  -> ``FrameStep.set_to_zero``
and the host synchronisations on that path: 68 ``sample_indices.tolist()`` per frame inside the reference cache
(modular_vibevoice_tokenizer.py:204,230,238), the per-sample ``.item()`` loop (:644-648) and the streamer's ``.cpu()``.
A step here costs: 3 small copies into the graph's input buffers, ONE graph launch (about 160 kernels), 2-4 small copies
out -- and no host synchronisation at all (tests run it under ``torch.cuda.set_sync_debug_mode("error")``).

The number of active streams changes from step to step (``diffusion_indices`` is a subset of the batch, :571); a graph
has a fixed shape, so a step is padded up to the next bucket with rows whose slot id is -1: the kernels decode them with
a zero context and save nothing (include/vvdec.h, vv_state_check), and their outputs are never returned.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch

from .decoder import B200AcousticDecoder, B200StreamingCache
from .pcm import pcm16


@dataclass
class FrameStepOutput:
    audio: torch.Tensor                       # [n, 1, T*hop], the latents' dtype      (audio_chunk, :635)
    semantic: Optional[torch.Tensor]          # [n, T, semantic vae_dim]               (semantic_features, :656-662)
    pcm: Optional[torch.Tensor]               # [n, T*hop] int16                       (convert_to_16_bit_wav per chunk)
    peak: Optional[torch.Tensor]              # [n] fp32 max|x| of each chunk


class FrameStep:
    def __init__(self, decoder: B200AcousticDecoder, semantic_encoder=None, pcm: bool = True, n_slots: int = 256,
                 bucket: int = 32, use_graph: bool = True, acoustic_cache: Optional[B200StreamingCache] = None,
                 semantic_cache: Optional[B200StreamingCache] = None):
        if decoder.strict_reference_quirks or (semantic_encoder is not None and semantic_encoder.strict_reference_quirks):
            raise ValueError("FrameStep keeps sample indices on the device; strict_reference_quirks needs a host view")
        self.dec, self.enc, self.with_pcm = decoder, semantic_encoder, bool(pcm)
        self.bucket, self.use_graph = max(1, int(bucket)), bool(use_graph)
        self.acoustic_cache = acoustic_cache or decoder.new_cache(n_slots)
        self.semantic_cache = (semantic_cache or semantic_encoder.new_cache(n_slots)) if semantic_encoder is not None else None
        self.n_slots = self.acoustic_cache.n_slots
        self._plans: Dict[Tuple[int, int, torch.dtype], dict] = {}
        self.graph_launches = 0

    # -- reference cache API, applied to both tokenizers' state (:540-544) ----------------------------------------------
    def set_to_zero(self, sample_indices) -> None:
        self.acoustic_cache.set_to_zero(sample_indices)
        if self.semantic_cache is not None:
            self.semantic_cache.set_to_zero(sample_indices)

    def check(self) -> None:
        """Raise IndexError if an earlier step carried a sample index outside [0, n_slots) (see B200StreamingCache.check)."""
        self.acoustic_cache.check()
        if self.semantic_cache is not None:
            self.semantic_cache.check()

    # -- one bucket = one set of static buffers + one captured graph ----------------------------------------------------
    def _run(self, plan: dict) -> None:
        lat, idx = plan["lat"], plan["idx"]
        audio = self.dec.decode(lat, cache=self.acoustic_cache, sample_indices=idx, use_cache=True, workspace=plan["ws_dec"])
        plan["audio"] = audio
        if self.enc is not None:
            plan["semantic"] = self.enc.encode(audio, cache=self.semantic_cache, sample_indices=idx, use_cache=True,
                                               workspace=plan["ws_enc"]).mean
        if self.with_pcm:
            plan["pcm"], plan["peak"] = pcm16(audio.reshape(audio.shape[0], -1), return_peak=True)

    def _plan(self, nb: int, T: int, dtype: torch.dtype) -> dict:
        key = (nb, T, dtype)
        plan = self._plans.get(key)
        if plan is not None:
            return plan
        dev = self.dec.device
        plan = {"lat": torch.zeros((nb, T, self.dec.vae_dim), dtype=dtype, device=dev),
                "idx": torch.full((nb,), -1, dtype=torch.int32, device=dev),
                "ws_dec": torch.empty(self.dec.workspace_bytes(nb, T) + 2048, dtype=torch.uint8, device=dev),
                "ws_enc": (torch.empty(self.enc.workspace_bytes(nb, T * self.dec.hop) + 2048, dtype=torch.uint8, device=dev)
                           if self.enc is not None else None),
                "graph": None}
        if self.use_graph:
            # warm-up on a side stream with padding rows only (slot -1: no state is touched), then capture
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(2):
                    self._run(plan)
            torch.cuda.current_stream(dev).wait_stream(side)
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                self._run(plan)
            plan["graph"] = graph
        self._plans[key] = plan
        return plan

    @torch.no_grad()
    def __call__(self, latents: torch.Tensor, sample_indices: torch.Tensor) -> FrameStepOutput:
        """latents [n, T, vae_dim] (T = 1 in generate()) on the decoder's device; sample_indices: int tensor [n] ON THE
        DEVICE (``diffusion_indices``), distinct values in [0, n_slots)."""
        if not (isinstance(sample_indices, torch.Tensor) and sample_indices.is_cuda):
            sample_indices = torch.as_tensor(sample_indices, device=self.dec.device)     # a host list costs one H2D copy
        if latents.dim() != 3 or latents.shape[2] != self.dec.vae_dim:
            raise ValueError("latents must be [n, T, vae_dim]")
        n, T = int(latents.shape[0]), int(latents.shape[1])
        if sample_indices.numel() != n:
            raise AssertionError("len(sample_indices) must match batch size")             # :322-323
        if n == 0:
            raise ValueError("empty step")
        if latents.dtype not in (torch.float32, torch.bfloat16):
            latents = latents.float()
        nb = -(-n // self.bucket) * self.bucket
        plan = self._plan(nb, T, latents.dtype)
        plan["lat"][:n].copy_(latents)
        if n < nb:
            plan["idx"][n:].fill_(-1)
        plan["idx"][:n].copy_(sample_indices.reshape(-1))
        if plan["graph"] is not None:
            plan["graph"].replay()
            self.graph_launches += 1
        else:
            self._run(plan)
        sem = plan["semantic"][:n].clone() if self.enc is not None else None
        pcm = plan["pcm"][:n].clone() if self.with_pcm else None
        peak = plan["peak"][:n].clone() if self.with_pcm else None
        return FrameStepOutput(audio=plan["audio"][:n].clone(), semantic=sem, pcm=pcm, peak=peak)
