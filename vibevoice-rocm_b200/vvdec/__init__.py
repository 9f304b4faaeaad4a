"""vvdec — B200-native (sm_100a) VibeVoice acoustic-tokenizer decoder behind the reference's decode() API
(and the mirror encoder behind encode()).

Importing this package loads libvvdec_b200.so; there is no CPU / PyTorch fallback.
"""
from . import _capi
from .decoder import B200AcousticDecoder, B200StreamingCache, Geometry, patch
from .encoder import B200AcousticEncoder, B200TokenizerEncoderOutput, patch_encoder
from .framestep import FrameStep, FrameStepOutput
from .pcm import B200PcmStreamer, pcm16
from .sampler import B200DiffusionSampler, SamplerGeometry, patch_sampler
from .shard import plan_shards, Shard
from .sharded import decode_sharded, encode_sharded

__all__ = ["B200DiffusionSampler", "SamplerGeometry", "patch_sampler", "FrameStep", "FrameStepOutput", "decode_sharded", "encode_sharded", "B200PcmStreamer", "pcm16", "B200AcousticEncoder", "B200TokenizerEncoderOutput", "patch_encoder", "B200AcousticDecoder", "B200StreamingCache", "Geometry", "patch", "plan_shards", "Shard", "_capi"]
