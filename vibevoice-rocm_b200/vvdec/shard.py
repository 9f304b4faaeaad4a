"""Multi-GPU partitioning of decode work: by utterance and by time chunk with a recomputed causal halo.

The decoder is causal FIR + per-time-step ops, so an output sample depends on at most ``halo`` latent
frames of left context (58 for the shipped geometry).  Rank g decodes frames
[start - halo, end) of an utterance and drops the first halo*hop samples: no exchange step, no
collective on the hot path; NCCL is used only to gather the finished waveform slices.
Pure host logic (no torch, no CUDA) so it is unit-tested on CPU, including under gloo with 2 ranks.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence


@dataclass(frozen=True)
class Shard:
    rank: int
    utt: int          # utterance (batch row) index
    start: int        # first frame this shard OWNS
    end: int          # one past the last owned frame
    halo: int         # recomputed frames before ``start`` (0 at the utterance start)

    @property
    def frames(self) -> int:
        return self.end - self.start

    @property
    def work(self) -> int:
        return self.frames + self.halo


def plan_shards(lengths: Sequence[int], world: int, halo: int, min_chunk: int = 2048) -> List[Shard]:
    """Assign every frame of every utterance to exactly one rank.

    Utterances are split into time chunks only when there are fewer work items than ranks would
    need to balance (long-form decode); chunks never get shorter than ``min_chunk`` frames unless the
    utterance itself is shorter, which keeps the halo overhead at most halo/min_chunk.
    Greedy longest-processing-time assignment balances total work (owned + halo frames).
    """
    if world < 1:
        raise ValueError("world must be >= 1")
    total = sum(int(n) for n in lengths)
    target = max(1, -(-total // world))                   # ideal frames per rank
    items = []
    for u, n in enumerate(lengths):
        n = int(n)
        if n <= 0:
            continue
        # split so that no item is much larger than the per-rank target
        parts = max(1, min(-(-n // max(min_chunk, 1)), -(-n // target))) if n > target else 1
        parts = max(1, min(parts, max(1, n // max(1, min(min_chunk, n)))))
        base, extra = divmod(n, parts)
        s = 0
        for p in range(parts):
            e = s + base + (1 if p < extra else 0)
            items.append((u, s, e, min(halo, s)))
            s = e
    items.sort(key=lambda it: -(it[2] - it[1] + it[3]))
    load = [0] * world
    out: List[Shard] = []
    for u, s, e, h in items:
        r = min(range(world), key=lambda i: (load[i], i))
        load[r] += (e - s) + h
        out.append(Shard(rank=r, utt=u, start=s, end=e, halo=h))
    out.sort(key=lambda sh: (sh.utt, sh.start))
    return out
