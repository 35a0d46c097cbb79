"""Utterance sharding across ranks (one process per GPU).

The reference shards `wav.scp` into `nj` pieces with `utils/split_scp.pl` and runs one Python process per
piece (recipes/timit/local_pyspeech/make_FDLPspectrum_feats.sh:126-157).  Work per utterance is
proportional to its SEGMENT count (every utterance pays one extra leading segment), so utterances are
sorted by segment count and assigned greedily to the least-loaded rank; within a rank they are cut into
batches bounded by a segment budget (scratch memory is proportional to segments)."""
from __future__ import annotations

import heapq
from typing import List, Sequence


def assign_to_ranks(n_segments: Sequence[int], world_size: int) -> List[List[int]]:
    """Greedy longest-processing-time assignment; returns the utterance indices of every rank."""
    order = sorted(range(len(n_segments)), key=lambda u: (-int(n_segments[u]), u))
    heap = [(0, r) for r in range(world_size)]
    heapq.heapify(heap)
    shards: List[List[int]] = [[] for _ in range(world_size)]
    for u in order:
        load, r = heapq.heappop(heap)
        shards[r].append(u)
        heapq.heappush(heap, (load + int(n_segments[u]), r))
    for s in shards:
        s.sort()
    return shards


def make_batches(utts: Sequence[int], n_segments: Sequence[int], max_segments: int) -> List[List[int]]:
    """Length-bucketed batches: utterances sorted by segment count, cut when the segment budget is reached."""
    order = sorted(utts, key=lambda u: (int(n_segments[u]), u))
    batches, cur, load = [], [], 0
    for u in order:
        n = int(n_segments[u])
        if cur and load + n > max_segments:
            batches.append(cur)
            cur, load = [], 0
        cur.append(u)
        load += n
    if cur:
        batches.append(cur)
    return batches
