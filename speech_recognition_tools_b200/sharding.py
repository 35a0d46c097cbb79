"""Utterance sharding across ranks (one process per GPU).

The reference shards `wav.scp` into `nj` pieces with `utils/split_scp.pl` and runs one Python process per
piece (recipes/timit/local_pyspeech/make_FDLPspectrum_feats.sh:126-157).  Work per utterance is
proportional to its SEGMENT count (every utterance pays one extra leading segment), so utterances are
sorted by segment count and assigned greedily to the least-loaded rank; within a rank they are cut into
batches bounded by a segment budget (scratch memory is proportional to segments)."""
from __future__ import annotations

import heapq
import os
from typing import List, Optional, Sequence, Set


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


def bind_host_to_gpu(device_index: int) -> Optional[Set[int]]:
    """Restrict the calling process to the CPUs NVML reports as local to CUDA device `device_index`, so that the
    pinned staging buffers allocated afterwards (first touch) sit on the GPU's NUMA node.  With one process per GPU
    the host<->device copies of eight ranks otherwise share one socket's memory controllers.  Returns the previous
    affinity (to restore with os.sched_setaffinity) or None when nothing was changed (no NVML, one node, ...)."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(device_index).uuid)
        handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
        cpus = {64 * i + b for i, m in enumerate(mask) for b in range(64) if (int(m) >> b) & 1}
        old = os.sched_getaffinity(0)
        cpus &= old
        if not cpus or cpus == old:
            return None
        os.sched_setaffinity(0, cpus)
        return old
    except Exception:  # placement is an optimisation only
        return None
