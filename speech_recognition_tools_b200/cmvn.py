"""Global CMVN: the one cross-shard reduction of the path.

Reference: Kaldi `compute-cmvn-stats scp:data/$train/feats.scp data/$train/cmvn.ark`
(e2e/wsj/run_fdlp_e1.sh:280), a serial single-process pass over every shard's output, followed by ESPnet
`dump.sh` applying it (e2e/wsj/run_fdlp_e1.sh:293-303).  Here the overlap-add kernel accumulates the
per-GPU partial sums [sum x, count; sum x^2, 0] (float64, 2 x (D+1)) while it writes the features, the
ranks all-reduce them over NCCL, and `apply` normalises in place with a CUDA kernel (cmvn_apply)."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L


def new_accumulator(dim: int, device) -> torch.Tensor:
    return torch.zeros((2, dim + 1), dtype=torch.float64, device=device)


def all_reduce(stats: torch.Tensor, group=None) -> torch.Tensor:
    """Sum the partial stats over all ranks (NCCL on GPUs, gloo on CPU); no-op without a process group."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats


def apply(feats: torch.Tensor, stats: torch.Tensor, norm_vars: bool = True) -> torch.Tensor:
    """In place (x - mean) / sqrt(var) with the global stats, on the device (Kaldi apply-cmvn --norm-vars)."""
    if not feats.is_cuda or feats.dtype != torch.float32 or not feats.is_contiguous():
        raise ValueError("feats must be a contiguous float32 CUDA tensor")
    lib = L.load()
    stats = stats.to(device=feats.device, dtype=torch.float64).contiguous()
    stream = torch.cuda.current_stream(feats.device).cuda_stream
    with torch.cuda.device(feats.device):
        L.check(lib.cmvn_apply(C.c_void_p(feats.data_ptr()), feats.shape[0], feats.shape[1],
                               C.c_void_p(stats.data_ptr()), 1 if norm_vars else 0, C.c_void_p(stream)))
    return feats
