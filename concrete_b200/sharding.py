"""Ciphertext sharding across GPUs / ranks.

Ciphertexts are independent and keys are read-only, so a batch of S ciphertexts is split
contiguously: S_g = floor(S/G) (+1 for g < S mod G) — the rule of the reference's
`get_num_inputs_on_gpu` (third_party/tfhe-rs/backends/tfhe-cuda-backend/cuda/src/utils/
helper_multi_gpu.cu:53-83) and of the per-device host threads in wrappers.cpp::for_each_shard.
There is no data-path collective: keys are broadcast once, outputs are gathered by plain copies.
"""
from __future__ import annotations

from typing import List, Tuple


def shard_counts(total: int, parts: int) -> List[int]:
    if parts < 1:
        raise ValueError("parts must be >= 1")
    if total < 0:
        raise ValueError("total must be >= 0")
    base, rem = divmod(total, parts)
    return [base + (1 if g < rem else 0) for g in range(parts)]


def shard_range(total: int, parts: int, index: int) -> Tuple[int, int]:
    """(begin, count) of shard `index` of `parts` over `total` items."""
    if not 0 <= index < parts:
        raise ValueError("index out of range")
    counts = shard_counts(total, parts)
    return sum(counts[:index]), counts[index]


def broadcast_keys(tensors, src: int = 0):
    """Broadcast key tensors (Fourier BSK, KSK) from `src` once; NCCL over NVLink on GPUs,
    gloo in the CPU tests.  No-op when torch.distributed is not initialised."""
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return tensors
    for t in tensors:
        dist.broadcast(t, src=src)
    return tensors
