"""Sharding of the (batch, head) groups over GPUs -- the multi-GPU form of configs 2-4.

The packed ``(3, G, L, D)`` tensor of the reference API holds ``G`` independent attention problems
(the reference flattens batch and heads into G before the op,
``flash_attention_ipu/_impl/scaled_dot_product_attention.py:57``), so N GPUs split the groups and
never exchange anything: rank ``r`` runs the ordinary public op on groups ``shard_groups(G, N, r)``.
The split is contiguous and balanced; every group has the same causal cost, so equal counts are
equal work.  Inside a GPU the kernels then order tiles heaviest-first (the persistent backward
hands KV tiles out by an atomic counter), which is what keeps 16 groups x 64 causal tiles from
leaving a tail on 148 SMs.
"""
from typing import Tuple

import torch

__all__ = ["shard_groups", "local_groups", "sharded_flash_attention_qkv_packed"]


def shard_groups(G: int, world: int, rank: int) -> Tuple[int, int]:
    """[lo, hi) of the groups rank ``rank`` of ``world`` owns: contiguous, sizes differ by at most 1."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} of world {world}")
    base, rem = divmod(G, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def local_groups(qkv: torch.Tensor, world: int, rank: int) -> torch.Tensor:
    """Zero-copy view ``qkv[:, lo:hi]`` of a packed (3, G, L, D) tensor for this rank."""
    if qkv.ndim != 4 or qkv.shape[0] != 3:
        raise ValueError(f"expected a packed (3, G, L, D) tensor, got {tuple(qkv.shape)}")
    lo, hi = shard_groups(qkv.shape[1], world, rank)
    return qkv[:, lo:hi]


def sharded_flash_attention_qkv_packed(qkv: torch.Tensor, world: int, rank: int, num_chunks_q: int = 1,
                                       num_chunks_kv: int = 1, **kw) -> torch.Tensor:
    """This rank's rows ``[lo:hi]`` of ``flash_attention_qkv_packed(qkv, ...)`` for a replicated or
    host-resident global ``qkv``; no collective.  (A rank that already holds only its groups just
    calls the public op.)"""
    from ._impl.flash_attention_qkv_packed import flash_attention_qkv_packed
    return flash_attention_qkv_packed(local_groups(qkv, world, rank), num_chunks_q, num_chunks_kv, **kw)
