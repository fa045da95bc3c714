"""Sequence-sharded causal attention over a ring of GPUs (BASELINE.json config 5).

The reference is single-device: its chunked loop (``cpp/flash_attention_qkv_packed.cpp:229-328``)
walks KV blocks on one IPU and keeps a running (max, sum, O).  Here the same recurrence runs
ACROSS devices: every rank keeps its own Q rows, the K/V shards travel around a ring
(``torch.distributed`` P2P over NCCL / NVLink) and each hop's partial result is folded into fp32
accumulators with the log-sum-exp merge (``fa_b200_merge_acc``), which is the cross-device form of
the running-softmax update at ``:281-300``.

Layout: the sequence is cut into ``2*P`` chunks of ``C`` rows and rank ``r`` holds chunks
``r`` ("lo") and ``2P-1-r`` ("hi") -- the zig-zag split that gives every rank the same causal work.
With that split every hop is exactly ONE kernel launch of equal cost (``2*C*C`` score entries):

    visiting shard from rank src == r : causal,  Q = [lo;hi]  K/V = [lo;hi]   (local block)
    src <  r                          : full,    Q = [lo;hi]  K/V = lo of src
    src >  r                          : full,    Q = hi       K/V = [lo;hi] of src

Backward: a second ring in which the fp32 dK/dV accumulators travel with their K/V shard and
come home after ``P`` hops; dQ is accumulated locally.  The per-block gradient kernel is the
ordinary one -- it only needs the GLOBAL lse and O (delta = rowsum(O o dO)), which is why the
forward keeps both.

Two drivers share the per-hop code: :func:`ring_attention` (one process per GPU, real P2P) and
:func:`ring_attention_emulated` (all ranks stepped in lock-step inside one process; used to
check the schedule and the merge on a single GPU).
"""
from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

__all__ = ["zigzag_shard", "zigzag_unshard", "hop_plan", "ring_attention", "ring_attention_emulated",
           "NativeBlocks"]


# ------------------------------------------------------------------ layout helpers
def zigzag_shard(x: torch.Tensor, rank: int, world: int, dim: int = 2) -> torch.Tensor:
    """Rows of ``x`` (sequence on ``dim``) that rank ``rank`` owns: chunks ``rank`` and ``2*world-1-rank``."""
    if x.shape[dim] % (2 * world):
        raise ValueError(f"sequence length {x.shape[dim]} is not divisible by 2*world = {2 * world}")
    chunks = x.chunk(2 * world, dim=dim)
    return torch.cat([chunks[rank], chunks[2 * world - 1 - rank]], dim=dim).contiguous()


def zigzag_unshard(shards: Sequence[torch.Tensor], dim: int = 2) -> torch.Tensor:
    """Inverse of :func:`zigzag_shard` given every rank's shard in rank order."""
    world = len(shards)
    chunks: List[Optional[torch.Tensor]] = [None] * (2 * world)
    for r, s in enumerate(shards):
        lo, hi = s.chunk(2, dim=dim)
        chunks[r], chunks[2 * world - 1 - r] = lo, hi
    return torch.cat(chunks, dim=dim)


def hop_plan(rank: int, src: int) -> Tuple[str, str, bool]:
    """(query rows, key rows, causal) of the single block a rank computes against ``src``'s shard."""
    if src == rank:
        return "all", "all", True
    if src < rank:
        return "all", "lo", False
    return "hi", "all", False


def _rows(x: torch.Tensor, which: str) -> torch.Tensor:
    """Sub-range of the sequence dim (-2) of a local shard: 'all' | 'lo' | 'hi' (zero-copy view)."""
    if which == "all":
        return x
    C = x.shape[-2] // 2
    return x[..., :C, :] if which == "lo" else x[..., C:, :]


# ------------------------------------------------------------------ per-block kernels
class NativeBlocks:
    """The CUDA kernels behind one hop (the only implementation the product uses; tests of the
    host-side schedule on CPU substitute an object with the same three methods)."""

    def fwd_block(self, q, k, v, causal: bool, scale: float):
        from ._impl import ops
        return ops.attention_fwd(q, k, v, scale, causal, 0, 0)  # (o, lse)

    def bwd_block(self, dout, q, k, v, out, lse, causal: bool, scale: float):
        from ._impl import ops
        return ops.attention_bwd(dout, q, k, v, out, lse, scale, causal, 0, 0)  # (dq, dk, dv)

    def bwd_block_accum(self, dout, q, k, v, out, lse, causal: bool, scale: float, dq_acc):
        """Backward of one block with dQ reduce-added in place into the fp32 accumulator rows
        ``dq_acc`` (WITHOUT softmax_scale: the ring applies it once at the end), through the C ABI
        (``fa_b200_bwd_dq_accum``).  Returns (dk, dv) in the kernel dtype, or None when the
        problem is off the tensor path (the caller then uses :meth:`bwd_block` and adds)."""
        from ._impl import native
        if q.dtype not in (torch.bfloat16, torch.float16) or q.shape[-1] not in (64, 128):
            return None
        prob = native.make_problem(q, k, causal, scale)
        if native.kernel_family(prob) != 1:
            return None
        lse = lse.contiguous()
        dk, dv = torch.empty_like(k, memory_format=torch.contiguous_format), torch.empty_like(v, memory_format=torch.contiguous_format)
        ws = torch.empty(native.bwd_workspace_bytes(prob), dtype=torch.uint8, device=q.device)
        native.bwd_dq_accum(prob, q, k, v, out, dout, lse, dq_acc, dk, dv, ws)
        return dk, dv

    def merge_acc(self, o_acc, lse_acc, o_b, lse_b, row0: int) -> None:
        from ._impl import native
        G = o_acc.shape[0] * o_acc.shape[1]
        native.merge_acc(o_acc.view(G, *o_acc.shape[2:]), lse_acc.view(G, -1),
                         o_b.reshape(G, *o_b.shape[2:]), lse_b.reshape(G, -1), row0)


def _check_inputs(q, k, v) -> None:
    if q.ndim != 4 or q.shape != k.shape or k.shape != v.shape:
        raise ValueError(f"ring attention wants equal (B, H, 2C, D) shards, got {q.shape} {k.shape} {v.shape}")
    if q.shape[2] % 2:
        raise ValueError("local shard must hold two chunks (even number of rows)")


def _fwd_hop(blocks, rank: int, src: int, q, kv, o_acc, lse_acc, scale: float) -> None:
    """Attention of the local Q rows against the visiting shard ``kv`` (stacked K, V of rank
    ``src``), folded into the fp32 accumulators."""
    qr, kr, causal = hop_plan(rank, src)
    o_b, lse_b = blocks.fwd_block(_rows(q, qr), _rows(kv[0], kr), _rows(kv[1], kr), causal, scale)
    row0 = q.shape[2] // 2 if qr == "hi" else 0
    blocks.merge_acc(o_acc, lse_acc, o_b, lse_b, row0)


def _bwd_hop(blocks, rank: int, src: int, q, kv, out, dout, lse, lse_hi, dq_acc, scale: float):
    """Gradients of one block.  dQ is added to the local fp32 accumulator; the (dK, dV) partials
    are returned with the key-row range they belong to (they are added to the travelling
    accumulator once it has arrived)."""
    qr, kr, causal = hop_plan(rank, src)
    lse_q = lse_hi if qr == "hi" else lse
    args = (_rows(dout, qr), _rows(q, qr), _rows(kv[0], kr), _rows(kv[1], kr), _rows(out, qr), lse_q, causal, scale)
    fused = getattr(blocks, "bwd_block_accum", None)
    res = fused(*args, _rows(dq_acc, qr)) if fused is not None else None
    if res is not None:   # dQ went straight into the accumulator (unscaled: see _finish_dq)
        dk, dv = res
    else:
        dq, dk, dv = blocks.bwd_block(*args)
        _rows(dq_acc, qr).add_(dq, alpha=1.0 / scale)  # keep the accumulator in the unscaled convention
    return kr, dk, dv


def _finish_dq(dq_acc, scale: float, dtype):
    """The fp32 dQ accumulator holds sum(dS K) over all hops; softmax_scale is applied once here."""
    return dq_acc.mul_(scale).to(dtype)


# ------------------------------------------------------------------ ring exchange (real P2P)
class _Exchange:
    """One in-flight "send to the next rank, receive from the previous rank" of a list of
    tensors.  NCCL runs it on its own stream, so it overlaps the block kernel enqueued after
    ``start``; ``wait`` orders the current stream after the transfer."""

    def __init__(self, group, send: Sequence[torch.Tensor], recv: Sequence[torch.Tensor]):
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        nxt = dist.get_global_rank(group, (rank + 1) % world) if group is not None else (rank + 1) % world
        prv = dist.get_global_rank(group, (rank - 1) % world) if group is not None else (rank - 1) % world
        ops = []
        for s, r in zip(send, recv):
            ops.append(dist.P2POp(dist.isend, s, nxt, group))
            ops.append(dist.P2POp(dist.irecv, r, prv, group))
        self._reqs = dist.batch_isend_irecv(ops)

    def wait(self) -> None:
        for r in self._reqs:
            r.wait()


def _ring_forward(blocks, group, q, k, v, scale: float):
    """Forward of one rank.  Returns (out in q.dtype, lse fp32 (B,H,2C))."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    B, H, L2, D = q.shape
    o_acc = torch.zeros((B, H, L2, D), dtype=torch.float32, device=q.device)
    lse_acc = torch.full((B, H, L2), float("-inf"), dtype=torch.float32, device=q.device)
    bufs = [torch.stack([k, v]), None]  # (2, B, H, 2C, D): K and V travel as one message
    if world > 1:
        bufs[1] = torch.empty_like(bufs[0])
    for s in range(world):
        cur = bufs[s % 2]
        xch = _Exchange(group, [cur], [bufs[(s + 1) % 2]]) if s + 1 < world else None
        _fwd_hop(blocks, rank, (rank - s) % world, q, cur, o_acc, lse_acc, scale)
        if xch is not None:
            xch.wait()
    return o_acc.to(q.dtype), lse_acc


def _ring_backward(blocks, group, q, k, v, out, lse, dout, scale: float):
    """Backward of one rank.  K/V shards make the same trip as in the forward; each carries its
    fp32 (dK, dV) accumulator, which is sent on after the local block has been added and reaches
    its owner with the P-th exchange."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    C = q.shape[2] // 2
    lse_hi = lse[:, :, C:].contiguous()
    dq_acc = torch.zeros(q.shape, dtype=torch.float32, device=q.device)
    kv = [torch.stack([k, v]), None]
    dkv = [torch.zeros((2,) + tuple(k.shape), dtype=torch.float32, device=q.device), None]
    if world > 1:
        kv[1] = torch.empty_like(kv[0])
        dkv[1] = torch.empty_like(dkv[0])
    dkv_xch = None
    for s in range(world):
        cur = kv[s % 2]
        kv_xch = _Exchange(group, [cur], [kv[(s + 1) % 2]]) if s + 1 < world else None
        kr, dk, dv = _bwd_hop(blocks, rank, (rank - s) % world, q, cur, out, dout, lse, lse_hi, dq_acc, scale)
        if dkv_xch is not None:
            dkv_xch.wait()  # accumulator of the shard we are holding, with all earlier hops' terms
        acc = dkv[s % 2]
        _rows(acc[0], kr).add_(dk)
        _rows(acc[1], kr).add_(dv)
        if world > 1:
            dkv_xch = _Exchange(group, [acc], [dkv[(s + 1) % 2]])
        if kv_xch is not None:
            kv_xch.wait()
    if dkv_xch is not None:
        dkv_xch.wait()
    home = dkv[world % 2] if world > 1 else dkv[0]
    return _finish_dq(dq_acc, scale, q.dtype), home[0].to(k.dtype), home[1].to(v.dtype)


class _RingAttention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, group, scale, blocks):
        out, lse = _ring_forward(blocks, group, q, k, v, scale)
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.group, ctx.scale, ctx.blocks = group, scale, blocks
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse = ctx.saved_tensors
        dq, dk, dv = _ring_backward(ctx.blocks, ctx.group, q, k, v, out, lse, dout.contiguous(), ctx.scale)
        return dq, dk, dv, None, None, None


def ring_attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, group=None,
                   softmax_scale: Optional[float] = None, blocks=None) -> torch.Tensor:
    """Causal attention over a sequence sharded zig-zag across the ranks of ``group``.

    ``q, k, v``: this rank's (B, H, 2C, D) shard (see :func:`zigzag_shard`).  Returns this rank's
    rows of the output; differentiable.  One process per GPU; K/V move by P2P send/recv only."""
    _check_inputs(q, k, v)
    if blocks is None:
        if not q.is_cuda:
            raise RuntimeError("ring_attention runs on CUDA (sm_100a) tensors only; there is no CPU path")
        blocks = NativeBlocks()
    scale = q.shape[-1] ** -0.5 if softmax_scale is None else float(softmax_scale)
    return _RingAttention.apply(q.contiguous(), k.contiguous(), v.contiguous(), group, scale, blocks)


# ------------------------------------------------------------------ single-process emulation
def ring_attention_emulated(q_shards: Sequence[torch.Tensor], k_shards: Sequence[torch.Tensor],
                            v_shards: Sequence[torch.Tensor], dout_shards: Optional[Sequence[torch.Tensor]] = None,
                            softmax_scale: Optional[float] = None, blocks=None):
    """The same hop schedule with every rank stepped in lock-step inside one process (the shard a
    rank would have received is simply indexed).  Returns per-rank outputs, and per-rank
    (dq, dk, dv) as well when ``dout_shards`` is given.  Used to check schedule + merge on one GPU."""
    world = len(q_shards)
    for q, k, v in zip(q_shards, k_shards, v_shards):
        _check_inputs(q, k, v)
    if blocks is None:
        if not q_shards[0].is_cuda:
            raise RuntimeError("ring attention runs on CUDA (sm_100a) tensors only; there is no CPU path")
        blocks = NativeBlocks()
    scale = q_shards[0].shape[-1] ** -0.5 if softmax_scale is None else float(softmax_scale)
    kv = [torch.stack([k, v]) for k, v in zip(k_shards, v_shards)]
    outs, lses = [], []
    for r, q in enumerate(q_shards):
        B, H, L2, D = q.shape
        o_acc = torch.zeros((B, H, L2, D), dtype=torch.float32, device=q.device)
        lse_acc = torch.full((B, H, L2), float("-inf"), dtype=torch.float32, device=q.device)
        for s in range(world):
            src = (r - s) % world
            _fwd_hop(blocks, r, src, q, kv[src], o_acc, lse_acc, scale)
        outs.append(o_acc.to(q.dtype))
        lses.append(lse_acc)
    if dout_shards is None:
        return outs
    dkv = [torch.zeros(t.shape, dtype=torch.float32, device=t.device) for t in kv]
    dqs = []
    for r, q in enumerate(q_shards):
        C = q.shape[2] // 2
        dq_acc = torch.zeros(q.shape, dtype=torch.float32, device=q.device)
        lse_hi = lses[r][:, :, C:].contiguous()
        for s in range(world):
            src = (r - s) % world
            kr, dk, dv = _bwd_hop(blocks, r, src, q, kv[src], outs[r], dout_shards[r].contiguous(), lses[r],
                                  lse_hi, dq_acc, scale)
            _rows(dkv[src][0], kr).add_(dk)
            _rows(dkv[src][1], kr).add_(dv)
        dqs.append(_finish_dq(dq_acc, scale, q.dtype))
    grads = [(dqs[r], dkv[r][0].to(k_shards[r].dtype), dkv[r][1].to(v_shards[r].dtype)) for r in range(world)]
    return outs, grads
