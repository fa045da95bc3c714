"""``torch.library`` ops over the C ABI -- the B200 stand-in for the reference's popart custom op
``("ai.graphcore", "FlashAttentionQKVPacked", 1)`` and its grad op
(``flash_attention_ipu/cpp/flash_attention_qkv_packed.cpp:516-517, 522-607``).

Autograd contract: the forward op returns ``(out, lse)`` and the backward op consumes the saved
``out``/``lse``.  The reference's grad op instead receives only ``(dOut, qkv)`` and re-runs the
whole forward (``cpp/flash_attention_qkv_packed.cpp:357-359, 536-541``); that recompute is
deliberately not replicated.
"""
from typing import Tuple

import torch

from . import native


def _as_bhld(t: torch.Tensor) -> torch.Tensor:
    """(B,H,L,D) view the TMA path can address: unit stride on D, 16-byte aligned strides."""
    es = t.element_size()
    ok = t.stride(-1) == 1 and t.data_ptr() % 16 == 0 and all((s * es) % 16 == 0 for s in t.stride()[:-1])
    return t if ok else t.contiguous()


# ------------------------------------------------------------------ separate q, k, v  (SDPA path)
@torch.library.custom_op("flash_attention_ipu::attention_fwd", mutates_args=())
def attention_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, softmax_scale: float,
                  causal: bool, block_q: int, block_kv: int) -> Tuple[torch.Tensor, torch.Tensor]:
    q, k, v = _as_bhld(q), _as_bhld(k), _as_bhld(v)
    B, H, Lq, D = q.shape
    out = torch.empty_like(q)  # keeps q's (possibly (B,L,H,D)-transposed) dense layout
    if out.stride(-1) != 1:
        out = torch.empty(q.shape, dtype=q.dtype, device=q.device)
    lse = torch.empty((B, H, Lq), dtype=torch.float32, device=q.device)
    prob = native.make_problem(q, k, causal, softmax_scale, block_q, block_kv)
    native.fwd(prob, q, k, v, out, lse)
    return out, lse


@attention_fwd.register_fake
def _(q, k, v, softmax_scale, causal, block_q, block_kv):
    return torch.empty_like(q), q.new_empty(q.shape[:3], dtype=torch.float32)


@torch.library.custom_op("flash_attention_ipu::attention_bwd", mutates_args=())
def attention_bwd(dout: torch.Tensor, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor,
                  out: torch.Tensor, lse: torch.Tensor, softmax_scale: float, causal: bool,
                  block_q: int, block_kv: int) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    q, k, v, out, dout = _as_bhld(q), _as_bhld(k), _as_bhld(v), _as_bhld(out), _as_bhld(dout)
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    prob = native.make_problem(q, k, causal, softmax_scale, block_q, block_kv)
    ws = torch.empty(native.bwd_workspace_bytes(prob), dtype=torch.uint8, device=q.device)
    native.bwd(prob, q, k, v, out, dout, lse, dq, dk, dv, ws)
    return dq, dk, dv


@attention_bwd.register_fake
def _(dout, q, k, v, out, lse, softmax_scale, causal, block_q, block_kv):
    return torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)


def _attention_setup(ctx, inputs, output):
    q, k, v, ctx.softmax_scale, ctx.causal, ctx.block_q, ctx.block_kv = inputs
    out, lse = output
    ctx.save_for_backward(q, k, v, out, lse)


def _attention_backward(ctx, dout, dlse):
    q, k, v, out, lse = ctx.saved_tensors
    dq, dk, dv = attention_bwd(dout, q, k, v, out, lse, ctx.softmax_scale, ctx.causal,
                               ctx.block_q, ctx.block_kv)
    return dq, dk, dv, None, None, None, None


attention_fwd.register_autograd(_attention_backward, setup_context=_attention_setup)


# ------------------------------------------------------------------ packed qkv  (reference API)
@torch.library.custom_op("flash_attention_ipu::qkv_packed_fwd", mutates_args=())
def qkv_packed_fwd(qkv: torch.Tensor, softmax_scale: float, causal: bool, block_q: int,
                   block_kv: int) -> Tuple[torch.Tensor, torch.Tensor]:
    qkv = qkv.contiguous()
    _, G, L, D = qkv.shape
    q, k, v = (qkv[i].unsqueeze(0) for i in range(3))  # zero-copy (1,G,L,D) views
    out = torch.empty((G, L, D), dtype=qkv.dtype, device=qkv.device)
    lse = torch.empty((G, L), dtype=torch.float32, device=qkv.device)
    prob = native.make_problem(q, k, causal, softmax_scale, block_q, block_kv)
    native.fwd(prob, q, k, v, out.unsqueeze(0), lse)
    return out, lse


@qkv_packed_fwd.register_fake
def _(qkv, softmax_scale, causal, block_q, block_kv):
    return qkv.new_empty(qkv.shape[1:]), qkv.new_empty(qkv.shape[1:3], dtype=torch.float32)


@torch.library.custom_op("flash_attention_ipu::qkv_packed_bwd", mutates_args=())
def qkv_packed_bwd(dout: torch.Tensor, qkv: torch.Tensor, out: torch.Tensor, lse: torch.Tensor,
                   softmax_scale: float, causal: bool, block_q: int, block_kv: int) -> torch.Tensor:
    qkv, dout, out = qkv.contiguous(), dout.contiguous(), out.contiguous()
    q, k, v = (qkv[i].unsqueeze(0) for i in range(3))
    dqkv = torch.empty_like(qkv)  # gradients land directly in the packed (3,G,L,D) buffer
    dq, dk, dv = (dqkv[i].unsqueeze(0) for i in range(3))
    prob = native.make_problem(q, k, causal, softmax_scale, block_q, block_kv)
    ws = torch.empty(native.bwd_workspace_bytes(prob), dtype=torch.uint8, device=qkv.device)
    native.bwd(prob, q, k, v, out.unsqueeze(0), dout.unsqueeze(0), lse, dq, dk, dv, ws)
    return dqkv


@qkv_packed_bwd.register_fake
def _(dout, qkv, out, lse, softmax_scale, causal, block_q, block_kv):
    return torch.empty_like(qkv)


def _packed_setup(ctx, inputs, output):
    qkv, ctx.softmax_scale, ctx.causal, ctx.block_q, ctx.block_kv = inputs
    out, lse = output
    ctx.save_for_backward(qkv, out, lse)


def _packed_backward(ctx, dout, dlse):
    qkv, out, lse = ctx.saved_tensors
    dqkv = qkv_packed_bwd(dout, qkv, out, lse, ctx.softmax_scale, ctx.causal, ctx.block_q, ctx.block_kv)
    return dqkv, None, None, None, None


qkv_packed_fwd.register_autograd(_packed_backward, setup_context=_packed_setup)
