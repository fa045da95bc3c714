"""Variable-length (packed-token) attention over the C ABI.

The reference has no such entry point (its inputs are dense ``(3, G, L, D)`` tensors,
``_impl/flash_attention_qkv_packed.py:7-19``; SURVEY.md section 8, row f4 lists it as a gap).  Sequences of
different lengths are stored back to back along the token axis,

    q (total_q, H, D)   k, v (total_k, H_kv, D)   cu_seqlens_q / cu_seqlens_k = [0, l_0, l_0 + l_1, ...]

and attention runs inside each sequence only.  Every sequence is one ``(1, H, L_s, D)`` problem whose operands
are strided VIEWS of the packed tensors (the kernels take any 16-byte aligned row stride through their tensor
maps), so nothing is copied or padded: one forward launch per sequence writes straight into the packed output,
and the backward writes dq / dk / dv straight into the packed gradients.  The launches are queued back to back
on the current stream (programmatic dependent launch overlaps their prologues); a single kernel that walks all
sequences is not built.

CUDA tensors only -- there is no CPU path.
"""
from typing import Sequence

import torch

from ._impl import native

__all__ = ["flash_attention_varlen"]


def _bounds(cu) -> list:
    b = [int(x) for x in (cu.tolist() if isinstance(cu, torch.Tensor) else cu)]
    if len(b) < 2 or b[0] != 0 or any(y < x for x, y in zip(b, b[1:])):
        raise ValueError("cu_seqlens must start at 0 and be non-decreasing")
    return b


def _view(t: torch.Tensor, lo: int, hi: int) -> torch.Tensor:
    # (L_s, H, D) rows of the packed tensor as a (1, H, L_s, D) problem: a view, no copy
    return t[lo:hi].permute(1, 0, 2).unsqueeze(0)


class _VarlenAttention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, bq, bk, causal, scale):
        out = torch.empty_like(q)
        lse = torch.full((q.shape[1], q.shape[0]), float("-inf"), device=q.device, dtype=torch.float32)  # (H, total_q)
        for s in range(len(bq) - 1):
            q0, q1, k0, k1 = bq[s], bq[s + 1], bk[s], bk[s + 1]
            if q1 == q0:
                continue
            if k1 == k0:  # no keys: out = 0, lse = -inf (the kernels' convention for rows that see nothing)
                out[q0:q1].zero_()
                continue
            qs, ks, vs = _view(q, q0, q1), _view(k, k0, k1), _view(v, k0, k1)
            prob = native.make_problem(qs, ks, causal, scale)
            lse_s = torch.empty(1, q.shape[1], q1 - q0, device=q.device, dtype=torch.float32)
            native.fwd(prob, qs, ks, vs, _view(out, q0, q1), lse_s)
            lse[:, q0:q1] = lse_s[0]
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.bq, ctx.bk, ctx.causal, ctx.scale = bq, bk, causal, scale
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse = ctx.saved_tensors
        dout = dout.contiguous()
        dq, dk, dv = torch.zeros_like(q), torch.zeros_like(k), torch.zeros_like(v)
        for s in range(len(ctx.bq) - 1):
            q0, q1, k0, k1 = ctx.bq[s], ctx.bq[s + 1], ctx.bk[s], ctx.bk[s + 1]
            if q1 == q0 or k1 == k0:
                continue
            qs, ks, vs = _view(q, q0, q1), _view(k, k0, k1), _view(v, k0, k1)
            prob = native.make_problem(qs, ks, ctx.causal, ctx.scale)
            ws = torch.empty(max(native.bwd_workspace_bytes(prob), 256), device=q.device, dtype=torch.uint8)
            native.bwd(prob, qs, ks, vs, _view(out, q0, q1), _view(dout, q0, q1), lse[:, q0:q1].contiguous().unsqueeze(0),
                       _view(dq, q0, q1), _view(dk, k0, k1), _view(dv, k0, k1), ws)
        return dq, dk, dv, None, None, None, None


def flash_attention_varlen(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, cu_seqlens_q: Sequence[int],
                           cu_seqlens_k: Sequence[int], causal: bool = True,
                           softmax_scale: "float | None" = None) -> torch.Tensor:
    """Attention inside each of the sequences packed along the token axis.

    q -- (total_q, H, D); k, v -- (total_k, H_kv, D) with H a multiple of H_kv; cu_seqlens_* -- sequence
    boundaries (a Python sequence or an int tensor; read on the host).  With ``causal`` the diagonal is
    aligned at the START of each sequence (query r sees keys <= r of its own sequence).
    returns -- (total_q, H, D)
    """
    if q.ndim != 3 or k.ndim != 3 or v.shape != k.shape or q.shape[2] != k.shape[2]:
        raise ValueError("flash_attention_varlen expects q (total_q, H, D) and k, v (total_k, H_kv, D)")
    bq, bk = _bounds(cu_seqlens_q), _bounds(cu_seqlens_k)
    if len(bq) != len(bk) or bq[-1] != q.shape[0] or bk[-1] != k.shape[0]:
        raise ValueError("cu_seqlens_q / cu_seqlens_k must describe the same number of sequences and cover q / k")
    if not (q.is_cuda and k.is_cuda and v.is_cuda):  # argument errors first, as in the reference's public op
        raise RuntimeError("flash_attention_varlen: q, k, v must be CUDA tensors (B200 / sm_100a); there is no CPU path")
    scale = float(q.shape[2] ** -0.5 if softmax_scale is None else softmax_scale)
    return _VarlenAttention.apply(q.contiguous(), k.contiguous(), v.contiguous(), bq, bk, bool(causal), scale)
