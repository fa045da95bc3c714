"""CPU oracle for the flash-attention hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product package
(``flash_attention_ipu_b200`` / ``flash_attention_ipu``) never imports it and
has no CPU fallback.

Parity status: **pinned** against outputs of the reference's own Python code run
in the build container (``oracle/gen_golden.py`` imports the reference's
``_impl/flash_attention_qkv_packed.py`` with a stub ``poptorch`` so that its
non-IPU branch executes, and writes ``tests/golden/*.npz``).  The reference ships
no stored golden vectors of its own.  The non-causal extension
(BASELINE.json config 4) is *unpinned*: the reference refuses that case.

All citations are relative to /root/reference.
"""
from __future__ import annotations

import math
from typing import Tuple

import torch

MASK_VALUE = -10000.0  # flash_attention_ipu/cpp/flash_attention_qkv_packed.cpp:141,238


def packed_attention_reference(qkv: torch.Tensor, causal: bool = True) -> torch.Tensor:
    """Restates flash_attention_ipu/_impl/flash_attention_qkv_packed.py:48-53.

    ``softmax(q @ k^T + triu(-10000, 1)) @ v`` with *no* softmax scale (the caller
    pre-scales q, _impl/scaled_dot_product_attention.py:55).  ``causal=False`` is
    the config-4 extension (no mask at all).
    """
    q, k, v = qkv
    attn = q @ k.permute(0, 2, 1)
    if causal:
        L = q.shape[1]
        mask = torch.triu(torch.full((L, L), MASK_VALUE, dtype=attn.dtype), 1)
        attn = attn + mask
    attn = torch.nn.functional.softmax(attn, dim=-1)
    return attn @ v


def packed_attention_reference_with_lse(
    qkv: torch.Tensor, causal: bool = True
) -> Tuple[torch.Tensor, torch.Tensor]:
    """As above but also returns logSumExp = m + log(l) per row
    (cpp/flash_attention_qkv_packed.cpp:316; AttentionOutputWithStash, .hpp:10-13)."""
    q, k, v = qkv
    attn = q @ k.permute(0, 2, 1)
    if causal:
        L = q.shape[1]
        attn = attn + torch.triu(torch.full((L, L), MASK_VALUE, dtype=attn.dtype), 1)
    lse = torch.logsumexp(attn, dim=-1)
    return torch.exp(attn - lse[..., None]) @ v, lse


def vanilla_attention_grad(
    grad: torch.Tensor, qkv: torch.Tensor, causal: bool = True
) -> torch.Tensor:
    """Analytic backward, restating cpp/vanilla_attention.cpp:68-124:
    dV = P^T dO; dP = dO V^T; dS = P*dP - P*rowsum(P*dP); dQ = dS K; dK = dS^T Q."""
    q, k, v = qkv
    attn = q @ k.permute(0, 2, 1)
    if causal:
        L = q.shape[1]
        attn = attn + torch.triu(torch.full((L, L), MASK_VALUE, dtype=attn.dtype), 1)
    p = torch.softmax(attn, dim=-1)
    dv = p.permute(0, 2, 1) @ grad
    dp = (grad @ v.permute(0, 2, 1)) * p
    s = dp.sum(dim=-1, keepdim=True)
    ds = dp - p * s
    dq = ds @ k
    dk = ds.permute(0, 2, 1) @ q
    return torch.stack([dq, dk, dv])


def chunked_attention_forward(
    qkv: torch.Tensor, num_chunks_q: int, num_chunks_kv: int
) -> Tuple[torch.Tensor, torch.Tensor]:
    """Literal restatement of the chunked forward,
    cpp/flash_attention_qkv_packed.cpp:229-328 (loop structure, block skip :244,
    diagonal-block mask :268-278, online softmax :281-300, epilogue :314-316).
    Pure-Python loops: small cases only."""
    q, k, v = qkv
    G, L, D = q.shape
    cq, ck = L // num_chunks_q, L // num_chunks_kv
    out = torch.zeros_like(q)
    lse = torch.zeros(G, L, dtype=q.dtype)
    for i in range(num_chunks_q):
        q_i = q[:, i * cq:(i + 1) * cq]
        o_i = torch.zeros(G, cq, D, dtype=q.dtype)
        l_i = torch.zeros(G, cq, dtype=q.dtype)
        m_i = torch.full((G, cq), MASK_VALUE, dtype=q.dtype)
        for j in range(num_chunks_kv):
            if (i + 1) * cq > j * ck:  # block has at least one live element
                k_j = k[:, j * ck:(j + 1) * ck]
                v_j = v[:, j * ck:(j + 1) * ck]
                s = q_i @ k_j.permute(0, 2, 1)
                if i * cq < (j + 1) * ck - 1:  # block crosses the diagonal
                    rows = torch.arange(i * cq, (i + 1) * cq)[:, None]
                    cols = torch.arange(j * ck, (j + 1) * ck)[None, :]
                    s = s + torch.where(cols > rows, MASK_VALUE, 0.0).to(s.dtype)
                m_new = torch.maximum(m_i, s.max(dim=-1).values)
                c = torch.exp(m_i - m_new)
                p = torch.exp(s - m_new[..., None])
                l_i = l_i * c + p.sum(dim=-1)
                o_i = o_i * c[..., None] + p @ v_j
                m_i = m_new
        out[:, i * cq:(i + 1) * cq] = o_i / l_i[..., None]
        lse[:, i * cq:(i + 1) * cq] = m_i + torch.log(l_i)
    return out, lse


def chunked_attention_backward(
    grad: torch.Tensor, qkv: torch.Tensor, num_chunks_q: int, num_chunks_kv: int
) -> torch.Tensor:
    """Literal restatement of cpp/flash_attention_qkv_packed.cpp:348-512:
    recompute fwd (:357), D=rowsum(O*dO) (:374-377), KV-outer/Q-inner loop with
    P=exp(S-LSE) (:457-459), dV (:462), dP,dS (:465-467), dQ,dK (:470-472)."""
    q, k, v = qkv
    G, L, D = q.shape
    cq, ck = L // num_chunks_q, L // num_chunks_kv
    out, lse = chunked_attention_forward(qkv, num_chunks_q, num_chunks_kv)
    delta = (out * grad).sum(dim=-1)
    dq, dk, dv = torch.zeros_like(q), torch.zeros_like(k), torch.zeros_like(v)
    for j in range(num_chunks_kv):
        ks = slice(j * ck, (j + 1) * ck)
        for i in range(num_chunks_q):
            if (i + 1) * cq > j * ck:
                qs = slice(i * cq, (i + 1) * cq)
                s = q[:, qs] @ k[:, ks].permute(0, 2, 1)
                if i * cq < (j + 1) * ck - 1:
                    rows = torch.arange(i * cq, (i + 1) * cq)[:, None]
                    cols = torch.arange(j * ck, (j + 1) * ck)[None, :]
                    s = s + torch.where(cols > rows, MASK_VALUE, 0.0).to(s.dtype)
                p = torch.exp(s - lse[:, qs, None])
                dv[:, ks] += p.permute(0, 2, 1) @ grad[:, qs]
                dp = grad[:, qs] @ v[:, ks].permute(0, 2, 1)
                ds = (dp - delta[:, qs, None]) * p
                dq[:, qs] += ds @ k[:, ks]
                dk[:, ks] += ds.permute(0, 2, 1) @ q[:, qs]
    return torch.stack([dq, dk, dv])


def sdpa_reference(q, k, v, is_causal=True, scale=None):
    """The SDPA-path ground truth the reference's own tests use:
    tests/python/test_scaled_dot_product_attention.py:77-84 (torch CPU SDPA)."""
    return torch.nn.functional.scaled_dot_product_attention(
        q, k, v, is_causal=is_causal, scale=scale
    )


def heuristic_chunks(L: int, D: int) -> Tuple[int, int]:
    """_impl/scaled_dot_product_attention.py:52-53."""
    r = max(L, D) // D
    return int(2 ** math.ceil(math.log2(r) / 2)), int(2 ** math.floor(math.log2(r) / 2))


def merge_partials(o_a, lse_a, o_b, lse_b):
    """LSE merge of two partial attention results over disjoint KV sets (ring step;
    SURVEY.md §8e).  Not in the reference (it has no distributed path): plain algebra."""
    lse = torch.logaddexp(lse_a, lse_b)
    wa = torch.exp(lse_a - lse)[..., None]
    wb = torch.exp(lse_b - lse)[..., None]
    return o_a * wa + o_b * wb, lse


def attention_flops(G: int, L: int, D: int, causal: bool, fwd: bool = True, bwd: bool = True) -> float:
    """Algorithmic FLOPs: fwd 4*G*L^2*D, bwd 2.5x, halved if causal."""
    f = 4.0 * G * L * L * D * (0.5 if causal else 1.0)
    return f * ((1.0 if fwd else 0.0) + (2.5 if bwd else 0.0))


_M32 = 0xFFFFFFFF


def _drop_mix(x):
    """lowbias32 integer finaliser on uint64 numpy arrays holding 32-bit values
    (csrc/sm100_ptx.cuh: drop_mix)."""
    x = x ^ (x >> 16)
    x = (x * 0x7FEB352D) & _M32
    x = x ^ (x >> 15)
    x = (x * 0x846CA68B) & _M32
    return x ^ (x >> 16)


def dropout_keep_mask(seed: int, G: int, Lq: int, Lk: int, dropout_p: float, q_pos0: int = 0, kv_pos0: int = 0):
    """The kernels' dropout mask, restated (no reference counterpart: the reference refuses
    dropout_p > 0, _impl/scaled_dot_product_attention.py:25-28).  Returns (keep [G, Lq, Lk] bool
    tensor, scale): key k of query q in group g is kept iff byte (k & 3) of
    drop_hash(seed_g, q, k >> 2) >= round(256 p); survivors are scaled by 256 / (256 - round(256 p)).
    Follows csrc/sm100_ptx.cuh (drop_hash, drop_seed_of_group) and csrc/fa_api.cu (drop_thresh,
    drop_seed32); q, k are global positions (q_pos0 + row, kv_pos0 + key)."""
    import numpy as np
    p32 = float(np.float32(dropout_p))  # the C ABI carries dropout_p as a float
    thresh = min(255, max(1, int(math.floor(p32 * 256.0 + 0.5)))) if p32 > 0 else 0
    seed &= 0xFFFFFFFFFFFFFFFF
    seed32 = (seed & _M32) ^ (((seed >> 32) * 0x9E3779B1) & _M32)
    g = np.arange(G, dtype=np.uint64)
    seed_g = _drop_mix(np.uint64(seed32) ^ ((g * 0xC2B2AE3D + 0x27D4EB2F) & _M32))      # [G]
    q = ((np.arange(Lq, dtype=np.uint64) + np.uint64(q_pos0)) * 0x9E3779B1) & _M32        # [Lq]
    kpos = np.arange(Lk, dtype=np.uint64) + np.uint64(kv_pos0)
    k4 = ((kpos >> 2) * 0x85EBCA77) & _M32                                                # [Lk]
    h = _drop_mix(seed_g[:, None, None] ^ q[None, :, None] ^ k4[None, None, :])
    byte = (h >> ((kpos & 3) * 8)[None, None, :]) & 0xFF
    keep = torch.from_numpy(byte >= thresh)
    return keep, 256.0 / (256.0 - thresh)
