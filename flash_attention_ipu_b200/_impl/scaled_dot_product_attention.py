"""Override of ``torch.nn.functional.scaled_dot_product_attention`` -- counterpart of the
reference's ``_impl/scaled_dot_product_attention.py:11-61``.

Kept: the patching mechanism (``patch_function``), the ``NotImplementedError`` gates for
attn_mask / dropout / q,k,v shape mismatch (:21-41, applied before the device check exactly as
in the reference; one boolean key-padding mask form and dropout on the tensor path are taken), the chunk heuristic (:52-53, now only a tile hint) and the pass-through to the
original function for tensors that are not on our device (:60-61).
Changed (documented deviations): ``is_causal=False`` is routed to the
kernel instead of raising (needed by BASELINE.json config 4); ``attn_mask is not None`` replaces
the reference's ``if attn_mask:`` (which breaks on real tensors); torch >= 2.1's ``scale=`` and
``enable_gqa=`` keywords are accepted, and with ``enable_gqa=True`` grouped-query / multi-query
shapes (fewer K/V heads, 16-bit, D in {64,128}) go to the kernels (the reference raises, :33-37); q is not pre-multiplied by D**-0.5 and stacked (:55-57) --
the scale is folded into the kernel and q, k, v are consumed in place through their strides.
"""
from math import ceil, floor, log2
from typing import Callable, Optional

import torch

from . import ops
from .utils import patch_function, unpatch_function

_SUPPORTED = (torch.bfloat16, torch.float16, torch.float32)
KEY_PADDING_FAST_PATH = True  # boolean (B, 1, 1, Lk) masks are taken as prefix (right-padding) masks -> kv_lens


def _chunk_heuristic(L: int, D: int):
    r = max(L, D) // D
    return int(2 ** (ceil(log2(r) / 2))), int(2 ** (floor(log2(r) / 2)))


def _key_padding_lengths(mask: torch.Tensor, query: torch.Tensor, key: torch.Tensor):
    """(B*H,) int32 key counts of a boolean key-padding mask, or None if `mask` is not of that form
    or the problem is off the tensor path."""
    if (mask.dtype != torch.bool or query.ndim != 4 or not query.is_cuda or not mask.is_cuda
            or query.dtype not in (torch.bfloat16, torch.float16) or query.shape[-1] not in (64, 128)):
        return None
    B, H, Lk = query.shape[0], query.shape[1], key.shape[-2]
    if mask.shape == (B, Lk):
        rows = mask
    elif mask.shape == (B, 1, 1, Lk):
        rows = mask.view(B, Lk)
    else:
        return None
    lens = rows.sum(-1, dtype=torch.int32)
    ar = torch.arange(Lk, device=mask.device, dtype=torch.int32)
    torch._assert_async((rows == (ar[None, :] < lens[:, None])).all(),
                        "attn_mask is not a key-padding mask (True entries must be a prefix of every row)")
    return lens.repeat_interleave(H).contiguous()


def _additive_bias(mask: torch.Tensor, query: torch.Tensor, key: torch.Tensor):
    """The mask as an additive score bias the kernels take -- q's dtype, 4-D, broadcastable to
    (B, H, Lq, Lk), unit stride on the keys -- or None if the problem is off the tensor path.  Boolean
    masks become 0 / -inf (True = attend, as in torch); float masks are added as they are."""
    if (query.ndim != 4 or not query.is_cuda or not mask.is_cuda or mask.ndim > 4 or mask.requires_grad
            or query.dtype not in (torch.bfloat16, torch.float16) or query.shape[-1] not in (64, 128)
            or key.shape[-2] % 8 != 0):
        return None
    B, H, Lq, Lk = query.shape[0], query.shape[1], query.shape[2], key.shape[-2]
    m = mask.reshape((1,) * (4 - mask.ndim) + tuple(mask.shape))
    if any(have not in (1, want) for have, want in zip(m.shape, (B, H, Lq, Lk))):
        return None
    if m.dtype == torch.bool:
        m = torch.zeros(m.shape, dtype=query.dtype, device=m.device).masked_fill_(~m, float("-inf"))
    elif m.dtype != query.dtype:
        m = m.to(query.dtype)
    if m.shape[3] != Lk:  # one value for every key: the kernels need unit stride on the keys
        m = m.expand(m.shape[0], m.shape[1], m.shape[2], Lk).contiguous()
    return m.expand(m.shape[0], m.shape[1], Lq, Lk)  # a broadcast query dim stays a stride of 0


def _sdpa_impl(
    orig_fn: Callable,
    query: torch.Tensor,
    key: torch.Tensor,
    value: torch.Tensor,
    attn_mask: Optional[torch.Tensor] = None,
    dropout_p: float = 0.0,
    is_causal: bool = False,
    scale: Optional[float] = None,
    enable_gqa: bool = False,
) -> torch.Tensor:
    kv_lens = bias = None
    if attn_mask is not None:
        # The reference refuses every mask (:21-24).  Here: (1) a boolean key-padding mask, (B, 1, 1, Lk)
        # or (B, Lk) shaped, whose True entries are a prefix of each row, is the per-batch number of
        # existing keys (``fa_b200_problem.kv_lens``: padded KV tiles are skipped; the prefix property is
        # asserted on the device without a host synchronisation; KEY_PADDING_FAST_PATH = False sends such
        # masks down path 2 instead, e.g. for left padding); (2) any other boolean or float mask that
        # broadcasts to (B, H, Lq, Lk) is added to the scores inside the kernels (``fa_b200_problem.bias``).
        # Both on the tensor path only (bf16 / fp16, D in {64, 128}); everything else keeps the error.
        if KEY_PADDING_FAST_PATH:
            kv_lens = _key_padding_lengths(attn_mask, query, key)
        if kv_lens is None:
            bias = _additive_bias(attn_mask, query, key)
        if kv_lens is None and bias is None:
            raise NotImplementedError(
                "flash_attention_ipu supports attn_mask only for 4-D bf16 / fp16 CUDA tensors with D in {64, 128}: "
                "a boolean or float mask broadcastable to (B, H, Lq, Lk) with Lk % 8 == 0 and no gradient")
    dropout_seed = 0
    if dropout_p > 0.0:
        # The reference refuses dropout (:25-28).  Here the tcgen05 kernels apply it (16-bit, D in {64, 128});
        # everything else keeps the reference's error.  The seed comes from torch's CPU generator (no device
        # synchronisation; repeatable under torch.manual_seed) and travels to the backward through autograd.
        if not (query.ndim == 4 and query.is_cuda and query.dtype in (torch.bfloat16, torch.float16)
                and query.shape[-1] in (64, 128) and dropout_p < 1.0):
            raise NotImplementedError(
                "flash_attention_ipu does not currently support non-zero dropout_p "
                "(only for 4-D bf16 / fp16 CUDA tensors with D in {64, 128} and dropout_p < 1)")
        dropout_seed = int(torch.randint(0, 2 ** 62, (1,)).item())
    gqa = (enable_gqa and query.ndim == 4 and key.ndim == 4 and key.shape == value.shape
           and query.shape[0] == key.shape[0] and query.shape[2:] == key.shape[2:]
           and key.shape[1] > 0 and query.shape[1] % key.shape[1] == 0
           and query.is_cuda and query.dtype in (torch.bfloat16, torch.float16) and query.shape[-1] in (64, 128))
    if query.shape != key.shape and not enable_gqa:
        raise NotImplementedError(
            "flash_attention_ipu does not currently support Grouped- or Multi-query "
            "attention (query.shape != key.shape)"
        )
    if key.shape != value.shape:
        raise NotImplementedError("flash_attention_ipu does not currently support value.shape != key.shape")

    on_device = (query.is_cuda and query.dtype in _SUPPORTED and query.ndim >= 2
                 and key.dtype == query.dtype and value.dtype == query.dtype
                 and (query.shape[-1] <= 256) and (query.shape == key.shape or gqa)
                 and (scale is None or scale > 0))  # the kernels want a positive scale; torch accepts any
    if not on_device:
        kwargs = {} if scale is None else {"scale": scale}
        if enable_gqa:
            kwargs["enable_gqa"] = True
        return orig_fn(query, key, value, attn_mask, dropout_p, is_causal, **kwargs)

    L, D = query.shape[-2:]
    batch_shape = query.shape[:-2]
    num_chunks_q, num_chunks_kv = _chunk_heuristic(L, D)
    softmax_scale = float(D ** -0.5 if scale is None else scale)

    if query.ndim == 4:
        q, k, v = query, key, value  # (B, H, L, D), any strides with unit stride on D
    elif query.ndim == 3:
        q, k, v = query.unsqueeze(0), key.unsqueeze(0), value.unsqueeze(0)
    else:
        q, k, v = (t.reshape(1, -1, L, D) for t in (query, key, value))
    out, _lse = ops.attention_fwd(q, k, v, softmax_scale, bool(is_causal),
                                  L // num_chunks_q, L // num_chunks_kv, kv_lens, float(dropout_p), dropout_seed, bias)
    return out.reshape(*batch_shape, L, D) if query.ndim != 4 else out


def uninstall() -> None:
    """Restore torch's own ``scaled_dot_product_attention``."""
    unpatch_function(torch.nn.functional, "scaled_dot_product_attention")


def install() -> Callable:
    """(Re-)install the override on ``torch.nn.functional`` (idempotent)."""
    uninstall()
    return patch_function(torch.nn.functional.scaled_dot_product_attention, [torch.nn.functional])(_sdpa_impl)


# importing this module is the patch, as in the reference (decorator at module level, :11)
_scaled_dot_product_attention_b200 = install()
