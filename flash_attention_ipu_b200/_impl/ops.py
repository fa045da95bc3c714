"""The package's ``torch.library`` ops -- the B200 stand-in for the reference's popart custom op
``("ai.graphcore", "FlashAttentionQKVPacked", 1)`` and its grad op
(``flash_attention_ipu/cpp/flash_attention_qkv_packed.cpp:516-517, 522-607``).

Like the reference's, the ops are registered natively: ``csrc/torch_ops.cpp`` (built into
``build/libfa_torch_ops.so``) defines the schemas, the implementation over the C ABI of
``include/fa_b200.h`` and the autograd functions; loading that library is what registers them
(the reference's plugin registers its op from static initialisers the same way,
``cpp/flash_attention_qkv_packed.cpp:618-625``).  This module loads it, adds the shape-only
("fake") implementations that tracing needs, and exposes the four ops under their names:

    flash_attention_ipu::attention_fwd(q, k, v, softmax_scale, causal, block_q, block_kv,
                                       kv_lens=None, dropout_p=0.0, dropout_seed=0)       -> (out, lse)
    flash_attention_ipu::attention_bwd(dout, q, k, v, out, lse, ...)                      -> (dq, dk, dv)
    flash_attention_ipu::qkv_packed_fwd(qkv, softmax_scale, causal, block_q, block_kv)    -> (out, lse)
    flash_attention_ipu::qkv_packed_bwd(dout, qkv, out, lse, ...)                         -> dqkv

Autograd contract: the forward ops return ``(out, lse)`` and the backward ops consume the saved
``out``/``lse``.  The reference's grad op instead receives only ``(dOut, qkv)`` and re-runs the
whole forward (``cpp/flash_attention_qkv_packed.cpp:357-359, 536-541``); that recompute is
deliberately not replicated.

There is no Python implementation behind these names: without the built library the import
fails (``ImportError``), and the ops refuse non-CUDA tensors.
"""
import pathlib

import torch

from .. import _lib_path  # the C-ABI library the ops library links against (already loaded)

_OPS_LIB = "libfa_torch_ops.so"


def _load_ops_library() -> str:
    path = pathlib.Path(_lib_path).parent / _OPS_LIB
    if not path.exists():
        raise ImportError(f"Cannot find {_OPS_LIB} next to {_lib_path} (run `make`)")
    torch.ops.load_library(str(path))
    return str(path)


ops_library_path = _load_ops_library()

attention_fwd = torch.ops.flash_attention_ipu.attention_fwd
attention_bwd = torch.ops.flash_attention_ipu.attention_bwd
qkv_packed_fwd = torch.ops.flash_attention_ipu.qkv_packed_fwd
qkv_packed_bwd = torch.ops.flash_attention_ipu.qkv_packed_bwd


def _like_kernel_view(t):
    """Shape/stride twin of what csrc/torch_ops.cpp allocates next to an operand: the operand is
    used in place when D has unit stride and every other dim of size > 1 has a positive, 16-byte
    aligned stride (``as_bhld``), and the result is ``empty_like`` of that view; otherwise the operand
    is made contiguous first and the result is contiguous."""
    es = t.element_size()
    ok = t.ndim == 4 and t.stride(-1) == 1 and all(
        t.size(i) == 1 or (t.stride(i) > 0 and (t.stride(i) * es) % 16 == 0) for i in range(t.ndim - 1))
    if not ok:
        return t.new_empty(t.shape)
    like = torch.empty_like(t)
    return like if like.stride(-1) == 1 else t.new_empty(t.shape)


@torch.library.register_fake("flash_attention_ipu::attention_fwd")
def _(q, k, v, softmax_scale, causal, block_q, block_kv, kv_lens=None, dropout_p=0.0, dropout_seed=0, bias=None):
    return _like_kernel_view(q), q.new_empty(q.shape[:3], dtype=torch.float32)


@torch.library.register_fake("flash_attention_ipu::attention_bwd")
def _(dout, q, k, v, out, lse, softmax_scale, causal, block_q, block_kv, kv_lens=None, dropout_p=0.0, dropout_seed=0,
      bias=None):
    return _like_kernel_view(q), _like_kernel_view(k), _like_kernel_view(v)


@torch.library.register_fake("flash_attention_ipu::qkv_packed_fwd")
def _(qkv, softmax_scale, causal, block_q, block_kv, kv_lens=None, dropout_p=0.0, dropout_seed=0):
    return qkv.new_empty(qkv.shape[1:]), qkv.new_empty(qkv.shape[1:3], dtype=torch.float32)


@torch.library.register_fake("flash_attention_ipu::qkv_packed_bwd")
def _(dout, qkv, out, lse, softmax_scale, causal, block_q, block_kv, kv_lens=None, dropout_p=0.0, dropout_seed=0):
    return torch.empty_like(qkv)
