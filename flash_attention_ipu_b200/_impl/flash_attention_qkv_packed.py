"""Public op ``flash_attention_qkv_packed`` -- same name, positional order, return value and
``ValueError`` conditions as the reference (``flash_attention_ipu/_impl/flash_attention_qkv_packed.py:7-55``).

Differences, all at the dispatch point (reference lines 38-53):
  * ``poptorch.isRunningOnIpu()`` -> "qkv is a CUDA tensor"; the op lowers to the
    ``flash_attention_ipu::qkv_packed_fwd`` torch.library op over the C ABI (sm_100a kernels).
  * there is no CPU branch: a non-CUDA tensor raises ``RuntimeError`` (the reference's six
    lines of CPU math live on only as the test oracle, ``oracle/attention_oracle.py``).
  * ``num_chunks_q`` / ``num_chunks_kv`` are tile-size *hints* (``L // num_chunks``); the result
    does not depend on them, which is also true of the reference
    (``tests/cpp/test_flash_attention.cpp:245-259``).
  * keyword-only extensions ``causal=True`` and ``softmax_scale=1.0`` keep the reference
    semantics by default (always causal, caller pre-scales q).
"""
import torch

from . import ops


def flash_attention_qkv_packed(
    qkv: torch.Tensor, num_chunks_q: int, num_chunks_kv: int, *, causal: bool = True,
    softmax_scale: float = 1.0, kv_lens: "torch.Tensor | None" = None
) -> torch.Tensor:
    """
    Memory-efficient causally masked multi-head attention from a packed qkv tensor

    Computes `nn.softmax(Q@K.T, dim=-1)@V` without materialising the full attention matrix

    qkv -- shape (3, N, L, D)
    returns -- shape (N, L, D)

    Keyword extensions of the B200 build (the reference has none of them): ``causal=False``,
    ``softmax_scale``, and ``kv_lens`` -- an int32 CUDA tensor with one entry per group n: only the
    first ``kv_lens[n]`` keys / values of group n exist, the rest is padding (its dk, dv are zero).
    """
    if qkv.ndim != 4:
        raise ValueError("flash_attention_qkv_packed expects qkv input to have 4 dimensions")
    if qkv.shape[0] != 3:
        raise ValueError("flash_attention_qkv_packed expects qkv input to have size 3 at dimension 0")
    if qkv.shape[2] % num_chunks_q != 0:
        raise ValueError(
            "flash_attention_qkv_packed expects qkv size at dimension 2 to be divisible by num_chunks_q"
        )
    if qkv.shape[2] % num_chunks_kv != 0:
        raise ValueError(
            "flash_attention_qkv_packed expects qkv size at dimension 2 to be divisible by num_chunks_kv"
        )
    if not qkv.is_cuda:
        raise RuntimeError(
            "flash_attention_qkv_packed: qkv must be a CUDA tensor (B200 / sm_100a); "
            "this build has no CPU path"
        )
    L = qkv.shape[2]
    out, _lse = ops.qkv_packed_fwd(qkv, float(softmax_scale), bool(causal),
                                   L // num_chunks_q, L // num_chunks_kv, kv_lens)
    return out


__all__ = ["flash_attention_qkv_packed"]
