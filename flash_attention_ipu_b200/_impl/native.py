"""ctypes binding of the C ABI declared in ``include/fa_b200.h``.

This is the only place Python touches the native library.  Tensors cross the boundary as raw
device pointers + element strides; the stream is torch's current CUDA stream.  Errors come back
as negative status codes and are raised as ``RuntimeError`` with ``fa_b200_last_error()``.
"""
import ctypes
from ctypes import POINTER, Structure, byref, c_char_p, c_float, c_int, c_int32, c_int64, c_size_t, c_uint64, c_void_p

import torch

from .. import _lib, _lib_path  # loaded (or ImportError raised) by the package __init__

FA_B200_F16, FA_B200_BF16, FA_B200_F32 = 0, 1, 2
_DTYPES = {torch.float16: FA_B200_F16, torch.bfloat16: FA_B200_BF16, torch.float32: FA_B200_F32}

EXPORTED_SYMBOLS = (
    "fa_b200_fwd", "fa_b200_bwd", "fa_b200_bwd_workspace_bytes", "fa_b200_bwd_dq_accum", "fa_b200_bwd_noalloc", "fa_b200_bwd_sm_clock",
    "fa_b200_merge",
    "fa_b200_merge_acc",
    "fa_b200_kernel_family", "fa_b200_dropout_keep_host", "fa_b200_set_deterministic", "fa_b200_launch_count", "fa_b200_tmap_cache_stats", "fa_b200_last_error",
    "fa_b200_version",
)


class Tensor(Structure):  # fa_b200_tensor
    _fields_ = [("ptr", c_void_p), ("stride_b", c_int64), ("stride_h", c_int64), ("stride_l", c_int64)]


class Problem(Structure):  # fa_b200_problem
    _fields_ = [
        ("B", c_int32), ("H", c_int32), ("Lq", c_int32), ("Lk", c_int32), ("D", c_int32),
        ("dtype", c_int32), ("causal", c_int32), ("softmax_scale", c_float),
        ("block_q_hint", c_int32), ("block_kv_hint", c_int32), ("q_pos0", c_int32), ("kv_pos0", c_int32),
        ("H_kv", c_int32), ("kv_lens", c_void_p), ("dropout_p", c_float), ("dropout_seed", c_uint64),
        ("bias", c_void_p), ("bias_stride_b", c_int64), ("bias_stride_h", c_int64), ("bias_stride_l", c_int64),
    ]


_PT, _PP = POINTER(Tensor), POINTER(Problem)
_lib.fa_b200_fwd.argtypes = [_PP, _PT, _PT, _PT, _PT, c_void_p, c_void_p]
_lib.fa_b200_fwd.restype = c_int
_lib.fa_b200_bwd.argtypes = [_PP, _PT, _PT, _PT, _PT, _PT, c_void_p, _PT, _PT, _PT, c_void_p, c_size_t, c_void_p]
_lib.fa_b200_bwd.restype = c_int
_lib.fa_b200_bwd_workspace_bytes.argtypes = [_PP]
_lib.fa_b200_bwd_workspace_bytes.restype = c_size_t
_lib.fa_b200_bwd_dq_accum.argtypes = _lib.fa_b200_bwd.argtypes
_lib.fa_b200_bwd_dq_accum.restype = c_int
_lib.fa_b200_bwd_noalloc.argtypes = [_PP, _PT, _PT, _PT, _PT, _PT, c_void_p, _PT, _PT, _PT, c_void_p]
_lib.fa_b200_bwd_noalloc.restype = c_int
_lib.fa_b200_bwd_sm_clock.argtypes = [_PP, c_void_p, POINTER(ctypes.c_uint64), POINTER(ctypes.c_uint64)]
_lib.fa_b200_bwd_sm_clock.restype = c_int
_lib.fa_b200_merge.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32, c_void_p]
_lib.fa_b200_merge.restype = c_int
_lib.fa_b200_merge_acc.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64,
                                   c_int32, c_int32, c_void_p]
_lib.fa_b200_merge_acc.restype = c_int
_lib.fa_b200_dropout_keep_host.argtypes = [_PP, c_int32, c_int32, c_int32, c_int32, c_void_p, POINTER(c_float)]
_lib.fa_b200_dropout_keep_host.restype = c_int
_lib.fa_b200_set_deterministic.argtypes = [c_int]
_lib.fa_b200_set_deterministic.restype = c_int
_lib.fa_b200_kernel_family.argtypes = [_PP]
_lib.fa_b200_kernel_family.restype = c_int
_lib.fa_b200_launch_count.argtypes = []
_lib.fa_b200_launch_count.restype = c_int64
_lib.fa_b200_tmap_cache_stats.argtypes = [POINTER(c_int64), POINTER(c_int64)]
_lib.fa_b200_tmap_cache_stats.restype = None
_lib.fa_b200_last_error.argtypes = []
_lib.fa_b200_last_error.restype = c_char_p
_lib.fa_b200_version.argtypes = []
_lib.fa_b200_version.restype = c_int


def library_path() -> str:
    return _lib_path


def launch_count() -> int:
    return int(_lib.fa_b200_launch_count())


def tmap_cache_stats() -> tuple:
    """(hits, misses) of the shim's TMA-descriptor cache since the library was loaded."""
    hits, misses = c_int64(0), c_int64(0)
    _lib.fa_b200_tmap_cache_stats(byref(hits), byref(misses))
    return int(hits.value), int(misses.value)


def set_deterministic(on: bool) -> bool:
    """Ordered (bit-reproducible) dQ accumulation for every later tensor-family backward of this
    process; returns the previous setting."""
    return bool(_lib.fa_b200_set_deterministic(int(bool(on))))


def version() -> int:
    return int(_lib.fa_b200_version())


def _check(rc: int, what: str) -> None:
    if rc != 0:
        msg = _lib.fa_b200_last_error().decode(errors="replace")
        raise RuntimeError(f"{what} failed (status {rc}): {msg}")


def _tensor(t: torch.Tensor) -> Tensor:
    # t is (B, H, L, D) with stride(-1) == 1
    return Tensor(t.data_ptr(), t.stride(0), t.stride(1), t.stride(2))


def _require_cuda(*ts: torch.Tensor) -> None:
    for t in ts:
        if not t.is_cuda:
            raise RuntimeError(
                "flash_attention_ipu_b200 runs on CUDA (sm_100a) tensors only; there is no CPU path"
            )


def _bias_fields(bias, q, k):
    """(pointer, stride_b, stride_h, stride_l) of an additive score bias broadcastable to (B, H, Lq, Lk)."""
    if bias is None:
        return None, 0, 0, 0
    B, H, Lq, _ = q.shape
    Lk = k.shape[2]
    if not (bias.is_cuda and bias.dtype == q.dtype and bias.ndim == 4 and tuple(bias.shape[2:]) == (Lq, Lk)
            and bias.shape[0] in (1, B) and bias.shape[1] in (1, H) and bias.stride(3) == 1):
        raise RuntimeError("bias must be a CUDA tensor of q's dtype, shaped (B|1, H|1, Lq, Lk), unit stride on Lk")
    return (bias.data_ptr(), bias.stride(0) if bias.shape[0] > 1 else 0, bias.stride(1) if bias.shape[1] > 1 else 0,
            bias.stride(2))


def make_problem(q: torch.Tensor, k: torch.Tensor, causal: bool, scale: float,
                 block_q: int = 0, block_kv: int = 0, q_pos0: int = 0, kv_pos0: int = 0,
                 kv_lens: "torch.Tensor | None" = None, dropout_p: float = 0.0, dropout_seed: int = 0,
                 bias: "torch.Tensor | None" = None) -> Problem:
    if q.dtype not in _DTYPES:
        raise RuntimeError(f"unsupported dtype {q.dtype}")
    B, H, Lq, D = q.shape
    H_kv = k.shape[1]  # fewer K/V heads than query heads = grouped-query attention
    if H % H_kv:
        raise RuntimeError(f"{H} query heads are not a multiple of {H_kv} key/value heads")
    lens_ptr = None
    if kv_lens is not None:  # the caller keeps the tensor alive until the launches that use the problem are enqueued
        if not (kv_lens.is_cuda and kv_lens.dtype == torch.int32 and kv_lens.is_contiguous() and kv_lens.numel() == B * H):
            raise RuntimeError("kv_lens must be a contiguous CUDA int32 tensor with B*H entries")
        lens_ptr = kv_lens.data_ptr()
    return Problem(B, H, Lq, k.shape[2], D, _DTYPES[q.dtype], int(bool(causal)), float(scale),
                   int(block_q), int(block_kv), int(q_pos0), int(kv_pos0), 0 if H_kv == H else H_kv, lens_ptr,
                   float(dropout_p), int(dropout_seed) & 0xFFFFFFFFFFFFFFFF, *_bias_fields(bias, q, k))


def dropout_keep_host(prob: Problem, group: int, q_row: int, k_first: int = 0, n: "int | None" = None):
    """(keep, scale): the dropout mask bits of keys ``k_first .. k_first + n`` of one query row, computed on the
    host by the functions the kernels compile; ``keep`` is a uint8 CPU tensor."""
    n = prob.Lk - k_first if n is None else n
    keep = torch.empty(n, dtype=torch.uint8)
    scale = c_float(0.0)
    _check(_lib.fa_b200_dropout_keep_host(byref(prob), group, q_row, k_first, n, keep.data_ptr(), byref(scale)),
           "fa_b200_dropout_keep_host")
    return keep, float(scale.value)


def kernel_family(prob: Problem) -> int:
    return int(_lib.fa_b200_kernel_family(byref(prob)))


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def fwd(prob: Problem, q, k, v, o, lse) -> None:
    _require_cuda(q, k, v, o, lse)
    with torch.cuda.device(q.device):
        rc = _lib.fa_b200_fwd(byref(prob), byref(_tensor(q)), byref(_tensor(k)), byref(_tensor(v)),
                              byref(_tensor(o)), lse.data_ptr(), _stream(q))
    _check(rc, "fa_b200_fwd")


def bwd_workspace_bytes(prob: Problem) -> int:
    return int(_lib.fa_b200_bwd_workspace_bytes(byref(prob)))


def bwd(prob: Problem, q, k, v, o, dout, lse, dq, dk, dv, workspace) -> None:
    _require_cuda(q, k, v, o, dout, lse, dq, dk, dv, workspace)
    with torch.cuda.device(q.device):
        rc = _lib.fa_b200_bwd(byref(prob), byref(_tensor(q)), byref(_tensor(k)), byref(_tensor(v)),
                              byref(_tensor(o)), byref(_tensor(dout)), lse.data_ptr(),
                              byref(_tensor(dq)), byref(_tensor(dk)), byref(_tensor(dv)),
                              workspace.data_ptr(), workspace.numel() * workspace.element_size(),
                              _stream(q))
    _check(rc, "fa_b200_bwd")


NOALLOC_SCRATCH_BYTES = 32 << 20  # FA_B200_NOALLOC_SCRATCH_BYTES of include/fa_b200.h


def bwd_noalloc(prob: Problem, q, k, v, o, dout, lse, dq, dk, dv) -> None:
    """fa_b200_bwd_noalloc: :func:`bwd` with the library's static device scratch as workspace (no
    allocation inside the call; problems needing more than NOALLOC_SCRATCH_BYTES are refused; calls
    that overlap on one GPU race on the shared scratch)."""
    _require_cuda(q, k, v, o, dout, lse, dq, dk, dv)
    with torch.cuda.device(q.device):
        rc = _lib.fa_b200_bwd_noalloc(byref(prob), byref(_tensor(q)), byref(_tensor(k)), byref(_tensor(v)),
                                      byref(_tensor(o)), byref(_tensor(dout)), lse.data_ptr(),
                                      byref(_tensor(dq)), byref(_tensor(dk)), byref(_tensor(dv)), _stream(q))
    _check(rc, "fa_b200_bwd_noalloc")


def bwd_dq_accum(prob: Problem, q, k, v, o, dout, lse, dq_acc, dk, dv, workspace) -> None:
    """fa_b200_bwd_dq_accum: ``dq_acc`` is an fp32 (B, H, Lq, D) accumulator that receives
    ``dS K`` (no softmax scale) by in-kernel reduce-add; dk, dv as in :func:`bwd`."""
    _require_cuda(q, k, v, o, dout, lse, dq_acc, dk, dv, workspace)
    if dq_acc.dtype != torch.float32:
        raise RuntimeError("dq_acc must be fp32")
    with torch.cuda.device(q.device):
        rc = _lib.fa_b200_bwd_dq_accum(byref(prob), byref(_tensor(q)), byref(_tensor(k)), byref(_tensor(v)),
                                       byref(_tensor(o)), byref(_tensor(dout)), lse.data_ptr(),
                                       byref(_tensor(dq_acc)), byref(_tensor(dk)), byref(_tensor(dv)),
                                       workspace.data_ptr(), workspace.numel() * workspace.element_size(),
                                       _stream(q))
    _check(rc, "fa_b200_bwd_dq_accum")


def bwd_sm_clock(prob: Problem, workspace) -> tuple:
    """(SM cycles, nanoseconds) CTA 0 of the last tensor-family backward run with `workspace` spent
    on its SM; cycles / ns * 1e3 = effective SM clock in MHz under load.  Synchronise first."""
    cyc, ns = ctypes.c_uint64(0), ctypes.c_uint64(0)
    _check(_lib.fa_b200_bwd_sm_clock(byref(prob), workspace.data_ptr(), byref(cyc), byref(ns)), "fa_b200_bwd_sm_clock")
    return int(cyc.value), int(ns.value)


def merge(o_a, lse_a, o_b, lse_b) -> None:
    """In-place LSE merge of two partial attention results (ring hop)."""
    _require_cuda(o_a, lse_a, o_b, lse_b)
    D = o_a.shape[-1]
    rows = o_a.numel() // D
    with torch.cuda.device(o_a.device):
        rc = _lib.fa_b200_merge(o_a.data_ptr(), lse_a.data_ptr(), o_b.data_ptr(), lse_b.data_ptr(),
                                rows, D, _DTYPES[o_a.dtype], _stream(o_a))
    _check(rc, "fa_b200_merge")


def merge_acc(o_acc, lse_acc, o_b, lse_b, row0: int = 0) -> None:
    """Fold a hop's partial result into the fp32 ring accumulators, in place.

    o_acc (G, La, D) fp32, lse_acc (G, La) fp32, o_b (G, Lb, D) kernel dtype, lse_b (G, Lb) fp32,
    all contiguous; partial row r of a group lands on accumulator row ``row0 + r``."""
    _require_cuda(o_acc, lse_acc, o_b, lse_b)
    assert o_acc.dtype == torch.float32 and lse_acc.dtype == torch.float32 and lse_b.dtype == torch.float32
    assert o_acc.is_contiguous() and lse_acc.is_contiguous() and o_b.is_contiguous() and lse_b.is_contiguous()
    G, La, D = o_acc.shape
    Gb, Lb, Db = o_b.shape
    assert G == Gb and D == Db and row0 + Lb <= La
    with torch.cuda.device(o_acc.device):
        rc = _lib.fa_b200_merge_acc(o_acc.data_ptr(), lse_acc.data_ptr(), o_b.data_ptr(), lse_b.data_ptr(),
                                    G, Lb, La, row0, D, _DTYPES[o_b.dtype], _stream(o_acc))
    _check(rc, "fa_b200_merge_acc")
