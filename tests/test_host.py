"""CPU tests of the host side: the C-ABI library loads and exports every symbol that
include/fa_b200.h declares, argument errors are reported without touching a GPU, and the Python
surface mirrors the reference's (names, ValueError / NotImplementedError conditions, patching)."""
import ctypes
import re

import pytest
import torch

from conftest import ROOT


def test_library_exports_every_declared_symbol(built):
    header = (ROOT / "include" / "fa_b200.h").read_text()
    declared = set(re.findall(r"\b(fa_b200_[a-z_]+)\s*\(", header))
    assert {"fa_b200_fwd", "fa_b200_bwd", "fa_b200_merge"} <= declared
    lib = ctypes.CDLL(str(built / "libflash_attention.so"))
    for sym in declared:
        assert hasattr(lib, sym), sym
    from flash_attention_ipu_b200._impl import native
    assert set(native.EXPORTED_SYMBOLS) == declared
    assert native.version() >= 100


def test_abi_rejects_bad_arguments_without_gpu(built):
    from flash_attention_ipu_b200._impl import native
    lib = ctypes.CDLL(str(built / "libflash_attention.so"))
    lib.fa_b200_last_error.restype = ctypes.c_char_p
    prob = native.Problem(1, 4, 0, 128, 128, native.FA_B200_BF16, 1, 1.0, 0, 0, 0, 0)
    t = native.Tensor(0, 0, 0, 0)
    rc = lib.fa_b200_fwd(ctypes.byref(prob), ctypes.byref(t), ctypes.byref(t), ctypes.byref(t),
                         ctypes.byref(t), None, None)
    assert rc == -1 and b"non-positive" in lib.fa_b200_last_error()
    prob.Lq = 128
    rc = lib.fa_b200_fwd(ctypes.byref(prob), ctypes.byref(t), ctypes.byref(t), ctypes.byref(t),
                         ctypes.byref(t), None, None)
    assert rc == -1 and b"null tensor" in lib.fa_b200_last_error()
    assert native.kernel_family(prob) == 1
    prob.dtype = native.FA_B200_F32
    assert native.kernel_family(prob) == 0
    prob.D = 512
    assert native.kernel_family(prob) == -2
    prob.D, prob.dtype = 128, native.FA_B200_BF16
    lib.fa_b200_bwd_workspace_bytes.restype = ctypes.c_size_t
    ws = lib.fa_b200_bwd_workspace_bytes(ctypes.byref(prob))
    assert ws >= 4 * 128 * 4 + 4 * 128 * 128 * 4


def test_packed_api_value_errors():
    # _impl/flash_attention_qkv_packed.py:20-35
    import flash_attention_ipu as fa
    with pytest.raises(ValueError, match="4 dimensions"):
        fa.flash_attention_qkv_packed(torch.zeros(3, 4, 8), 1, 1)
    with pytest.raises(ValueError, match="size 3 at dimension 0"):
        fa.flash_attention_qkv_packed(torch.zeros(2, 1, 4, 8), 1, 1)
    with pytest.raises(ValueError, match="num_chunks_q"):
        fa.flash_attention_qkv_packed(torch.zeros(3, 1, 6, 8), 4, 1)
    with pytest.raises(ValueError, match="num_chunks_kv"):
        fa.flash_attention_qkv_packed(torch.zeros(3, 1, 6, 8), 1, 4)
    assert fa.__all__ == ["flash_attention_qkv_packed"]


def test_no_cpu_fallback():
    import flash_attention_ipu as fa
    with pytest.raises(RuntimeError, match="no CPU path"):
        fa.flash_attention_qkv_packed(torch.zeros(3, 1, 8, 8), 1, 1)


def test_product_never_imports_oracle():
    for path in list((ROOT / "flash_attention_ipu_b200").rglob("*.py")) + list((ROOT / "flash_attention_ipu").rglob("*.py")) + \
            list((ROOT / "flash_attention_ipu_b200" / "csrc").glob("*")):
        text = path.read_text()
        assert "import oracle" not in text and "from oracle" not in text and "vanilla_attention.c" not in text, path


def test_sdpa_patch_and_gates():
    import torch.nn.functional as F
    import flash_attention_ipu.auto  # noqa: F401
    import flash_attention_ipu_b200._impl.scaled_dot_product_attention as mod
    mod.install()
    try:
        patched = F.scaled_dot_product_attention
        assert hasattr(patched, "__wrapped__")  # _impl/utils.py:29 (functools.wraps)
        q = torch.randn(2, 8, 4)
        # _impl/scaled_dot_product_attention.py:21-41 -- gates fire before any device check
        with pytest.raises(NotImplementedError, match="attn_mask"):
            F.scaled_dot_product_attention(q, q, q, attn_mask=torch.ones(8, 8, dtype=torch.bool))
        with pytest.raises(NotImplementedError, match="dropout_p"):
            F.scaled_dot_product_attention(q, q, q, dropout_p=0.1, is_causal=True)
        with pytest.raises(NotImplementedError, match="Multi-query"):
            F.scaled_dot_product_attention(torch.randn(2, 2, 8, 4), torch.randn(2, 1, 8, 4), torch.randn(2, 1, 8, 4), is_causal=True)
        # off-device tensors pass through to the original function (:60-61)
        out = F.scaled_dot_product_attention(q, q, q, is_causal=True)
        ref = patched.__wrapped__(q, q, q, is_causal=True)
        assert torch.equal(out, ref)
    finally:
        mod.uninstall()
    assert not hasattr(F.scaled_dot_product_attention, "__wrapped__")


def test_chunk_heuristic_matches_reference_formula():
    from flash_attention_ipu_b200._impl.scaled_dot_product_attention import _chunk_heuristic
    from oracle.attention_oracle import heuristic_chunks
    for L in (64, 256, 512, 1024, 4096, 8192, 131072):
        for D in (64, 128):
            assert _chunk_heuristic(L, D) == heuristic_chunks(L, D)


def test_ops_library_registers_the_four_ops():
    """csrc/torch_ops.cpp: loading build/libfa_torch_ops.so defines the ops (no Python fallback)."""
    from flash_attention_ipu_b200._impl import ops
    assert ops.ops_library_path.endswith("libfa_torch_ops.so")
    for name in ("attention_fwd", "attention_bwd", "qkv_packed_fwd", "qkv_packed_bwd"):
        assert hasattr(torch.ops.flash_attention_ipu, name)
    x = torch.randn(1, 2, 8, 8)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.attention_fwd(x, x, x, 1.0, True, 0, 0)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.qkv_packed_bwd(x[0], torch.randn(3, 2, 8, 8), x[0], torch.randn(2, 8), 1.0, True, 0, 0)


def test_varlen_argument_checks():
    """flash_attention_ipu_b200/varlen.py refuses malformed boundaries and CPU tensors (there is no CPU path)."""
    from flash_attention_ipu_b200.varlen import flash_attention_varlen
    q = torch.randn(10, 2, 64)
    with pytest.raises(RuntimeError, match="no CPU path"):
        flash_attention_varlen(q, q, q, [0, 4, 10], [0, 4, 10])
    with pytest.raises(ValueError):
        flash_attention_varlen(q, q, q[:, :, :32], [0, 10], [0, 10])
    with pytest.raises(ValueError):
        flash_attention_varlen(q, q, q, [1, 10], [0, 10])
