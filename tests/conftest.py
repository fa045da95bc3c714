import ctypes
import pathlib
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
GOLDEN = ROOT / "tests" / "golden"
SEED = 1123581321  # the reference tests' seed (tests/python/test_flash_attention_qkv_packed.py:59)


_LIBS = ("libflash_attention.so", "libfa_torch_ops.so", "liboracle.so")


def _ensure_built():
    """The package refuses to import without its native libraries (no fallback), so a fresh
    checkout is built once here (nvcc / g++ cross-compile without a GPU)."""
    missing = [name for name in _LIBS if not (ROOT / "build" / name).exists()]
    if missing:
        subprocess.check_call(["make", "-s"] + [f"build/{name}" for name in missing], cwd=ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    _ensure_built()


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def built():
    """Make sure the native libraries exist (nvcc cross-compiles without a GPU)."""
    _ensure_built()
    return ROOT / "build"


@pytest.fixture(scope="session")
def c_oracle(built):
    lib = ctypes.CDLL(str(built / "liboracle.so"))
    fp = ctypes.POINTER(ctypes.c_float)
    lib.oracle_vanilla_attention_fwd.argtypes = [fp, fp, fp] + [ctypes.c_int] * 4
    lib.oracle_vanilla_attention_bwd.argtypes = [fp, fp, fp] + [ctypes.c_int] * 4

    def ptr(a):
        return a.ctypes.data_as(fp)

    class Oracle:
        @staticmethod
        def fwd(qkv: np.ndarray, causal=True):
            _, G, L, D = qkv.shape
            qkv = np.ascontiguousarray(qkv, dtype=np.float32)
            out = np.empty((G, L, D), np.float32)
            lse = np.empty((G, L), np.float32)
            assert lib.oracle_vanilla_attention_fwd(ptr(qkv), ptr(out), ptr(lse), G, L, D, int(causal)) == 0
            return out, lse

        @staticmethod
        def bwd(grad: np.ndarray, qkv: np.ndarray, causal=True):
            _, G, L, D = qkv.shape
            qkv = np.ascontiguousarray(qkv, dtype=np.float32)
            grad = np.ascontiguousarray(grad, dtype=np.float32)
            dqkv = np.empty_like(qkv)
            assert lib.oracle_vanilla_attention_bwd(ptr(grad), ptr(qkv), ptr(dqkv), G, L, D, int(causal)) == 0
            return dqkv

    return Oracle


def packed_inputs(G, L, D, prescale=True):
    """Inputs exactly as oracle/gen_golden.py made them (seeded torch CPU generator)."""
    torch.manual_seed(SEED)
    qkv = torch.randn(3, G, L, D)
    grad = torch.randn(G, L, D)
    if prescale:
        qkv[0] *= D ** -0.5
    return qkv, grad


def sdpa_inputs(batch, L, D):
    torch.manual_seed(SEED)
    return torch.randn(3, *batch, L, D), torch.randn(*batch, L, D)
