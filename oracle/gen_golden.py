"""Generate tests/golden/*.npz by running the REFERENCE's own Python code.

Run in the build container only (needs /root/reference):
    python oracle/gen_golden.py

How the reference is executed: ``flash_attention_ipu/__init__.py:4-20`` refuses to
import without libflash_attention.so and ``_impl/*.py`` import ``poptorch`` at
module top, neither of which exists here.  We therefore load the two ``_impl``
source files *unmodified, from where they lie*, by file path, with a stub
``poptorch`` module in ``sys.modules`` whose ``isRunningOnIpu()`` returns False,
so the reference's non-IPU branch (``_impl/flash_attention_qkv_packed.py:47-53``)
and its SDPA pass-through (``_impl/scaled_dot_product_attention.py:60-61``) run.
Gradients come from torch autograd through that reference code.

Inputs are regenerated in the tests from the same seeds (torch's CPU generator is
deterministic); the fixtures store outputs + gradients and a checksum of inputs.
Seeds/shapes follow tests/python/test_flash_attention_qkv_packed.py:59-62 and
tests/python/test_scaled_dot_product_attention.py:62-66 (seed 1123581321).
"""
import importlib.util
import pathlib
import sys
import types

import numpy as np
import torch

REF = pathlib.Path("/root/reference/flash_attention_ipu")
OUT = pathlib.Path(__file__).resolve().parent.parent / "tests" / "golden"
SEED = 1123581321

# (name, G, L, D) -- kept small so the fixtures stay ~1 MB in total
PACKED_CASES = [("packed_g2_l256_d128", 2, 256, 128), ("packed_g3_l128_d64", 3, 128, 64),
                ("packed_g4_l6_d2", 4, 6, 2)]
SDPA_CASES = [("sdpa_b2x3_l256_d64", (2, 3), 256, 64)]


def load_reference():
    stub = types.ModuleType("poptorch")
    stub.isRunningOnIpu = lambda: False
    sys.modules["poptorch"] = stub
    pkg = types.ModuleType("refpkg")
    pkg.__path__ = [str(REF / "_impl")]
    sys.modules["refpkg"] = pkg
    mods = {}
    for name in ("utils", "flash_attention_qkv_packed", "scaled_dot_product_attention"):
        spec = importlib.util.spec_from_file_location(f"refpkg.{name}", REF / "_impl" / f"{name}.py")
        m = importlib.util.module_from_spec(spec)
        sys.modules[f"refpkg.{name}"] = m
        if name == "scaled_dot_product_attention":
            saved = torch.nn.functional.scaled_dot_product_attention
        spec.loader.exec_module(m)
        mods[name] = m
    return mods, saved


def make_packed_inputs(G, L, D):
    torch.manual_seed(SEED)
    qkv = torch.randn(3, G, L, D)
    grad = torch.randn(G, L, D)
    return qkv, grad


def make_sdpa_inputs(batch, L, D):
    torch.manual_seed(SEED)
    qkv = torch.randn(3, *batch, L, D)
    grad = torch.randn(*batch, L, D)
    return qkv, grad


def main():
    mods, orig_sdpa = load_reference()
    ref_packed = mods["flash_attention_qkv_packed"].flash_attention_qkv_packed
    OUT.mkdir(parents=True, exist_ok=True)
    for name, G, L, D in PACKED_CASES:
        qkv, grad = make_packed_inputs(G, L, D)
        # the reference's only in-repo caller pre-scales q (scaled_dot_product_attention.py:55)
        qkv[0] *= D ** -0.5
        x = qkv.clone().requires_grad_(True)
        out = ref_packed(x, 1, 1)
        (out * grad).sum().backward()
        np.savez(OUT / f"{name}.npz", out=out.detach().numpy(), dqkv=x.grad.numpy(),
                 in_checksum=np.float64(qkv.double().sum().item() + grad.double().sum().item()),
                 shape=np.array([G, L, D]))
        print(name, tuple(out.shape), float(out.abs().max()))
    patched = torch.nn.functional.scaled_dot_product_attention  # reference's wrapper (auto.py:3)
    assert patched is not orig_sdpa and hasattr(patched, "__wrapped__")
    for name, batch, L, D in SDPA_CASES:
        qkv, grad = make_sdpa_inputs(batch, L, D)
        x = qkv.clone().requires_grad_(True)
        out = patched(x[0], x[1], x[2], is_causal=True)
        (out * grad).sum().backward()
        np.savez(OUT / f"{name}.npz", out=out.detach().numpy(), dqkv=x.grad.numpy(),
                 in_checksum=np.float64(qkv.double().sum().item() + grad.double().sum().item()),
                 shape=np.array([*batch, L, D]))
        print(name, tuple(out.shape), float(out.abs().max()))
    torch.nn.functional.scaled_dot_product_attention = orig_sdpa


if __name__ == "__main__":
    main()
