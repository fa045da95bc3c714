"""Packaging of the B200-native FlashAttention behind the flash-attention-ipu API.

Same shape as the reference's packaging (``setup.py:10-40`` there: a ``build_ext`` that shells out
to ``make OUT=...``): the native libraries are built by the repo's Makefile with nvcc for sm_100a
and land NEXT TO the Python package, where ``flash_attention_ipu_b200/__init__.py`` looks for them
(third entry of its search list).  ``pip install .`` therefore needs the CUDA toolkit (nvcc) and
PyTorch at build time -- use ``--no-build-isolation`` so that the installed torch is the one the
op library is compiled against.  No Triton, no multi-arch fatbins, no CPU fallback.
"""
import subprocess
from pathlib import Path

import setuptools
import setuptools.command.build_ext

ROOT = Path(__file__).parent.resolve()
PKG = "flash_attention_ipu_b200"
LIBS = ("libflash_attention.so", "libfa_torch_ops.so")


class make_ext(setuptools.command.build_ext.build_ext):
    def build_extension(self, ext: setuptools.Extension) -> None:
        if ext.name != f"{PKG}.libflash_attention":
            return super().build_extension(ext)
        out = (ROOT / PKG) if self.inplace else (Path(self.build_lib).resolve() / PKG)
        out.mkdir(parents=True, exist_ok=True)
        subprocess.check_call(["make", "-C", str(ROOT), f"OUT={out}"] + [f"{out}/{lib}" for lib in LIBS])

    def get_outputs(self):
        out = (ROOT / PKG) if self.inplace else (Path(self.build_lib) / PKG)
        return [str(out / lib) for lib in LIBS]


setuptools.setup(
    name="flash-attention-ipu-b200",
    version="0.2.0",
    description="FlashAttention for NVIDIA B200 (sm_100a) behind the flash-attention-ipu Python API",
    python_requires=">=3.9",
    install_requires=["torch"],
    packages=[PKG, f"{PKG}._impl", "flash_attention_ipu"],
    package_data={PKG: ["csrc/*.cu", "csrc/*.cuh", "csrc/*.cpp"]},
    ext_modules=[
        setuptools.Extension(
            f"{PKG}.libflash_attention",
            sorted(str(p.relative_to(ROOT)) for p in (ROOT / PKG / "csrc").glob("*.c*")),
        )
    ],
    cmdclass=dict(build_ext=make_ext),
    zip_safe=False,
)
