"""B200-native FlashAttention behind the flash-attention-ipu Python API.

Same loading scheme as the reference (``flash_attention_ipu/__init__.py:4-20``): find
``build/libflash_attention.so`` (or an installed copy next to the package), ``ctypes``-load it and
raise ``ImportError`` if it is missing -- there is no CPU or eager fallback.  Unlike the
reference's plugin, whose static initialisers register a popart op, this library exports the C
ABI of ``include/fa_b200.h``; ``_impl/native.py`` binds it and ``_impl/ops.py`` wraps it as
``torch.library`` ops.
"""


def _load_native_library():
    import ctypes
    import pathlib
    import sysconfig

    root = pathlib.Path(__file__).parent.parent.absolute()
    name = "libflash_attention.so"

    paths = [
        root / "build" / name,
        (root / name).with_suffix(sysconfig.get_config_vars()["EXT_SUFFIX"] or ".so"),
        pathlib.Path(__file__).parent / name,
    ]
    for path in paths:
        if path.exists():
            return ctypes.cdll.LoadLibrary(str(path)), str(path)
    raise ImportError(f"Cannot find {name} - tried {[str(p) for p in paths]} (run `make`)")


_lib, _lib_path = _load_native_library()

from ._impl.flash_attention_qkv_packed import *  # noqa: E402,F401,F403
from ._impl.flash_attention_qkv_packed import __all__  # noqa: E402,F401


def set_deterministic(on: bool) -> bool:
    """Bit-reproducible gradients: dQ partials are added in a fixed order (``fa_b200_set_deterministic``)
    for every later backward of this process, at a throughput cost.  Returns the previous setting.
    ``torch.use_deterministic_algorithms(True)`` selects the same mode for the torch ops."""
    from ._impl import native
    return native.set_deterministic(on)
