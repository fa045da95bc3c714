"""``patch_function`` -- same behaviour as the reference's ``_impl/utils.py:10-39``: install a
wrapper that receives the original function first, keep ``__wrapped__`` (functools.wraps) so the
patch can be undone, and log one info line per patched module."""
import functools
import logging
from typing import Any, Callable, List

logger = logging.getLogger(__name__)


def patch_function(orig_fn: Callable, modules_to_patch: List[Any]):
    def decorator(patched_fn: Callable):
        @functools.wraps(orig_fn)
        def patch_wrapper(*args, **kwargs):
            return patched_fn(orig_fn, *args, **kwargs)

        for module in modules_to_patch:
            logger.info(f"flash_attention_ipu: patching {orig_fn.__name__}!")
            setattr(module, orig_fn.__name__, patch_wrapper)
        return patch_wrapper

    return decorator


def unpatch_function(module: Any, name: str) -> None:
    """Undo ``patch_function`` (tests use this; the reference relies on ``__wrapped__`` by hand)."""
    fn = getattr(module, name)
    while hasattr(fn, "__wrapped__"):
        fn = fn.__wrapped__
    setattr(module, name, fn)
