"""Drop-in alias: ``import flash_attention_ipu`` / ``import flash_attention_ipu.auto`` keep working
for users of graphcore-research/flash-attention-ipu; everything lives in ``flash_attention_ipu_b200``."""
from flash_attention_ipu_b200 import *  # noqa: F401,F403
from flash_attention_ipu_b200 import __all__  # noqa: F401
