"""``import flash_attention_ipu.auto`` -- same side effect as the reference's ``auto.py:3``."""
from flash_attention_ipu_b200.auto import _scaled_dot_product_attention_b200  # noqa: F401
