from .registry import Registry
from .separate import denormalize_batch, normalize_batch, power_order_sources, select_elem_at_random, shuffle_sources

__all__ = ["Registry", "normalize_batch", "denormalize_batch"]
