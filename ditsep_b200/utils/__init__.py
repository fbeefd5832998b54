from .registry import Registry
from .separate import denormalize_batch, normalize_batch

__all__ = ["Registry", "normalize_batch", "denormalize_batch"]
