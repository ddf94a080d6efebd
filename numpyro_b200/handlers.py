"""``numpyro.handlers`` surface that applies to the registered families: ``reparam``."""
from .models import reparam

__all__ = ["reparam"]
