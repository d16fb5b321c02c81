"""kgp_b200 -- B200-native GP inference engine for keras-gp's backend slot.

``GP_BACKEND`` mirrors ``kgp.backend.GP_BACKEND`` (kgp/backend/__init__.py:26-30):
set ``GP_BACKEND=b200`` semantics by importing it from here instead of kgp.backend.
"""
from .engine import B200GP, GP_BACKEND  # noqa: F401

__all__ = ['B200GP', 'GP_BACKEND']
