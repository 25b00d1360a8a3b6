"""TEST INFRASTRUCTURE ONLY -- stand-in for flax (see ../jax/__init__.py)."""
from . import struct  # noqa: F401
