"""beifong_b200 — B200-native radar path-integration hot path of JacobMackay/beifong behind a C ABI.

`FlatScene` wraps the ABI handle; `configs` holds the BASELINE.json configurations.  The native
library is loaded on first use and its absence is an error (no CPU fallback).
"""
from . import _native  # noqa: F401
from .scene import FlatScene, BeifongError  # noqa: F401

__all__ = ["FlatScene", "BeifongError"]
