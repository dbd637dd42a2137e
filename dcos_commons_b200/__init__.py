"""B200-native offer-evaluation engine for the DC/OS SDK scheduler's offer-matching path.

The product is the C-ABI library built from dcos_commons_b200/csrc (include/oe_abi.h).  This package is the thin
Python plumbing around it: ctypes bindings (abi), table assembly (tables), synthetic clusters (synth) and the
multi-GPU driver (sharded).  Importing the package does not load the library; creating an Engine does, and fails
loudly when the library is missing or no B200 is present.
"""
from .abi import CycleTables, Engine, EngineError, engine_library, unpack_bits  # noqa: F401

__all__ = ["CycleTables", "Engine", "EngineError", "engine_library", "unpack_bits"]
