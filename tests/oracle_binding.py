"""Test-side binding of the CPU oracle (oracle/liboe_oracle.so).  Only tests, smoke() and bench.py's CPU baseline
legs import this; nothing under dcos_commons_b200/ does."""
import ctypes as C
import os
import subprocess

import numpy as np

from dcos_commons_b200.abi import BoundLibrary, Handle, OeStats

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_LIB = os.path.join(ORACLE_DIR, "liboe_oracle.so")
_lib = None


def oracle_library() -> BoundLibrary:
    global _lib
    if _lib is None:
        src = os.path.join(ORACLE_DIR, "oe_oracle.cpp")
        if not os.path.exists(ORACLE_LIB) or os.path.getmtime(ORACLE_LIB) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])
        _lib = BoundLibrary(ORACLE_LIB, "oeo_")
    return _lib


class Oracle(Handle):
    def __init__(self, cached=False, threads=1):
        """cached=True: the oracle's second mode (tallies built once per cycle, see oracle/oe_oracle.cpp) — pinned to the
        literal mode by tests/test_oracle_modes.py; for tables the literal walk cannot finish."""
        super().__init__(oracle_library())
        if cached or threads > 1:
            self.set_mode(cached, threads)

    def set_mode(self, cached: bool, threads: int = 1):
        f = self.b.fn("set_mode")
        f.argtypes = [C.c_void_p, C.c_int32, C.c_int32]
        self._check(f(self.h, 1 if cached else 0, int(threads)))

    def feasibility_row(self, j: int) -> np.ndarray:
        """One matrix row ([stride] words) without allocating the whole matrix."""
        f = self.b.fn("feasibility_row")
        f.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
        row = np.zeros(self.stride(), np.uint32)
        self._check(f(self.h, int(j), row.ctypes.data))
        return row

    def stage_failures(self):
        """[n_reqs, n_offers] uint64 failed-stage masks and the [n_reqs, stride] bit matrix."""
        f = self.b.fn("feasibility_ex")
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        M, N = self.tables.n_reqs, self.tables.n_offers
        bits = np.zeros((M, self.stride()), np.uint32)
        fails = np.zeros((M, N), np.uint64)
        self._check(f(self.h, bits.ctypes.data, fails.ctypes.data, 0, M))
        return fails, bits

    def feasibility_rows(self, j0, j1):
        f = self.b.fn("feasibility_ex")
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64]
        bits = np.zeros((self.tables.n_reqs, self.stride()), np.uint32)
        self._check(f(self.h, bits.ctypes.data, None, j0, j1))
        return bits[j0:j1]

    def pairs_evaluated(self) -> int:
        f = self.b.fn("pairs_evaluated")
        f.argtypes, f.restype = [C.c_void_p], C.c_int64
        return int(f(self.h))
