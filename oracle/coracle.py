"""ctypes front-end of oracle/chronos_oracle.c.  TEST INFRASTRUCTURE ONLY."""
import ctypes
from ctypes import POINTER, c_char_p, c_int, c_uint8, c_uint32, c_uint64, c_void_p

import numpy as np

from . import build_oracle

REC_BYTES = 288
_lib = None


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(build_oracle.build_oracle())
        L.co_create.restype = c_void_p
        L.co_destroy.argtypes = [c_void_p]
        L.co_feed.argtypes = [c_void_p, c_void_p, c_uint64, c_uint64]
        L.co_feed.restype = c_int
        L.co_clear_outputs.argtypes = [c_void_p]
        for name, rt in [
            ("co_n_survivors", c_uint64), ("co_survivors", POINTER(c_uint64)),
            ("co_n_windows", c_uint64), ("co_win_pid", POINTER(c_uint32)),
            ("co_win_trigger", POINTER(c_uint64)), ("co_win_flags", POINTER(c_uint32)),
            ("co_win_off", POINTER(c_uint64)), ("co_n_members", c_uint64),
            ("co_win_members", POINTER(c_uint64)), ("co_n_chains", c_uint64),
            ("co_chain_pid", POINTER(c_uint32)), ("co_chain_idx", POINTER(c_uint64)),
            ("co_state_npids", c_uint64), ("co_state_nmembers", c_uint64),
        ]:
            f = getattr(L, name)
            f.argtypes = [c_void_p]
            f.restype = rt
        L.co_state_dump.argtypes = [c_void_p, c_void_p, c_void_p, c_void_p]
        L.co_classify_batch.argtypes = [c_void_p, c_uint64, c_void_p, c_void_p]
        L.co_utf8_valid.argtypes = [c_char_p, c_int]
        L.co_utf8_valid.restype = c_int
        L.co_openat_noise.argtypes = [c_char_p]
        L.co_openat_noise.restype = c_int
        _lib = L
    return _lib


def _as_u8(records):
    if isinstance(records, (bytes, bytearray, memoryview)):
        return np.frombuffer(records, dtype=np.uint8)
    a = np.ascontiguousarray(records).view(np.uint8).reshape(-1)
    return a


def _arr(ptr, n, dtype):
    if n == 0:
        return np.zeros(0, dtype=dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True)


def classify(records):
    a = _as_u8(records)
    n = a.size // REC_BYTES
    keep = np.zeros(n, dtype=np.uint8)
    meta = np.zeros(n, dtype=np.uint8)
    lib().co_classify_batch(a.ctypes.data, n, keep.ctypes.data, meta.ctypes.data)
    return keep, meta


def utf8_valid(b: bytes) -> bool:
    return bool(lib().co_utf8_valid(b, len(b)))


def openat_noise(argv_field: bytes) -> bool:
    f = bytes(argv_field)[:256]
    f = f + b"\0" * (256 - len(f))
    return bool(lib().co_openat_noise(f))


class COracle:
    """Stateful engine (process_memory across feeds)."""

    def __init__(self):
        self.L = lib()
        self.h = self.L.co_create()

    def close(self):
        if self.h:
            self.L.co_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def feed(self, records, base=0):
        """Process one batch; returns this batch's outputs as numpy arrays."""
        a = _as_u8(records)
        n = a.size // REC_BYTES
        L, h = self.L, self.h
        L.co_clear_outputs(h)
        L.co_feed(h, a.ctypes.data, n, base)
        nw = L.co_n_windows(h)
        out = {
            "survivors": _arr(L.co_survivors(h), L.co_n_survivors(h), np.uint64),
            "win_pid": _arr(L.co_win_pid(h), nw, np.uint32),
            "win_trigger": _arr(L.co_win_trigger(h), nw, np.uint64),
            "win_flags": _arr(L.co_win_flags(h), nw, np.uint32),
            "win_off": _arr(L.co_win_off(h), nw + 1, np.uint64),
            "members": _arr(L.co_win_members(h), L.co_n_members(h), np.uint64),
            "chain_pid": _arr(L.co_chain_pid(h), L.co_n_chains(h), np.uint32),
            "chain_idx": _arr(L.co_chain_idx(h), L.co_n_chains(h), np.uint64),
        }
        return out

    def state(self):
        """Final process_memory as (pids ascending, offsets, member arrival indices)."""
        L, h = self.L, self.h
        k = L.co_state_npids(h)
        m = L.co_state_nmembers(h)
        pids = np.zeros(k, dtype=np.uint32)
        offs = np.zeros(k + 1, dtype=np.uint64)
        mem = np.zeros(max(m, 1), dtype=np.uint64)
        L.co_state_dump(h, pids.ctypes.data, offs.ctypes.data, mem.ctypes.data)
        return pids, offs, mem[:m]
