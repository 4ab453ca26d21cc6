"""Synthetic execve/openat traces for the BASELINE.json workloads.

The reference ships no trace (only attack_chain.sh and one screenshot), so the configs
are synthesised by csrc/tracegen.cuh: record i is a pure integer function of (cfg, i),
identical on the host and in the CUDA twin.
"""
import ctypes
from ctypes import byref, c_void_p

import numpy as np

from . import _ffi
from .records import RECORD_BYTES

# name -> generator parameters; sizes are the BASELINE.json ones, tests scale `n` down.
WORKLOADS = {
    # config 1: attack_chain.sh-style replay, ~10k events, 64 PIDs
    "attack_replay_10k": dict(n=10_000, seed=0xC0FFEE01, n_pids=64, pid_base=2700, pid_stride=1,
                              noise_permille=850, exec_permille=80, adv_permille=20, long_permille=30,
                              chain_period=0, chain_pid_base=0, zipf=0.0),
    # config 2: 10M events, 4K PIDs spread over the 22-bit pid space, 256-byte path fields
    "uniform_10m_4k": dict(n=10_000_000, seed=0xC0FFEE02, n_pids=4096, pid_base=300, pid_stride=1021,
                           noise_permille=900, exec_permille=60, adv_permille=10, long_permille=100,
                           chain_period=0, chain_pid_base=0, zipf=0.0),
    # config 3: 100M events, 1M PIDs (sharded over 8 GPUs by pid hash)
    "sharded_100m_1m": dict(n=100_000_000, seed=0xC0FFEE03, n_pids=1_000_000, pid_base=300, pid_stride=3,
                            noise_permille=900, exec_permille=60, adv_permille=10, long_permille=100,
                            chain_period=0, chain_pid_base=0, zipf=0.0),
    # config 4: 95% noise, Zipf(1.1) PIDs, planted single-PID dropper chains
    "skewed_zipf_chains": dict(n=10_000_000, seed=0xC0FFEE04, n_pids=4096, pid_base=300, pid_stride=1021,
                               noise_permille=950, exec_permille=40, adv_permille=10, long_permille=100,
                               chain_period=10_000, chain_pid_base=4_200_000, zipf=1.1),
    # config 5: stream of 64K-event micro-batches with persistent state
    "stream_64k": dict(n=65_536 * 1000, seed=0xC0FFEE05, n_pids=4096, pid_base=300, pid_stride=1021,
                       noise_permille=900, exec_permille=60, adv_permille=10, long_permille=100,
                       chain_period=0, chain_pid_base=0, zipf=0.0),
}


def zipf_cdf(n_pids, s):
    """u64 thresholds: index k is drawn when cdf[k-1] <= u < cdf[k] (cdf[n-1] = 2^64-1)."""
    w = 1.0 / np.power(np.arange(1, n_pids + 1, dtype=np.float64), s)
    c = np.cumsum(w)
    c /= c[-1]
    out = np.minimum(np.floor(c * 18446744073709551616.0), 18446744073709549568.0).astype(np.uint64)
    out[-1] = np.uint64(0xFFFFFFFFFFFFFFFF)
    return out


def make_cfg(spec):
    cfg = _ffi.TracegenCfg()
    cfg.seed = spec["seed"]
    cfg.n_pids = spec["n_pids"]
    cfg.pid_base = spec["pid_base"]
    cfg.pid_stride = spec["pid_stride"]
    cfg.noise_permille = spec["noise_permille"]
    cfg.exec_permille = spec["exec_permille"]
    cfg.adv_permille = spec["adv_permille"]
    cfg.long_permille = spec["long_permille"]
    cfg.chain_period = spec["chain_period"]
    cfg.chain_pid_base = spec["chain_pid_base"]
    cfg.use_cdf = 1 if spec.get("zipf", 0.0) > 0 else 0
    cdf = zipf_cdf(spec["n_pids"], spec["zipf"]) if cfg.use_cdf else None
    return cfg, cdf


def generate_host(spec, first=0, n=None):
    """Records [first, first+n) of the trace as a (n, 288) uint8 array (host twin)."""
    n = spec["n"] if n is None else int(n)
    cfg, cdf = make_cfg(spec)
    out = np.zeros((n, RECORD_BYTES), dtype=np.uint8)
    rc = _ffi.lib().chronos_tracegen_host(byref(cfg), cdf.ctypes.data if cdf is not None else None, int(first), n,
                                          out.ctypes.data)
    if rc != _ffi.OK:
        raise _ffi.ChronosError(rc, "chronos_tracegen_host")
    return out


def generate_host_nocuda(spec, first=0, n=None):
    """generate_host through the host-only helper library (tools/libchronos_tracegen_host.so):
    the same bytes without loading the CUDA library (bench.py's CPU reference arm)."""
    from . import _build

    n = spec["n"] if n is None else int(n)
    cfg, cdf = make_cfg(spec)
    L = ctypes.CDLL(_build.build_hostgen())
    L.chronos_hostgen.argtypes = [ctypes.POINTER(_ffi.TracegenCfg), c_void_p, ctypes.c_uint64, ctypes.c_uint64, c_void_p]
    out = np.zeros((n, RECORD_BYTES), dtype=np.uint8)
    rc = L.chronos_hostgen(byref(cfg), cdf.ctypes.data if cdf is not None else None, int(first), n, out.ctypes.data)
    if rc != 0:
        raise ValueError("chronos_hostgen: bad arguments")
    return out


class DeviceBuffer:
    """A raw device allocation made through the C ABI (no torch needed)."""

    def __init__(self, device, nbytes):
        self.device = int(device)
        self.nbytes = int(nbytes)
        p = c_void_p()
        rc = _ffi.lib().chronos_dev_alloc(self.device, self.nbytes, byref(p))
        if rc != _ffi.OK:
            raise _ffi.ChronosError(rc, f"device allocation of {nbytes} bytes failed")
        self.ptr = p.value

    def upload(self, arr):
        a = np.ascontiguousarray(arr)
        assert a.nbytes <= self.nbytes
        rc = _ffi.lib().chronos_memcpy_h2d(self.device, c_void_p(self.ptr), a.ctypes.data, a.nbytes)
        if rc != _ffi.OK:
            raise _ffi.ChronosError(rc, "h2d copy failed")

    def download(self, nbytes=None):
        nbytes = self.nbytes if nbytes is None else int(nbytes)
        out = np.zeros(nbytes, dtype=np.uint8)
        rc = _ffi.lib().chronos_memcpy_d2h(self.device, out.ctypes.data, c_void_p(self.ptr), nbytes)
        if rc != _ffi.OK:
            raise _ffi.ChronosError(rc, "d2h copy failed")
        return out

    def free(self):
        if self.ptr:
            _ffi.lib().chronos_dev_free(self.device, c_void_p(self.ptr))
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def generate_device(spec, device=0, first=0, n=None):
    """Same bytes generated by the CUDA twin straight into HBM; returns a DeviceBuffer."""
    n = spec["n"] if n is None else int(n)
    cfg, cdf = make_cfg(spec)
    buf = DeviceBuffer(device, max(n, 1) * RECORD_BYTES)
    cdf_buf = None
    if cdf is not None:
        cdf_buf = DeviceBuffer(device, cdf.nbytes)
        cdf_buf.upload(cdf)
    rc = _ffi.lib().chronos_tracegen_device(int(device), byref(cfg), c_void_p(cdf_buf.ptr) if cdf_buf else None,
                                            int(first), n, c_void_p(buf.ptr))
    if cdf_buf is not None:
        cdf_buf.free()
    if rc != _ffi.OK:
        raise _ffi.ChronosError(rc, "chronos_tracegen_device")
    return buf
