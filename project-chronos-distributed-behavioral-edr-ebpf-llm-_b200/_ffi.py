"""ctypes binding of include/chronos.h. There is no fallback: a missing or unloadable
CUDA library is an ImportError-grade failure for every entry point of the package."""
import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_uint8, c_uint32, c_uint64, c_void_p

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libchronos_b200.so")
RECORD_BYTES = 288

OK, E_ARG, E_CUDA, E_NOMEM, E_CAPACITY, E_STATE = 0, -1, -2, -3, -4, -5
META_TRIG, META_CURL, META_CHMOD, META_EXEC = 1, 2, 4, 8
NOALLOC_MAX_EVENTS = 262144


class ChronosError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"chronos error {code}: {msg}")
        self.code = code


class Window(Structure):
    _fields_ = [("trigger_idx", c_uint64), ("start", c_uint64), ("len", c_uint32), ("pid", c_uint32),
                ("flags", c_uint32), ("reserved", c_uint32)]


class Counts(Structure):
    _fields_ = [("n_events", c_uint64), ("n_survivors", c_uint64), ("n_windows", c_uint64), ("n_arena", c_uint64),
                ("n_pending", c_uint64), ("n_chains", c_uint64), ("n_pids_seen", c_uint64), ("reserved", c_uint64)]


class DeviceView(Structure):
    _fields_ = [("arena", c_void_p), ("windows", c_void_p), ("survivor_idx", c_void_p), ("survivor_meta", c_void_p),
                ("pending_pid", c_void_p), ("pending_slot", c_void_p), ("pending_arrival", c_void_p)]


class TracegenCfg(Structure):
    _fields_ = [("seed", c_uint64), ("n_pids", c_uint32), ("pid_base", c_uint32), ("pid_stride", c_uint32),
                ("noise_permille", c_uint32), ("exec_permille", c_uint32), ("adv_permille", c_uint32),
                ("long_permille", c_uint32), ("chain_period", c_uint32), ("chain_pid_base", c_uint32),
                ("use_cdf", c_uint32)]


# every symbol include/chronos.h declares: name -> (restype, argtypes)
IPC_HANDLE_BYTES = 64


class Download(Structure):
    _fields_ = [("n_survivors", c_uint64), ("n_windows", c_uint64), ("n_window_records", c_uint64),
                ("off_survivor_idx", c_uint64), ("off_survivor_meta", c_uint64), ("off_windows", c_uint64),
                ("off_window_records", c_uint64), ("total_bytes", c_uint64)]


SIGNATURES = {
    "chronos_device_ok": (c_int, [c_int]),
    "chronos_create": (c_int, [c_int, c_uint64, POINTER(c_void_p)]),
    "chronos_create_sized": (c_int, [c_int, c_uint64, c_uint64, POINTER(c_void_p)]),
    "chronos_destroy": (c_int, [c_void_p]),
    "chronos_reset": (c_int, [c_void_p]),
    "chronos_set_overlap": (c_int, [c_void_p, c_int]),
    "chronos_set_host_trim": (c_int, [c_void_p, c_int]),
    "chronos_last_upload_bytes": (c_int, [c_void_p, POINTER(c_uint64)]),
    "chronos_last_error": (c_char_p, [c_void_p]),
    "chronos_submit": (c_int, [c_void_p, c_void_p, c_uint64, c_uint64]),
    "chronos_fetch": (c_int, [c_void_p, POINTER(Counts)]),
    "chronos_copy_survivors": (c_int, [c_void_p, c_void_p, c_void_p, c_uint64]),
    "chronos_copy_windows": (c_int, [c_void_p, c_void_p, c_uint64]),
    "chronos_copy_arena": (c_int, [c_void_p, c_void_p, c_uint64]),
    "chronos_copy_window_records": (c_int, [c_void_p, c_void_p, c_uint64]),
    "chronos_window_digests": (c_int, [c_void_p, c_void_p, c_uint64]),
    "chronos_copy_chains": (c_int, [c_void_p, c_void_p, c_void_p, c_uint64]),
    "chronos_state": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_uint64]),
    "chronos_state_pids": (c_int, [c_void_p, c_void_p, c_uint64]),
    "chronos_get_device_view": (c_int, [c_void_p, POINTER(DeviceView)]),
    "chronos_last_timing": (c_int, [c_void_p, POINTER(c_float), POINTER(c_float)]),
    "chronos_last_stage_times": (c_int, [c_void_p, POINTER(c_float), c_uint32]),
    "chronos_last_launches": (c_int, [c_void_p, POINTER(c_uint64)]),
    "chronos_stream": (c_void_p, [c_void_p]),
    "chronos_owner_of": (c_uint32, [c_uint32, c_uint32]),
    "chronos_filter_partition": (c_int, [c_void_p, c_void_p, c_uint64, c_uint64, c_uint32, c_void_p, c_uint64,
                                         c_void_p, c_uint64, c_void_p, c_void_p]),
    "chronos_submit_survivors": (c_int, [c_void_p, c_void_p, c_void_p, c_uint64, c_uint64]),
    "chronos_checkpoint_bytes": (c_int, [c_void_p, POINTER(c_uint64)]),
    "chronos_checkpoint_save": (c_int, [c_void_p, c_void_p, c_uint64]),
    "chronos_checkpoint_load": (c_int, [c_void_p, c_void_p, c_uint64]),
    "chronos_download_begin": (c_int, [c_void_p, c_void_p, c_uint64, POINTER(Download)]),
    "chronos_download_wait": (c_int, [c_void_p]),
    "chronos_exchange_create": (c_int, [c_void_p, c_uint32, c_uint32, c_uint64, c_uint64, POINTER(c_void_p), c_void_p]),
    "chronos_exchange_connect": (c_int, [c_void_p, c_void_p, c_void_p]),
    "chronos_filter_exchange": (c_int, [c_void_p, c_void_p, c_uint64, c_uint64]),
    "chronos_submit_exchanged": (c_int, [c_void_p]),
    "chronos_exchange_set_timeout": (c_int, [c_void_p, ctypes.c_double]),
    "chronos_filter_noalloc": (c_int, [c_int, c_void_p, c_void_p, c_uint64, c_void_p, c_void_p, c_void_p, c_void_p,
                                       c_char_p, c_uint64]),
    "chronos_tracegen_host": (c_int, [POINTER(TracegenCfg), c_void_p, c_uint64, c_uint64, c_void_p]),
    "chronos_tracegen_device": (c_int, [c_int, POINTER(TracegenCfg), c_void_p, c_uint64, c_uint64, c_void_p]),
    "chronos_dev_alloc": (c_int, [c_int, c_uint64, POINTER(c_void_p)]),
    "chronos_dev_free": (c_int, [c_int, c_void_p]),
    "chronos_host_alloc_pinned": (c_int, [c_uint64, POINTER(c_void_p)]),
    "chronos_host_free_pinned": (c_int, [c_void_p]),
    "chronos_memcpy_h2d": (c_int, [c_int, c_void_p, c_void_p, c_uint64]),
    "chronos_memcpy_d2h": (c_int, [c_int, c_void_p, c_void_p, c_uint64]),
    "chronos_flush_l2": (c_int, [c_int]),
    "chronos_device_synchronize": (c_int, [c_int]),
    "chronos_device_mem_info": (c_int, [c_int, POINTER(c_uint64), POINTER(c_uint64)]),
}

_lib = None


def lib():
    """Load libchronos_b200.so (built by _build.build / __graft_entry__.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                "chronos_b200 has no CPU fallback.")
        # CHRONOS_LIB: a differently tuned build of the same sources (tools/ab_build.sh), experiments only
        L = ctypes.CDLL(os.environ.get("CHRONOS_LIB") or LIB_PATH)
        for name, (rt, at) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = rt
            fn.argtypes = at
        _lib = L
    return _lib
