"""chronos_b200 -- B200-native implementation of Project CHRONOS's telemetry hot path:
openat noise filter + validity/ignore gates + per-PID context windows
(reference: chronos_sensor.py:18-47, :74-92, :105, :124-157).

Host code is Python over a C ABI (include/chronos.h -> libchronos_b200.so, hand-written
sm_100a CUDA). There is no CPU fallback: without the built library or without a B200
every entry point raises.
"""
from . import _ffi
from ._ffi import ChronosError, META_CHMOD, META_CURL, META_EXEC, META_TRIG  # noqa: F401
from .engine import BatchResult, Context, WINDOW_DTYPE  # noqa: F401
from .records import RECORD_BYTES, RECORD_DTYPE, event_str, pack_record  # noqa: F401


def owner_of(pid, world):
    """Rank that owns a PID under PID-hash sharding (same function the kernels use)."""
    return int(_ffi.lib().chronos_owner_of(int(pid) & 0xFFFFFFFF, int(world)))
