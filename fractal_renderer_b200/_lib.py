"""ctypes loader for libfrb.so, the C-ABI library of include/frb.h.

There is no Python or CPU implementation behind this package: if the CUDA library is missing or no B200 is visible,
every render raises.  The library is built in-tree by `__graft_entry__.build()` (nvcc, sm_100a).
"""
from __future__ import annotations

import ctypes
import os

from .calc import FrbConfig

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FRB_LIB_PATH") or os.path.join(_HERE, "libfrb.so")  # FRB_LIB_PATH: the tuning / trace builds
TUNE_LIB_PATH = os.path.join(_HERE, "libfrb_tune.so")  # -DFRB_TUNING build: frb_set_tuning works (sweeps, two tests)

FRB_RGB24 = 3
FRB_RGBA32 = 4

FRB_OK = 0
FRB_ERR_NO_DEVICE = -1
FRB_ERR_INVALID = -2
FRB_ERR_CUDA = -3
FRB_ERR_TOO_LARGE = -4

KERNEL_AUTO = 0
KERNEL_STREAM7 = 2
KERNEL_SIMPLE = 3
KERNEL_DENSE6 = 4
KERNEL_DEFER6 = 5
KERNEL_TILE = 6
KERNEL_NAMES = {0: "auto", 2: "stream7", 3: "simple", 4: "dense6", 5: "defer6", 6: "tile"}

# every symbol include/frb.h declares; tests/test_abi.py checks the library exports exactly these
EXPORTS = (
    "frb_abi_version",
    "frb_config_default",
    "frb_rgb_new",
    "frb_shard_rows",
    "frb_shard_row_to_y",
    "frb_device_count",
    "frb_init",
    "frb_init_list",
    "frb_shutdown",
    "frb_set_kernel_variant",
    "frb_set_tuning",
    "frb_render",
    "frb_render_shard",
    "frb_render_bands",
    "frb_render_counts",
    "frb_render_frames",
    "frb_render_frames_strided",
    "frb_render_fern",
    "frb_fern_default_threads",
    "frb_render_device_async",
    "frb_render_static_async",
    "frb_sync_stats",
    "frb_async_stats",
    "frb_last_error",
    "frb_last_stats",
    "frb_measure_fp64_peak",
    "frb_host_alloc",
    "frb_host_free",
    "frb_host_register",
    "frb_host_unregister",
    "frb_server_create",
    "frb_server_destroy",
    "frb_server_request",
    "frb_server_poll",
    "frb_server_acquire",
    "frb_server_stats",
    "frb_server_last_error",
    "frb_shared_frame_open",
    "frb_shared_frame_data",
    "frb_shared_frame_pinned",
    "frb_shared_frame_barrier",
    "frb_shared_frame_close",
)


class FrbShard(ctypes.Structure):
    _fields_ = [("index", ctypes.c_uint32), ("count", ctypes.c_uint32), ("band_rows", ctypes.c_uint32),
                ("reserved", ctypes.c_uint32)]


class FrbStats(ctypes.Structure):
    _fields_ = [
        ("pixels", ctypes.c_uint64),
        ("iterations", ctypes.c_uint64),
        ("lane_iterations", ctypes.c_uint64),
        ("iterate_ms", ctypes.c_float),
        ("shade_ms", ctypes.c_float),
        ("device_ms", ctypes.c_float),
        ("total_ms", ctypes.c_float),
        ("kernel_launches", ctypes.c_uint32),
        ("devices", ctypes.c_uint32),
        ("kernel_variant", ctypes.c_uint32),
        ("reserved", ctypes.c_uint32),
    ]

    def as_dict(self) -> dict:
        return {k: getattr(self, k) for k, _ in self._fields_ if k != "reserved"}


class FrbFp64Peak(ctypes.Structure):
    _fields_ = [
        ("dfma_per_s", ctypes.c_double),
        ("dmul_per_s", ctypes.c_double),
        ("dadd_per_s", ctypes.c_double),
        ("mixed6_per_s", ctypes.c_double),
        ("sm_clock_mhz", ctypes.c_double),
        ("sm_count", ctypes.c_uint32),
        ("reserved", ctypes.c_uint32),
    ]


class FrbServerStats(ctypes.Structure):
    _fields_ = [("requests", ctypes.c_uint64), ("coalesced", ctypes.c_uint64), ("frames_rendered", ctypes.c_uint64),
                ("working", ctypes.c_uint32), ("try_redraw", ctypes.c_uint32)]


SERVER_RENDER_FN = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.POINTER(FrbConfig), ctypes.POINTER(ctypes.c_uint8), ctypes.c_size_t,
                                    ctypes.c_void_p)


class FrbError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"frb error {code}: {message}")
        self.code = code


_lib = None


def load() -> ctypes.CDLL:
    """Loads libfrb.so; raises (never falls back) if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FrbError(FRB_ERR_NO_DEVICE, f"{LIB_PATH} not built — run `python -c 'import __graft_entry__ as g; g.build()'` "
                                          "(nvcc, sm_100a). There is no CPU fallback.")
    L = ctypes.CDLL(LIB_PATH)
    P = ctypes.POINTER
    L.frb_abi_version.restype = ctypes.c_uint32
    L.frb_abi_version.argtypes = []
    L.frb_render_fern.restype = ctypes.c_int
    L.frb_render_fern.argtypes = [P(FrbConfig), ctypes.c_uint32, ctypes.c_uint64, P(FrbShard), ctypes.c_void_p, ctypes.c_size_t,
                                  ctypes.c_uint32, ctypes.c_int]
    L.frb_fern_default_threads.restype = ctypes.c_uint32
    L.frb_fern_default_threads.argtypes = []
    L.frb_config_default.restype = None
    L.frb_config_default.argtypes = [P(FrbConfig), ctypes.c_uint32]
    L.frb_shard_rows.restype = ctypes.c_uint32
    L.frb_shard_rows.argtypes = [ctypes.c_uint32, P(FrbShard)]
    L.frb_shard_row_to_y.restype = ctypes.c_uint32
    L.frb_shard_row_to_y.argtypes = [ctypes.c_uint32, P(FrbShard), ctypes.c_uint32]
    L.frb_device_count.restype = ctypes.c_int
    L.frb_device_count.argtypes = []
    L.frb_init.restype = ctypes.c_int
    L.frb_init.argtypes = [ctypes.c_uint32]
    L.frb_init_list.restype = ctypes.c_int
    L.frb_init_list.argtypes = [P(ctypes.c_int32), ctypes.c_uint32]
    L.frb_set_tuning.restype = ctypes.c_int
    L.frb_set_tuning.argtypes = [ctypes.c_char_p, ctypes.c_int64]
    L.frb_render_bands.restype = ctypes.c_int
    L.frb_render_bands.argtypes = [P(FrbConfig), P(FrbShard), ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_int]
    L.frb_shared_frame_open.restype = ctypes.c_void_p
    L.frb_shared_frame_open.argtypes = [ctypes.c_char_p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_uint32]
    L.frb_shared_frame_data.restype = ctypes.c_void_p
    L.frb_shared_frame_data.argtypes = [ctypes.c_void_p]
    L.frb_shared_frame_pinned.restype = ctypes.c_int
    L.frb_shared_frame_pinned.argtypes = [ctypes.c_void_p]
    L.frb_shared_frame_barrier.restype = ctypes.c_int
    L.frb_shared_frame_barrier.argtypes = [ctypes.c_void_p]
    L.frb_shared_frame_close.restype = None
    L.frb_shared_frame_close.argtypes = [ctypes.c_void_p]
    L.frb_shutdown.restype = None
    L.frb_shutdown.argtypes = []
    L.frb_set_kernel_variant.restype = ctypes.c_int
    L.frb_set_kernel_variant.argtypes = [ctypes.c_uint32]
    L.frb_render.restype = ctypes.c_int
    L.frb_render.argtypes = [P(FrbConfig), ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_int]
    L.frb_render_shard.restype = ctypes.c_int
    L.frb_render_shard.argtypes = [P(FrbConfig), P(FrbShard), ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_int]
    L.frb_render_counts.restype = ctypes.c_int
    L.frb_render_counts.argtypes = [P(FrbConfig), P(FrbShard), ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    L.frb_render_frames.restype = ctypes.c_int
    L.frb_render_frames.argtypes = [P(FrbConfig), ctypes.c_uint32, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_int]
    L.frb_render_frames_strided.restype = ctypes.c_int
    L.frb_render_frames_strided.argtypes = [P(FrbConfig), ctypes.c_uint32, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_int,
                                            ctypes.c_uint32, ctypes.c_uint32]
    L.frb_render_device_async.restype = ctypes.c_int
    L.frb_render_device_async.argtypes = [P(FrbConfig), P(FrbShard), ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32,
                                          ctypes.c_int, ctypes.c_void_p]
    L.frb_render_static_async.restype = ctypes.c_int
    L.frb_render_static_async.argtypes = [P(FrbConfig), ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_int,
                                          ctypes.c_void_p]
    L.frb_sync_stats.restype = ctypes.c_int
    L.frb_sync_stats.argtypes = [P(FrbStats)]
    L.frb_async_stats.restype = ctypes.c_int
    L.frb_async_stats.argtypes = [ctypes.c_uint32, P(FrbStats)]
    L.frb_last_error.restype = ctypes.c_char_p
    L.frb_last_error.argtypes = []
    L.frb_last_stats.restype = ctypes.c_int
    L.frb_last_stats.argtypes = [P(FrbStats)]
    L.frb_host_alloc.restype = ctypes.c_void_p
    L.frb_host_alloc.argtypes = [ctypes.c_size_t]
    L.frb_host_free.restype = None
    L.frb_host_free.argtypes = [ctypes.c_void_p]
    L.frb_host_register.restype = ctypes.c_int
    L.frb_host_register.argtypes = [ctypes.c_void_p, ctypes.c_size_t]
    L.frb_host_unregister.restype = ctypes.c_int
    L.frb_host_unregister.argtypes = [ctypes.c_void_p]
    L.frb_measure_fp64_peak.restype = ctypes.c_int
    L.frb_measure_fp64_peak.argtypes = [ctypes.c_int, P(FrbFp64Peak)]
    L.frb_server_create.restype = ctypes.c_void_p
    L.frb_server_create.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
    L.frb_server_destroy.restype = None
    L.frb_server_destroy.argtypes = [ctypes.c_void_p]
    L.frb_server_request.restype = ctypes.c_int
    L.frb_server_request.argtypes = [ctypes.c_void_p, P(FrbConfig)]
    L.frb_server_acquire.restype = ctypes.c_int
    L.frb_server_acquire.argtypes = [ctypes.c_void_p, P(ctypes.c_void_p), P(ctypes.c_uint32), P(ctypes.c_uint32), ctypes.c_int]
    L.frb_server_poll.restype = ctypes.c_int
    L.frb_server_poll.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, P(ctypes.c_uint32), P(ctypes.c_uint32),
                                  ctypes.c_int]
    L.frb_server_stats.restype = ctypes.c_int
    L.frb_server_stats.argtypes = [ctypes.c_void_p, P(FrbServerStats)]
    L.frb_server_last_error.restype = ctypes.c_char_p
    L.frb_server_last_error.argtypes = [ctypes.c_void_p]
    if L.frb_abi_version() != 1:
        raise FrbError(FRB_ERR_INVALID, f"ABI version mismatch: library {L.frb_abi_version()}, binding 1")
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != FRB_OK:
        raise FrbError(rc, load().frb_last_error().decode("utf-8", "replace"))
