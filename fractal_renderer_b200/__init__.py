"""fractal_renderer_b200 — B200-native drop-in for the escape-time path of Icelk/fractal-renderer.

Host-side mirror of the reference's public interface for this path (same names and argument meaning):

    reference (Rust)                                   here
    -------------------------------------------------  ---------------------------------------------
    calc::Config / Imaginary / RGB / Algo              Config / Imaginary / RGB / Algo   (calc.py)
    get_image(&Config) -> Vec<RGB>  (src/lib.rs:253)   get_image(config) -> (H*W, 3) uint8 array
    GUI's RGBA expansion            (src/gui.rs:62-75) get_image_rgba(config) -> (H, W, 4) uint8 array
    calc::recursive per pixel       (calc/src/lib.rs:245) render_counts(config) -> (iters, zfinal)
    get_image's BarnsleyFern arm    (src/lib.rs:271-319) render_fern(config, threads, seed)

All pixels are computed by the CUDA library behind include/frb.h (libfrb.so); nothing here computes a pixel on the
CPU, and nothing here imports oracle/.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Sequence

import numpy as np

from . import _lib
from ._lib import (FRB_RGB24, FRB_RGBA32, KERNEL_AUTO, KERNEL_DEFER6, KERNEL_DENSE6, KERNEL_TILE, KERNEL_NAMES, KERNEL_SIMPLE, KERNEL_STREAM7,
                   FrbError, FrbFp64Peak, FrbShard, FrbStats)
from .calc import RGB, Algo, Config, FrbConfig, Imaginary

__all__ = [
    "Algo", "Config", "Imaginary", "RGB", "FrbError", "Shard", "get_image", "get_image_rgba", "get_image_into",
    "render_shard", "render_counts", "render_frames", "render_device_async", "render_static_async", "sync_stats", "async_stats", "last_stats",
    "device_count", "init", "shutdown", "render_fern", "fern_default_threads", "set_kernel_variant", "measure_fp64_peak", "shard_rows", "shard_row_to_y",
    "init_list", "set_tuning", "render_bands", "SharedFrame",
    "FRB_RGB24", "FRB_RGBA32", "pinned_empty", "KERNEL_AUTO", "KERNEL_STREAM7", "KERNEL_SIMPLE", "KERNEL_DENSE6", "KERNEL_DEFER6", "KERNEL_TILE", "KERNEL_NAMES",
]


class Shard:
    """One device's share of a frame: interleaved bands of `band_rows` rows (include/frb.h FrbShard)."""

    def __init__(self, index: int = 0, count: int = 1, band_rows: int = 0):
        self.index, self.count, self.band_rows = index, count, band_rows

    def to_c(self) -> FrbShard:
        return FrbShard(self.index, self.count, self.band_rows, 0)

    def rows(self, height: int) -> int:
        return shard_rows(height, self)

    def image_rows(self, height: int) -> np.ndarray:
        """Image row of every local row of this shard, in output order."""
        n = self.rows(height)
        br = self.band_rows if self.band_rows else max(height, 1)
        lr = np.arange(n, dtype=np.int64)
        band = lr // br
        return ((band * self.count + self.index) * br + (lr - band * br)).astype(np.uint32)


def _shard_ptr(shard: Optional[Shard]):
    if shard is None:
        return None, None
    s = shard.to_c()
    return s, ctypes.byref(s)


def device_count() -> int:
    return int(_lib.load().frb_device_count())


def init(device_mask: int = 0) -> None:
    _lib.check(_lib.load().frb_init(device_mask))


def init_list(devices: Sequence[int]) -> None:
    """Explicit device list for device = -1 renders: shard i of len(devices) goes to devices[i]; a device may repeat."""
    arr = (ctypes.c_int32 * len(devices))(*devices)
    _lib.check(_lib.load().frb_init_list(arr, len(devices)))


def set_tuning(name: str, value: int) -> None:
    """Tuning build only (FRB_LIB_PATH=.../libfrb_tune.so): the product library rejects every knob."""
    _lib.check(_lib.load().frb_set_tuning(name.encode(), int(value)))


def shutdown() -> None:
    _lib.load().frb_shutdown()


def set_kernel_variant(variant: int) -> None:
    _lib.check(_lib.load().frb_set_kernel_variant(variant))


def shard_rows(height: int, shard: Optional[Shard]) -> int:
    _, p = _shard_ptr(shard)
    return int(_lib.load().frb_shard_rows(height, p))


def shard_row_to_y(height: int, shard: Optional[Shard], local_row: int) -> int:
    _, p = _shard_ptr(shard)
    return int(_lib.load().frb_shard_row_to_y(height, p, local_row))


def last_stats() -> dict:
    s = FrbStats()
    _lib.check(_lib.load().frb_last_stats(ctypes.byref(s)))
    return s.as_dict()


def get_image_into(config: Config, out: np.ndarray, fmt: int = FRB_RGB24, device: int = 0) -> np.ndarray:
    """get_image writing into a caller-owned host buffer (any C-contiguous uint8 array of >= W*H*fmt bytes; a
    pinned buffer avoids the driver's staging copy).  device = -1 spreads the frame over all GPUs."""
    if out.dtype != np.uint8 or not out.flags.c_contiguous:
        raise ValueError("out must be a C-contiguous uint8 array")
    c = config.to_c()
    if config.algo == Algo.BarnsleyFern:
        # get_image's other arm (src/lib.rs:271-319): rayon's thread count, fresh entropy — as the reference
        seed = int.from_bytes(os.urandom(8), "little")
        _lib.check(_lib.load().frb_render_fern(ctypes.byref(c), 0, seed, None, out.ctypes.data, out.nbytes, fmt, max(device, 0)))
        return out
    _lib.check(_lib.load().frb_render(ctypes.byref(c), out.ctypes.data, out.nbytes, fmt, device))
    return out


def pinned_empty(nbytes: int) -> np.ndarray:
    """A uint8 array in page-locked host memory (frb_host_alloc): as the destination of get_image_into / render_shard it is
    written by the GPU while the iterations run instead of being copied afterwards.  Freed when the array is collected."""
    import weakref
    lib = _lib.load()
    ptr = lib.frb_host_alloc(max(int(nbytes), 1))
    if not ptr:
        raise _lib.FrbError(_lib.FRB_ERR_CUDA, lib.frb_last_error().decode("utf-8", "replace"))
    raw = (ctypes.c_uint8 * max(int(nbytes), 1)).from_address(ptr)
    arr = np.frombuffer(raw, dtype=np.uint8, count=int(nbytes))
    weakref.finalize(raw, lib.frb_host_free, ptr)  # `arr` keeps `raw` alive through its base
    return arr


def get_image(config: Config, device: int = 0) -> np.ndarray:
    """`get_image(&Config) -> Vec<RGB>` (src/lib.rs:253): (H*W, 3) uint8, row-major, fields r,g,b.  Like the reference it
    dispatches on the algorithm: Mandelbrot | Julia -> the per-pixel path, BarnsleyFern -> the chaos game with rayon's
    thread count and a fresh entropy seed (src/lib.rs:271-319, 428)."""
    out = np.empty((config.height * config.width, 3), np.uint8)
    return get_image_into(config, out, FRB_RGB24, device)


def fern_default_threads() -> int:
    """rayon::current_num_threads() as the reference would see it on this host (RAYON_NUM_THREADS or the logical CPUs)."""
    return int(_lib.load().frb_fern_default_threads())


def render_fern(config: Config, threads: int = 0, seed: Optional[int] = None, shard: Optional[Shard] = None,
                fmt: int = FRB_RGB24, device: int = 0) -> np.ndarray:
    """get_image's BarnsleyFern arm (src/lib.rs:271-319): (rows, W, fmt) uint8.  `threads` = the reference's rayon thread
    count T (0 = this host's), `seed` replaces SmallRng::from_entropy (None = fresh entropy, as the reference)."""
    if seed is None:
        seed = int.from_bytes(os.urandom(8), "little")
    rows = shard_rows(config.height, shard)
    out = np.empty((rows, config.width, fmt), np.uint8)
    c = config.to_c()
    s, p = _shard_ptr(shard)
    _lib.check(_lib.load().frb_render_fern(ctypes.byref(c), threads, seed & 0xFFFFFFFFFFFFFFFF, p, out.ctypes.data, out.nbytes,
                                           fmt, device))
    return out


def get_image_rgba(config: Config, device: int = 0) -> np.ndarray:
    """The RGBA8 frame the reference GUI builds from get_image (src/gui.rs:62-75): (H, W, 4) uint8, alpha 255."""
    out = np.empty((config.height, config.width, 4), np.uint8)
    return get_image_into(config, out, FRB_RGBA32, device)


def render_shard(config: Config, shard: Shard, fmt: int = FRB_RGB24, device: int = 0,
                 out: Optional[np.ndarray] = None) -> np.ndarray:
    """One shard of the frame, compact: (rows, W, fmt) uint8."""
    rows = shard.rows(config.height)
    if out is None:
        out = np.empty((rows, config.width, fmt), np.uint8)
    c = config.to_c()
    s, p = _shard_ptr(shard)
    _lib.check(_lib.load().frb_render_shard(ctypes.byref(c), p, out.ctypes.data, out.nbytes, fmt, device))
    return out


def render_bands(config: Config, shard: Shard, frame: np.ndarray, fmt: int = FRB_RGB24, device: int = 0) -> np.ndarray:
    """One shard of the frame written into ITS PLACE in the whole image `frame` (>= H*W*fmt bytes, shared by all shards)."""
    if frame.dtype != np.uint8 or not frame.flags.c_contiguous:
        raise ValueError("frame must be a C-contiguous uint8 array")
    c = config.to_c()
    s, p = _shard_ptr(shard)
    _lib.check(_lib.load().frb_render_bands(ctypes.byref(c), p, frame.ctypes.data, frame.nbytes, fmt, device))
    return frame


class SharedFrame:
    """One host image for a one-process-per-GPU job (frb_shared_frame_*): POSIX shared memory mapped — and, with a GPU, pinned —
    by every rank, so that each rank's render_bands lands in the same frame; `barrier()` = the whole frame is on the host."""

    def __init__(self, name: str, nbytes: int, ranks: int, rank: int):
        self._lib = _lib.load()
        self._h = self._lib.frb_shared_frame_open(name.encode(), int(nbytes), ranks, rank)
        if not self._h:
            raise _lib.FrbError(_lib.FRB_ERR_INVALID, self._lib.frb_last_error().decode("utf-8", "replace"))
        ptr = self._lib.frb_shared_frame_data(self._h)
        raw = (ctypes.c_uint8 * max(int(nbytes), 1)).from_address(ptr)
        self.array = np.frombuffer(raw, dtype=np.uint8, count=int(nbytes))
        self.pinned = bool(self._lib.frb_shared_frame_pinned(self._h))

    def barrier(self) -> None:
        _lib.check(self._lib.frb_shared_frame_barrier(self._h))

    def close(self) -> None:
        if self._h:
            self.array = None
            self._lib.frb_shared_frame_close(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def render_counts(config: Config, shard: Optional[Shard] = None, device: int = 0, want_z: bool = True):
    """The raw result of calc::recursive for every pixel: (iters (rows, W) uint32, zfinal (rows, W, 2) float64)."""
    rows = shard_rows(config.height, shard)
    iters = np.empty((rows, config.width), np.uint32)
    z = np.empty((rows, config.width, 2), np.float64) if want_z else None
    c = config.to_c()
    s, p = _shard_ptr(shard)
    _lib.check(_lib.load().frb_render_counts(ctypes.byref(c), p, iters.ctypes.data, z.ctypes.data if want_z else None, device))
    return iters, z


def render_frames(configs: Sequence[Config], fmt: int = FRB_RGB24, device: int = -1,
                  out: Optional[np.ndarray] = None, first: int = 0, stride: int = 1) -> np.ndarray:
    """Independent frames (zoom animation), handed to the devices in order from a dynamic queue.  Returns the frames back to
    back as bytes.  first/stride: render only frames first, first + stride, ... (one rank of a multi-process job), each into
    its place in `out`, which addresses ALL frames."""
    n = len(configs)
    arr = (FrbConfig * max(n, 1))()
    total = 0
    for k, cfg in enumerate(configs):
        arr[k] = cfg.to_c()
        total += cfg.width * cfg.height * fmt
    if out is None:
        out = np.empty(total, np.uint8)
    _lib.check(_lib.load().frb_render_frames_strided(arr, n, out.ctypes.data, out.nbytes, fmt, device, first, stride))
    return out


def render_device_async(config: Config, dev_ptr: int, nbytes: int, fmt: int = FRB_RGB24, device: int = 0,
                        shard: Optional[Shard] = None, stream: int = 0) -> None:
    """Enqueue a render into DEVICE memory (`dev_ptr`, 16-byte aligned) on `stream` (cudaStream_t handle, 0 = the
    library's stream).  No host copies; complete when the stream drains."""
    c = config.to_c()
    s, p = _shard_ptr(shard)
    _lib.check(_lib.load().frb_render_device_async(ctypes.byref(c), p, ctypes.c_void_p(dev_ptr), nbytes, fmt, device,
                                                   ctypes.c_void_p(stream) if stream else None))


def render_static_async(config: Config, dev_ptr: int, nbytes: int, fmt: int = FRB_RGB24, device: int = 0, stream: int = 0) -> None:
    """Enqueue a whole-frame render into DEVICE memory through frb_render_static_async, the entry point that allocates
    nothing: for hosts that cannot call cudaMalloc when they need a frame (all of HBM already owned by the host's pool,
    no device-wide stall allowed mid-loop, allocator not re-entrant where we are called).  Scratch is a fixed static
    device array, hence frames of at most 1 048 576 pixels with width + height <= 16 384 (FrbError FRB_ERR_TOO_LARGE
    beyond), and calls that overlap on the GPU race: keep them on one stream.  `stream` 0 = the legacy default stream."""
    c = config.to_c()
    _lib.check(_lib.load().frb_render_static_async(ctypes.byref(c), ctypes.c_void_p(dev_ptr), nbytes, fmt, device,
                                                   ctypes.c_void_p(stream) if stream else None))


def sync_stats() -> dict:
    s = FrbStats()
    _lib.check(_lib.load().frb_sync_stats(ctypes.byref(s)))
    return s.as_dict()


def async_stats(back: int = 0) -> dict:
    """Statistics of the asynchronous render issued `back` calls before the last one by this thread."""
    s = FrbStats()
    _lib.check(_lib.load().frb_async_stats(back, ctypes.byref(s)))
    return s.as_dict()


def measure_fp64_peak(device: int = 0) -> dict:
    p = FrbFp64Peak()
    _lib.check(_lib.load().frb_measure_fp64_peak(device, ctypes.byref(p)))
    return {k: getattr(p, k) for k, _ in p._fields_ if k != "reserved"}
