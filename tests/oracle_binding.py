"""ctypes binding of the CPU oracle (oracle/calc_oracle.c).  Test infrastructure: imported only by tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs — never by the product package."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

from fractal_renderer_b200.calc import Config, FrbConfig, FrbRGB

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(_ROOT, "oracle", "libcalc_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    if force or not os.path.exists(_SO):
        subprocess.check_call(["make", "-C", os.path.join(_ROOT, "oracle"), "-s"])
    return _SO


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_SO)
        L.oracle_recursive.restype = ctypes.c_uint32
        L.oracle_recursive.argtypes = [ctypes.c_uint32] + [ctypes.c_double] * 5 + [ctypes.POINTER(ctypes.c_double)] * 2
        L.oracle_xy_to_imaginary.restype = None
        L.oracle_xy_to_imaginary.argtypes = [ctypes.POINTER(FrbConfig), ctypes.c_uint32, ctypes.c_uint32,
                                             ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
        L.oracle_shade.restype = FrbRGB
        L.oracle_shade.argtypes = [ctypes.POINTER(FrbConfig), ctypes.c_double, ctypes.c_double, ctypes.c_uint32]
        L.oracle_get_recursive_pixel.restype = FrbRGB
        L.oracle_get_recursive_pixel.argtypes = [ctypes.POINTER(FrbConfig), ctypes.c_uint32, ctypes.c_uint32]
        L.oracle_render_rows.restype = ctypes.c_uint64
        L.oracle_render_rows.argtypes = [ctypes.POINTER(FrbConfig), ctypes.c_void_p, ctypes.c_uint32,
                                         ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        L.oracle_get_image.restype = ctypes.c_uint64
        L.oracle_get_image.argtypes = [ctypes.POINTER(FrbConfig), ctypes.c_void_p, ctypes.c_int]
        L.oracle_config_new.restype = None
        L.oracle_config_new.argtypes = [ctypes.POINTER(FrbConfig), ctypes.c_uint32]
        L.oracle_max_threads.restype = ctypes.c_int
        _lib = L
    return _lib


def recursive(iterations: int, start: complex, c: complex, limit: float):
    """calc::recursive -> ((re, im), count)."""
    re, im = ctypes.c_double(), ctypes.c_double()
    n = lib().oracle_recursive(iterations, start.real, start.imag, c.real, c.imag, limit, ctypes.byref(re), ctypes.byref(im))
    return (re.value, im.value), n


def get_recursive_pixel(cfg: Config, x: int, y: int):
    c = cfg.to_c()
    v = lib().oracle_get_recursive_pixel(ctypes.byref(c), x, y)
    return (v.r, v.g, v.b)


def render_rows(cfg: Config, rows=None, want_rgb=True, want_iters=True, want_z=True, threads: int = 0):
    """Returns dict(rgb=(n,W,3) u8, iters=(n,W) u32, z=(n,W,2) f64, total=int) for image rows `rows`
    (None = all rows)."""
    c = cfg.to_c()
    if rows is None:
        rows_arr = None
        n = cfg.height
    else:
        rows_arr = np.ascontiguousarray(rows, dtype=np.uint32)
        n = int(rows_arr.size)
    W = cfg.width
    rgb = np.empty((n, W, 3), np.uint8) if want_rgb else None
    it = np.empty((n, W), np.uint32) if want_iters else None
    z = np.empty((n, W, 2), np.float64) if want_z else None
    p = lambda a: a.ctypes.data if a is not None else None
    total = lib().oracle_render_rows(ctypes.byref(c), p(rows_arr), n, p(rgb), p(it), p(z), threads)
    return dict(rgb=rgb, iters=it, z=z, total=int(total))


def get_image(cfg: Config, threads: int = 0) -> np.ndarray:
    """get_image(&Config) -> (H*W, 3) u8, the memory image of Vec<RGB>."""
    c = cfg.to_c()
    out = np.empty((cfg.height * cfg.width, 3), np.uint8)
    lib().oracle_get_image(ctypes.byref(c), out.ctypes.data, threads)
    return out


def max_threads() -> int:
    return int(lib().oracle_max_threads())


# ---- Barnsley fern (oracle/fern_oracle.c) --------------------------------------------------------------------------

def _fern_lib():
    L = lib()
    if not getattr(L, "_fern_ready", False):
        P = ctypes.POINTER
        L.fern_oracle_chains_per_thread.restype = ctypes.c_uint32
        L.fern_oracle_chains_per_thread.argtypes = [ctypes.c_uint32]
        for name in ("fern_oracle_counts_chained", "fern_oracle_counts_reference_style"):
            fn = getattr(L, name)
            fn.restype = None
            fn.argtypes = [P(FrbConfig), ctypes.c_uint32, ctypes.c_uint64, ctypes.c_void_p]
        L.fern_oracle_colour_after.restype = FrbRGB
        L.fern_oracle_colour_after.argtypes = [P(FrbConfig), ctypes.c_uint32]
        L.fern_oracle_compose.restype = None
        L.fern_oracle_compose.argtypes = [P(FrbConfig), ctypes.c_uint32, ctypes.c_void_p, ctypes.c_void_p]
        L.fern_oracle_get_image.restype = ctypes.c_uint64
        L.fern_oracle_get_image.argtypes = [P(FrbConfig), ctypes.c_uint32, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_int]
        L._fern_ready = True
    return L


def fern_counts(cfg: Config, threads: int, seed: int, reference_style: bool = False) -> np.ndarray:
    """Hit counts (threads, H, W) uint32 of the chaos game: the CUDA path's chain decomposition, or the reference's own
    structure (one uninterrupted orbit per thread)."""
    c = cfg.to_c()
    counts = np.zeros((threads, cfg.height, cfg.width), np.uint32)
    fn = _fern_lib().fern_oracle_counts_reference_style if reference_style else _fern_lib().fern_oracle_counts_chained
    fn(ctypes.byref(c), threads, seed & 0xFFFFFFFFFFFFFFFF, counts.ctypes.data)
    return counts


def fern_colour_after(cfg: Config, n: int):
    c = cfg.to_c()
    v = _fern_lib().fern_oracle_colour_after(ctypes.byref(c), n)
    return (v.r, v.g, v.b)


def fern_compose(cfg: Config, counts: np.ndarray) -> np.ndarray:
    c = cfg.to_c()
    counts = np.ascontiguousarray(counts, np.uint32)
    rgb = np.empty((cfg.height, cfg.width, 3), np.uint8)
    _fern_lib().fern_oracle_compose(ctypes.byref(c), counts.shape[0], counts.ctypes.data, rgb.ctypes.data)
    return rgb


def fern_get_image(cfg: Config, threads: int, seed: int, reference_style: bool = False) -> np.ndarray:
    c = cfg.to_c()
    rgb = np.empty((cfg.height, cfg.width, 3), np.uint8)
    _fern_lib().fern_oracle_get_image(ctypes.byref(c), threads, seed & 0xFFFFFFFFFFFFFFFF, rgb.ctypes.data, int(reference_style))
    return rgb
