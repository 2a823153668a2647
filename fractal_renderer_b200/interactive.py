"""Headless mirror of the reference GUI's interactive loop (src/gui.rs) — SURVEY.md §8(f) rank 1 and 2.

No window, no egui, no GL (no graphics stack on these machines): what remains is
  * `ViewState`   — the pan / zoom / algo-switch / 2x-screenshot state updates of `App::update` (src/gui.rs:281-341),
                    arithmetic reproduced as written (including its quirks, see the method docs);
  * `FrameServer` — the render thread with latest-wins request coalescing (src/gui.rs:37-48, 56-82, 110-117), a native
                    C++ thread inside libfrb.so (csrc/frb_server.cu) that produces RGBA8 frames with frb_render.
"""
from __future__ import annotations

import ctypes
import math
from typing import Callable, Optional

import numpy as np

from . import _lib
from .calc import Algo, Config, FrbConfig, Imaginary


def _f32(v: float) -> float:
    return float(np.float32(v))


class ViewState:
    """`App.state.config` plus the input handling of src/gui.rs:281-341."""

    def __init__(self, config: Optional[Config] = None):
        self.config = config if config is not None else Config.new(Algo.Mandelbrot)

    # -- movement, src/gui.rs:287-300 ----------------------------------------------------------------------
    def apply_keys(self, dt: float, left=False, right=False, up=False, down=False) -> None:
        """Arrow keys held for a frame of `dt` seconds.  As written in the reference: Left and Up use 1/scale.re,
        Right and Down use 1/scale.im (identical when the scale is isotropic, which the CLI enforces by default)."""
        c = self.config
        scale_x = 1.0 / c.scale.re
        scale_y = 1.0 / c.scale.im
        re, im = c.pos.re, c.pos.im
        if left:
            re -= scale_x * dt * 0.5
        if right:
            re += scale_y * dt * 0.5
        if up:
            im -= scale_x * dt * 0.5
        if down:
            im += scale_y * dt * 0.5
        self.config = c.with_(pos=Imaginary(re, im))

    # -- zoom, src/gui.rs:302-317 ---------------------------------------------------------------------------
    def apply_scroll(self, delta_y: float) -> None:
        """Scroll wheel: |delta| < 1 is ignored; zooming out shrinks by min(log10(-d/10 + 1)/2, 1) computed in f32,
        zooming in grows by 1 + d/80."""
        delta = _f32(delta_y)
        if not (delta >= 1.0 or delta <= -1.0):
            return
        if delta < 0.0:
            d = _f32(-delta)
            scale_diff = min(_f32(_f32(math.log10(_f32(_f32(d / np.float32(10.0)) + np.float32(1.0)))) / np.float32(2.0)), 1.0)
            factor = 1.0 - float(scale_diff)
        else:
            factor = 1.0 + (float(delta) / 80.0)
        self.config = self.config.with_(scale=self.config.scale * factor)

    # -- "Apply changes", src/gui.rs:331-341 ----------------------------------------------------------------
    def set_algo(self, algo: Algo) -> None:
        """Changing the algorithm resets every parameter to Config::new(algo)."""
        if algo != self.config.algo:
            self.config = Config.new(algo)

    def set_size(self, width: int, height: int) -> None:
        """The size widgets clamp to >= 16 (src/gui.rs:168-174)."""
        self.config = self.config.with_(width=max(16, int(width)), height=max(16, int(height)))

    # -- screenshot, src/gui.rs:319-328 ---------------------------------------------------------------------
    def screenshot_config(self) -> Config:
        """The `S` key renders the current view at twice the width and height."""
        return self.config.with_(width=self.config.width * 2, height=self.config.height * 2)


class FrameServer:
    """Render thread + latest-wins coalescing.  `request(config)` never blocks; `poll()` returns the newest finished
    RGBA frame (H, W, 4) or None.  `renderer`, if given, replaces the GPU render (used to test the coalescing logic
    without a device): `renderer(config) -> (H, W, 4) uint8 array`."""

    def __init__(self, device: int = 0, renderer: Optional[Callable[[Config], np.ndarray]] = None):
        self._lib = _lib.load()
        self._cb = None
        cb_ptr = None
        if renderer is not None:
            def _trampoline(cfg_p, rgba_p, length, _user):
                try:
                    cfg = Config.from_c(cfg_p.contents)
                    frame = np.ascontiguousarray(renderer(cfg), dtype=np.uint8)
                    if frame.nbytes != length:
                        return _lib.FRB_ERR_INVALID
                    ctypes.memmove(rgba_p, frame.ctypes.data, length)
                    return _lib.FRB_OK
                except Exception:  # never unwind across the C ABI
                    return _lib.FRB_ERR_INVALID

            self._cb = _lib.SERVER_RENDER_FN(_trampoline)
            cb_ptr = ctypes.cast(self._cb, ctypes.c_void_p)
        self._h = self._lib.frb_server_create(device, cb_ptr, None)
        if not self._h:
            raise _lib.FrbError(_lib.FRB_ERR_INVALID, "frb_server_create failed")

    def request(self, config: Config) -> None:
        c = config.to_c()
        _lib.check(self._lib.frb_server_request(self._h, ctypes.byref(c)))

    def poll(self, timeout_ms: int = 0, copy: bool = True) -> Optional[np.ndarray]:
        """The newest finished frame as (H, W, 4) uint8, or None.  copy=False hands out the server's own (pinned) frame without
        copying it (frb_server_acquire): the array is only valid until the next poll that returns a frame, or close()."""
        w, h = ctypes.c_uint32(), ctypes.c_uint32()
        if not copy:
            ptr = ctypes.c_void_p()
            rc = self._lib.frb_server_acquire(self._h, ctypes.byref(ptr), ctypes.byref(w), ctypes.byref(h), timeout_ms)
            if rc < 0:
                raise _lib.FrbError(rc, self._lib.frb_server_last_error(self._h).decode("utf-8", "replace"))
            if rc == 0:
                return None
            n = h.value * w.value * 4
            raw = (ctypes.c_uint8 * max(n, 1)).from_address(ptr.value)
            return np.frombuffer(raw, dtype=np.uint8, count=n).reshape(h.value, w.value, 4)
        for _ in range(8):
            # peek (waits, reports the size), then take; a newer frame of another size may slip in between: retry
            rc = self._lib.frb_server_poll(self._h, None, 0, ctypes.byref(w), ctypes.byref(h), timeout_ms)
            if rc < 0:
                raise _lib.FrbError(rc, self._lib.frb_server_last_error(self._h).decode("utf-8", "replace"))
            if rc == 0:
                return None
            out = np.empty((h.value, w.value, 4), np.uint8)
            rc = self._lib.frb_server_poll(self._h, out.ctypes.data, out.nbytes, ctypes.byref(w), ctypes.byref(h), 0)
            if rc == 1 and (h.value, w.value) == out.shape[:2]:
                return out
            if rc == 1:
                return out.reshape(-1)[: h.value * w.value * 4].reshape(h.value, w.value, 4)
        return None

    def stats(self) -> dict:
        s = _lib.FrbServerStats()
        _lib.check(self._lib.frb_server_stats(self._h, ctypes.byref(s)))
        return {k: int(getattr(s, k)) for k, _ in s._fields_}

    def close(self) -> None:
        if self._h:
            self._lib.frb_server_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
