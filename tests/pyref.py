"""A second, independent restatement of the path in pure Python floats (IEEE binary64, one rounding per operator,
CPython never fuses).  Written from SURVEY.md Appendix A, not from oracle/calc_oracle.c; used on tiny images to
cross-check the C oracle (tests/test_oracle.py).  Test infrastructure only."""
from __future__ import annotations

import math
import struct

from fractal_renderer_b200.calc import Algo, Config


def same_bits(a: float, b: float) -> bool:
    if math.isnan(a) and math.isnan(b):
        return True  # NaN sign/payload differs between x86 and CUDA; the class is what the reference defines
    return struct.pack("<d", a) == struct.pack("<d", float(b))


def _div(a: float, b: float) -> float:
    try:
        return a / b
    except ZeroDivisionError:
        if a != a or a == 0.0:
            return math.nan
        neg = (math.copysign(1.0, a) < 0) != (math.copysign(1.0, b) < 0)
        return -math.inf if neg else math.inf


def _mul(a: float, b: float) -> float:
    try:
        return a * b
    except OverflowError:  # pragma: no cover (Python floats do not raise on multiply)
        return math.inf


def _sat_u8(v: float) -> int:
    if v != v:
        return 0
    if v <= 0.0:
        return 0
    if v >= 255.0:
        return 255
    return int(v)


def _log2(v: float) -> float:
    if v != v:
        return math.nan
    if v == 0.0:
        return -math.inf
    if v < 0.0:
        return math.nan
    if v == math.inf:
        return math.inf
    return math.log2(v)


def recursive(iterations: int, start, c, limit: float):
    squared = limit * limit
    pre, pim = start
    cre, cim = c
    for i in range(iterations):
        nre = (pre * pre) - (pim * pim)
        nim = (2.0 * pre) * pim
        nre = nre + cre
        nim = nim + cim
        dist = nre * nre + nim * nim
        if dist > squared:
            return (nre, nim), i
        pre, pim = nre, nim
    return (pre, pim), iterations


def get_recursive_pixel(cfg: Config, x: int, y: int):
    w, h = float(cfg.width), float(cfg.height)
    re = _div(_div(float(x), h) - _div(w, h) / 2.0, cfg.scale.re) + cfg.pos.re
    im = _div(_div(float(y), h) - 0.5, cfg.scale.im) + cfg.pos.im
    if cfg.algo == Algo.Mandelbrot:
        (zr, zi), n = recursive(cfg.iterations, (re, im), (re, im), cfg.limit)
    elif cfg.algo == Algo.Julia:
        (zr, zi), n = recursive(cfg.iterations, (re, im), (cfg.julia_set.re, cfg.julia_set.im), cfg.limit)
    else:
        return (0.0, 0.0), 0, (0, 0, 0)
    dist = zr * zr + zi * zi
    P, Q = cfg.primary_color, cfg.secondary_color
    if dist > cfg.stable_limit:
        it = float(n)
        if cfg.smooth:
            log_zn = _log2(math.sqrt(dist)) / 2.0
            nu = _log2(log_zn)
            it = it + (1.0 - nu)
        m = _div(it, float(cfg.iterations)) * cfg.exposure
        rgb = (_sat_u8(P.r * m), _sat_u8(P.b * m), _sat_u8(P.g * m))
    elif cfg.inside:
        rgb = (_sat_u8(Q.r * dist), _sat_u8(Q.b * dist), _sat_u8(Q.g * dist))
    else:
        rgb = (0, 0, 0)
    return (zr, zi), n, rgb
