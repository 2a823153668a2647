"""Host-side mirror of the reference's `calc` value types (calc/src/lib.rs:19-179).

Same names, field order and defaults as the Rust crate so that code and tests written against the reference read the
same here: `Config`, `Imaginary`, `RGB`, `Algo`.  These are plain host objects; `Config.to_c()` produces the
`FrbConfig` that crosses the C ABI (include/frb.h).
"""
from __future__ import annotations

import ctypes
import enum
from dataclasses import dataclass, field, replace


class Algo(enum.IntEnum):
    """calc::Algo (calc/src/lib.rs:150-154); values are the C ABI's FRB_ALGO_*."""

    Mandelbrot = 0
    BarnsleyFern = 1
    Julia = 2

    @classmethod
    def from_str(cls, s: str) -> "Algo":
        """impl FromStr for Algo (calc/src/lib.rs:166-179): case-insensitive names, `fern` alias."""
        t = s.lower()
        if t == "mandelbrot":
            return cls.Mandelbrot
        if t in ("fern", "barnsleyfern"):
            return cls.BarnsleyFern
        if t == "julia":
            return cls.Julia
        raise ValueError("invalid algorithm name")


@dataclass(frozen=True)
class Imaginary:
    """calc::Imaginary (calc/src/lib.rs:77-117)."""

    re: float = 0.0
    im: float = 0.0

    def __mul__(self, rhs: float) -> "Imaginary":
        return Imaginary(self.re * rhs, self.im * rhs)


Imaginary.ZERO = Imaginary(0.0, 0.0)  # type: ignore[attr-defined]
Imaginary.ONE = Imaginary(1.0, 1.0)  # type: ignore[attr-defined]


@dataclass(frozen=True)
class RGB:
    """calc::RGB (calc/src/lib.rs:119-131): fields r,g,b as STORED.

    `RGB.new(r, b, g)` reproduces the reference constructor's argument order (second argument lands in `b`).
    """

    r: int = 0
    g: int = 0
    b: int = 0

    @staticmethod
    def new(r: int, b: int, g: int) -> "RGB":
        return RGB(r=r, g=g, b=b)

    @staticmethod
    def from_hex(s: str) -> "RGB":
        """CLI hex parsing (src/lib.rs:22-29): RGB::new(bytes[0], bytes[1], bytes[2])."""
        v = bytes.fromhex(s.lstrip("#"))
        if len(v) != 3:
            raise ValueError("colour must be 3 hex bytes")
        return RGB.new(v[0], v[1], v[2])


RGB.BLACK = RGB(0, 0, 0)  # type: ignore[attr-defined]


class FrbImaginary(ctypes.Structure):
    _fields_ = [("re", ctypes.c_double), ("im", ctypes.c_double)]


class FrbRGB(ctypes.Structure):
    _fields_ = [("r", ctypes.c_uint8), ("g", ctypes.c_uint8), ("b", ctypes.c_uint8)]


class FrbConfig(ctypes.Structure):
    """include/frb.h FrbConfig (104 bytes)."""

    _fields_ = [
        ("algo", ctypes.c_uint32),
        ("width", ctypes.c_uint32),
        ("height", ctypes.c_uint32),
        ("iterations", ctypes.c_uint32),
        ("limit", ctypes.c_double),
        ("stable_limit", ctypes.c_double),
        ("pos", FrbImaginary),
        ("scale", FrbImaginary),
        ("exposure", ctypes.c_double),
        ("inside", ctypes.c_uint8),
        ("smooth", ctypes.c_uint8),
        ("primary_color", FrbRGB),
        ("secondary_color", FrbRGB),
        ("color_weight", ctypes.c_double),
        ("julia_set", FrbImaginary),
    ]


assert ctypes.sizeof(FrbConfig) == 104


@dataclass(frozen=True)
class Config:
    """calc::Config (calc/src/lib.rs:21-37) with Config::new's defaults (calc/src/lib.rs:39-69)."""

    algo: Algo = Algo.Mandelbrot
    width: int = 2000
    height: int = 1000
    iterations: int = 50
    limit: float = 65536.0
    stable_limit: float = 2.0
    pos: Imaginary = Imaginary(0.0, 0.0)
    scale: Imaginary = Imaginary(0.4, 0.4)
    exposure: float = 2.0
    inside: bool = True
    smooth: bool = True
    primary_color: RGB = field(default_factory=lambda: RGB.new(40, 40, 255))
    secondary_color: RGB = field(default_factory=lambda: RGB.new(240, 170, 0))
    color_weight: float = 0.01
    julia_set: Imaginary = Imaginary(0.0, 0.0)

    @staticmethod
    def new(algo: Algo = Algo.Mandelbrot) -> "Config":
        """Config::new(algo) (calc/src/lib.rs:39-69), including the fern's different defaults."""
        if algo == Algo.BarnsleyFern:
            return Config(
                algo=algo,
                iterations=10_000_000,
                primary_color=RGB.new(4, 100, 3),
                secondary_color=RGB.new(240, 240, 240),
            )
        return Config(algo=algo)

    @staticmethod
    def cli(algo: Algo = Algo.Mandelbrot, **kw) -> "Config":
        """The configuration the reference CLI builds with no flags (src/lib.rs:31-234): 750x500, exposure 5, pos.re 0 for
        julia and -0.6 otherwise (`default_value_if("algo", Some("julia"), Some("0"))`, src/lib.rs:66-71); iterations and both
        colours come from Config::new(algo) (src/lib.rs:209-222) — so the fern keeps its 10 M points and green-on-grey."""
        ref = Config.new(algo)
        base = dict(
            algo=algo,
            width=750,
            height=500,
            iterations=ref.iterations,
            exposure=5.0,
            pos=Imaginary(0.0 if algo == Algo.Julia else -0.6, 0.0),
            primary_color=ref.primary_color,
            secondary_color=ref.secondary_color,
        )
        base.update(kw)
        return Config(**base)

    def with_(self, **kw) -> "Config":
        return replace(self, **kw)

    def to_c(self) -> FrbConfig:
        c = FrbConfig()
        c.algo = int(self.algo)
        c.width = self.width
        c.height = self.height
        c.iterations = self.iterations
        c.limit = self.limit
        c.stable_limit = self.stable_limit
        c.pos = FrbImaginary(self.pos.re, self.pos.im)
        c.scale = FrbImaginary(self.scale.re, self.scale.im)
        c.exposure = self.exposure
        c.inside = 1 if self.inside else 0
        c.smooth = 1 if self.smooth else 0
        c.primary_color = FrbRGB(self.primary_color.r, self.primary_color.g, self.primary_color.b)
        c.secondary_color = FrbRGB(self.secondary_color.r, self.secondary_color.g, self.secondary_color.b)
        c.color_weight = self.color_weight
        c.julia_set = FrbImaginary(self.julia_set.re, self.julia_set.im)
        return c

    @staticmethod
    def from_c(c: FrbConfig) -> "Config":
        return Config(
            algo=Algo(c.algo),
            width=c.width,
            height=c.height,
            iterations=c.iterations,
            limit=c.limit,
            stable_limit=c.stable_limit,
            pos=Imaginary(c.pos.re, c.pos.im),
            scale=Imaginary(c.scale.re, c.scale.im),
            exposure=c.exposure,
            inside=bool(c.inside),
            smooth=bool(c.smooth),
            primary_color=RGB(c.primary_color.r, c.primary_color.g, c.primary_color.b),
            secondary_color=RGB(c.secondary_color.r, c.secondary_color.g, c.secondary_color.b),
            color_weight=c.color_weight,
            julia_set=Imaginary(c.julia_set.re, c.julia_set.im),
        )
