"""The five workloads of BASELINE.json as concrete, frozen inputs.

BASELINE.json names the configs in prose; the exact numbers below are the survey's reading of them and are the ones
every parity test and bench line uses.  All: limit 65536, stable_limit 2, reference default colours, smooth on.
"""
from __future__ import annotations

from .calc import Algo, Config, Imaginary

# C4/C5 centre: end point of the Wikipedia zoom sequence that the reference's examples.md:20-29 walks.
DEEP_CENTRE = Imaginary(-0.743643887037158704752191506114774, 0.131825904205311970493132056385139)


def c1() -> Config:
    """C1: Mandelbrot full view 1920x1080, max_iter 256 (CLI: `1920 1080 -i 256`). 85 781 804 iterations."""
    return Config.cli(Algo.Mandelbrot, width=1920, height=1080, iterations=256)


def c2(scale: float = 500000.0) -> Config:
    """C2: the README screenshot view at 3000x3000 (north-star frame). 12 911 865 777 iterations.
    CLI: `3000 3000 -s 500000 -x -0.7436447860 -y 0.1318252536 -i 4000 -e 2 -d`."""
    return Config.cli(
        Algo.Mandelbrot,
        width=3000,
        height=3000,
        iterations=4000,
        pos=Imaginary(-0.7436447860, 0.1318252536),
        scale=Imaginary(scale, scale),
        exposure=2.0,
        inside=False,
    )


def c2b() -> Config:
    """C2b: C2 with the literal `-s 1000000`. 18 662 294 162 iterations."""
    return c2(1000000.0)


def c3() -> Config:
    """C3: Julia 4096x4096, max_iter 1000, c = (-0.7269, 0.1889) (examples.md:8). 1 341 751 717 iterations."""
    return Config.cli(
        Algo.Julia, width=4096, height=4096, iterations=1000, julia_set=Imaginary(-0.7269, 0.1889)
    )


def c4(width: int = 16384, height: int = 16384) -> Config:
    """C4: Mandelbrot 16384x16384 at 1e12x, max_iter 10000. 574 496 116 940 iterations."""
    return Config.cli(
        Algo.Mandelbrot,
        width=width,
        height=height,
        iterations=10000,
        pos=DEEP_CENTRE,
        scale=Imaginary(1e12, 1e12),
    )


C5_FRAMES = 240


def c5_scale(k: int, frames: int = C5_FRAMES) -> float:
    """scale_k = 10^(12 k / (frames-1)); computed ONCE here with Python's pow and fed as the same bits to the
    oracle and the GPU (pow is not correctly rounded across libms)."""
    return float(10.0 ** (12.0 * k / (frames - 1)))


def c5_frame(k: int, frames: int = C5_FRAMES) -> Config:
    """C5 frame k of the 240-frame zoom 1e0 -> 1e12 at 1920x1080, max_iter 10000."""
    s = c5_scale(k, frames)
    return Config.cli(
        Algo.Mandelbrot, width=1920, height=1080, iterations=10000, pos=DEEP_CENTRE, scale=Imaginary(s, s)
    )


def c5(frames: int = C5_FRAMES) -> list[Config]:
    return [c5_frame(k, frames) for k in range(frames)]


def examples_md() -> dict[str, Config]:
    """The 14 Julia / Mandelbrot command lines of the reference's examples.md:7-29 as Configs — the only inputs the reference
    documents.  CLI semantics (src/lib.rs:31-234): size 750x500 unless given, exposure 5, pos.re -0.6 (0 for julia), `-d` =
    inside off, iterations from Config::new (50) unless given."""
    j = lambda re, im, **kw: Config.cli(Algo.Julia, julia_set=Imaginary(re, im), **kw)
    m = lambda **kw: Config.cli(Algo.Mandelbrot, **kw)
    z = lambda s, x, y, **kw: m(scale=Imaginary(s, s), pos=Imaginary(x, y), **kw)
    return {
        "julia_1": j(-0.8, 0.156, iterations=2000, scale=Imaginary(0.6, 0.6), exposure=30.0, width=2000, height=1000),
        "julia_2": j(-0.7269, 0.1889, iterations=1000, width=3000, height=1500),
        "julia_3": j(-0.70176, 0.3842, iterations=400, exposure=25.0, width=3000, height=1500),
        "julia_4": j(0.285, 0.01, iterations=100, exposure=10.0, width=2500, height=3000),
        "julia_5": j(-0.2256, 0.65, iterations=500, exposure=12.0, pos=Imaginary(0.29449, -0.40460), width=2000, height=1000),
        "julia_6": j(0.36105, 0.35977, exposure=6.0, iterations=500),
        "mandel_classic": m(inside=False, width=3000, height=2000),
        "mandel_golden": m(),
        "mandel_golden_fringe": m(iterations=400),
        "zoom_1": z(400.0, -0.7435669, 0.1314023, iterations=5000, exposure=10.0),
        "zoom_2": z(2000.0, -0.74364990, 0.13188204, iterations=800),
        "zoom_3": z(12000.0, -0.74364085, 0.13182733, iterations=5000, exposure=1.0),
        "zoom_4": z(100000.0, -.743643135, .131825963, iterations=2000, inside=False, exposure=3.0),
        "zoom_5": z(500000.0, -.7436447860, .1318252536, iterations=4000, inside=False, exposure=5.0, width=4000, height=2000),
    }


BY_NAME = {"C1": c1, "C2": c2, "C2b": c2b, "C3": c3, "C4": c4}

# Exact loop-iteration totals, used as size-independent checksums at full size.
TOTAL_ITERATIONS = {
    "C1": 85_781_804,
    "C2": 12_911_865_777,
    "C2b": 18_662_294_162,
    "C3": 1_341_751_717,
    "C4": 574_496_116_940,
}
