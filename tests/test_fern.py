"""Barnsley fern: get_image's BarnsleyFern arm (reference src/lib.rs:271-319, 366-463).

The reference seeds its generators from OS entropy and splits the work over the host's rayon threads, so no two of its
runs agree: parity is statistical by nature.  What is checked instead:
  CPU  - the oracle's colour map against an independent pure-Python restatement of subtract_pixel;
       - the chain decomposition the CUDA path uses against the reference's own structure (one uninterrupted orbit per
         thread), statistically, on coarse histograms;
  GPU  - frb_render_fern against the oracle's restatement of the same decomposition, BIT FOR BIT (hit counts are integers
         and order independent), over views, thread counts, formats and shards.
"""
import ctypes
import math

import numpy as np
import pytest

import oracle_binding as ob
from fractal_renderer_b200 import _lib
from fractal_renderer_b200.calc import RGB, Algo, Config, Imaginary

FERN = Config.new(Algo.BarnsleyFern)


def py_sat_u8(v: float) -> int:
    if v != v or v <= 0.0:
        return 0
    return 255 if v >= 255.0 else int(v)


def py_colour_after(cfg: Config, n: int):
    """n applications of subtract_pixel (src/lib.rs:388-408) to secondary_color, in plain Python floats."""
    r, g, b = cfg.secondary_color.r, cfg.secondary_color.g, cfg.secondary_color.b
    v = cfg.primary_color

    def chan(pixel, value):
        try:
            inv = 1.0 / (value / 255.0)
        except ZeroDivisionError:
            inv = math.inf
        d = ((inv - 1.0) * cfg.color_weight) + 1.0
        try:
            return py_sat_u8(pixel * 1.0 / d)
        except ZeroDivisionError:
            return py_sat_u8(math.copysign(math.inf, d) if pixel else math.nan)

    for _ in range(n):
        nr, ng, nb = chan(r, v.r), chan(g, v.g), chan(b, v.b)
        r, b, g = nr, ng, nb  # RGB::new(nr, ng, nb) with new(r, b, g)
    return (r, g, b)


@pytest.mark.parametrize("cfg", [
    FERN,
    FERN.with_(color_weight=0.5),
    FERN.with_(color_weight=0.0),
    FERN.with_(color_weight=-0.004),
    FERN.with_(primary_color=RGB(255, 0, 128), secondary_color=RGB(255, 200, 7)),
    FERN.with_(primary_color=RGB(1, 254, 255), secondary_color=RGB(13, 250, 100), color_weight=0.003),
])
def test_colour_map_matches_python_restatement(cfg):
    for n in list(range(0, 40)) + [63, 64, 100, 255, 256, 511, 512, 1000, 1023]:
        assert ob.fern_colour_after(cfg, n) == py_colour_after(cfg, n), n
    # the state has period <= 2 long before 1022 hits (frb_fern.cuh relies on it for n >= 1024)
    assert py_colour_after(cfg, 1022) == py_colour_after(cfg, 1020) and py_colour_after(cfg, 1023) == py_colour_after(cfg, 1021)
    assert ob.fern_colour_after(cfg, 10_000_001) == py_colour_after(cfg, 1023)
    assert ob.fern_colour_after(cfg, 10_000_000) == py_colour_after(cfg, 1022)


def test_default_fern_config():
    assert FERN.iterations == 10_000_000 and FERN.color_weight == 0.01
    assert (FERN.primary_color.r, FERN.primary_color.g, FERN.primary_color.b) == (4, 3, 100)  # RGB::new(4, 100, 3)


def test_oracle_is_deterministic_and_seed_sensitive():
    cfg = FERN.with_(width=200, height=100, iterations=50_000)
    a = ob.fern_counts(cfg, 3, 42)
    assert np.array_equal(a, ob.fern_counts(cfg, 3, 42))
    assert not np.array_equal(a, ob.fern_counts(cfg, 3, 43))
    assert a.sum() == (50_000 // 3) * 3  # every plotted point of the default view lands on a pixel; the remainder is dropped
    assert [int(x) for x in a.reshape(3, -1).sum(1)] == [50_000 // 3] * 3


@pytest.mark.parametrize("view", [
    dict(),
    dict(pos=Imaginary(0.02, 0.1), scale=Imaginary(1.5, 1.5)),
])
def test_chain_decomposition_agrees_with_reference_structure(view):
    """Same distribution: 8x8-pixel bins of the hit histogram agree within sampling noise.  Successive points of an orbit
    are correlated, so dense bins scatter more than a Poisson count would (two reference-style runs with different seeds
    differ by up to ~1.5 % there): tolerance = 5 Poisson sigma + 3 % of the bin."""
    cfg = FERN.with_(width=320, height=160, iterations=4_000_000, **view)
    a = ob.fern_counts(cfg, 4, 7).sum(0).astype(np.int64)
    b = ob.fern_counts(cfg, 4, 1007, reference_style=True).sum(0).astype(np.int64)
    assert abs(int(a.sum()) - int(b.sum())) <= 0.01 * b.sum() + 64  # points off the image differ from run to run
    A = a.reshape(20, 8, 40, 8).sum((1, 3))
    B = b.reshape(20, 8, 40, 8).sum((1, 3))
    tol = 5.0 * np.sqrt(A + B + 1.0) + 0.03 * (A + B) / 2 + 8
    assert (np.abs(A - B) <= tol).all(), float((np.abs(A - B) / tol).max())
    # and two reference-style runs with different seeds differ from each other by about as much
    c = ob.fern_counts(cfg, 4, 2007, reference_style=True).sum(0).astype(np.int64)
    C = c.reshape(20, 8, 40, 8).sum((1, 3))
    assert (np.abs(C - B) <= 5.0 * np.sqrt(C + B + 1.0) + 0.03 * (C + B) / 2 + 8).all()


M64 = (1 << 64) - 1


class PyXoshiro:
    """xoshiro256++ seeded by splitmix64 (rand 0.8 SmallRng::seed_from_u64), in Python integers."""

    def __init__(self, state: int):
        self.s = []
        for _ in range(4):
            state = (state + 0x9E3779B97F4A7C15) & M64
            z = state
            z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
            z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
            self.s.append(z ^ (z >> 31))

    @staticmethod
    def rotl(x, k):
        return ((x << k) | (x >> (64 - k))) & M64

    def f64(self) -> float:
        s = self.s
        result = (self.rotl((s[0] + s[3]) & M64, 23) + s[0]) & M64
        t = (s[1] << 17) & M64
        s[2] ^= s[0]
        s[3] ^= s[1]
        s[1] ^= s[2]
        s[0] ^= s[3]
        s[2] ^= t
        s[3] = self.rotl(s[3], 45)
        return (result >> 11) * (1.0 / 9007199254740992.0)


def py_sat_usize(v: float) -> int:
    if v != v or v <= 0.0:
        return 0
    return M64 if v >= 18446744073709551616.0 else int(v)


def py_fern_counts(cfg: Config, threads: int, seed: int):
    """Pure-Python restatement of the chained chaos game (frb.h: frb_render_fern); also reports how many points landed
    exactly one column past the right edge (x == width, which pixel_mut lets through, src/lib.rs:379)."""
    W, Hh = cfg.width, cfg.height
    wf, hf = float(W), float(Hh)
    esx = 65.0 * cfg.scale.re * hf * 0.006
    esy = 37.0 * cfg.scale.im * hf * 0.006
    per = cfg.iterations // threads
    C = min(max((per + 511) // 512, 1), 65536)
    counts = np.zeros((threads, W * Hh), np.uint32)
    wrapped = 0
    for t in range(threads):
        for k in range(C):
            steps = per // C + (1 if k < per % C else 0)
            g = PyXoshiro((seed + (t * C + k + 1) * 0xD1342543DE82EF95) & M64)
            x, y = cfg.pos.re * wf, cfg.pos.im * hf
            for i in range((64 if k > 0 else 0) + steps):
                if i >= (64 if k > 0 else 0):
                    px = py_sat_usize(((x - cfg.pos.re) * esx) + wf / 2.0)
                    py = py_sat_usize(hf - ((y + (cfg.pos.im - 5.0) - 0.5) * esy + hf / 2.0))
                    if px <= W:
                        idx = (py * W + px) & M64
                        if idx < W * Hh:
                            counts[t, idx] += 1
                            wrapped += px == W
                r = g.f64()
                ox = x
                if r < 0.01:
                    x = 0.00 * x + 0.00 * y
                    y = 0.00 * ox + 0.16 * y + 0.00
                elif r < 0.86:
                    x = 0.85 * x + 0.04 * y
                    y = -0.04 * ox + 0.85 * y + 1.60
                elif r < 0.93:
                    x = 0.20 * x - 0.26 * y
                    y = 0.23 * ox + 0.22 * y + 1.60
                else:
                    x = -0.15 * x + 0.28 * y
                    y = 0.26 * ox + 0.24 * y + 0.44
    return counts.reshape(threads, Hh, W), wrapped


def test_oracle_counts_match_python_restatement_including_x_equal_width():
    """Second, independent restatement of the hit counts (generator, maps, pixel mapping, bounds).  The narrow view puts
    points exactly one column past the right edge: pixel_mut rejects x > width only, so they land on the next row."""
    cfg = FERN.with_(width=37, height=23, iterations=6_000)
    got = ob.fern_counts(cfg, 2, 99)
    ref, _ = py_fern_counts(cfg, 2, 99)
    assert np.array_equal(got, ref)
    narrow = FERN.with_(width=6, height=40, iterations=6_000, pos=Imaginary(-0.02, 0.0), scale=Imaginary(0.35, 0.4))
    got = ob.fern_counts(narrow, 3, 5)
    ref, wrapped = py_fern_counts(narrow, 3, 5)
    assert np.array_equal(got, ref)
    assert wrapped > 0, "the case x == width was not exercised; adjust the view"


def test_no_device_means_error_not_fallback():
    L = _lib.load()
    if L.frb_device_count() > 0:
        pytest.skip("a CUDA device is present")
    c = FERN.with_(width=16, height=16, iterations=100).to_c()
    out = np.zeros(16 * 16 * 3, np.uint8)
    rc = L.frb_render_fern(ctypes.byref(c), 1, 1, None, out.ctypes.data, out.nbytes, 3, 0)
    assert rc == _lib.FRB_ERR_NO_DEVICE
    assert L.frb_fern_default_threads() >= 1


# ------------------------------------------------------------ GPU -----------------------------------------------------

GPU_CASES = {
    "default_small": (FERN.with_(width=400, height=200, iterations=300_000), 4, 11),
    "one_thread": (FERN.with_(width=300, height=150, iterations=100_000), 1, 12),
    "sixteen_threads": (FERN.with_(width=256, height=128, iterations=1_000_003), 16, 13),
    "more_threads_than_points": (FERN.with_(width=64, height=32, iterations=5), 8, 14),
    "zoomed_panned": (FERN.with_(width=333, height=211, iterations=400_000, pos=Imaginary(0.02, 0.1), scale=Imaginary(1.5, 1.1)), 3, 15),
    "far_start": (FERN.with_(width=200, height=100, iterations=60_000, pos=Imaginary(3.0, -2.0)), 2, 16),
    "heavy_weight": (FERN.with_(width=128, height=64, iterations=2_000_000, color_weight=0.3), 2, 17),
    "negative_weight": (FERN.with_(width=128, height=64, iterations=200_000, color_weight=-0.004,
                                   secondary_color=RGB(10, 20, 30)), 2, 18),
    "colours_255_0": (FERN.with_(width=128, height=64, iterations=200_000, primary_color=RGB(255, 0, 128)), 2, 19),
    "tiny_image_many_hits": (FERN.with_(width=4, height=2, iterations=500_000), 2, 20),
    "one_by_n": (FERN.with_(width=1, height=64, iterations=10_000), 1, 21),
    "huge_scale": (FERN.with_(width=64, height=64, iterations=10_000, scale=Imaginary(1e300, 1e300)), 1, 22),
    "nan_pos": (FERN.with_(width=64, height=64, iterations=1_000, pos=Imaginary(float("nan"), 0.0)), 1, 23),
}


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(GPU_CASES))
def test_gpu_fern_matches_oracle_bit_for_bit(gpu, name):
    cfg, threads, seed = GPU_CASES[name]
    ref = ob.fern_get_image(cfg, threads, seed)
    got = gpu.render_fern(cfg, threads=threads, seed=seed)
    assert got.shape == ref.shape
    assert np.array_equal(got, ref), f"{name}: {np.count_nonzero((got != ref).any(-1))} pixels differ"
    st = gpu.last_stats()
    assert st["iterations"] == (cfg.iterations // threads) * threads
    rgba = gpu.render_fern(cfg, threads=threads, seed=seed, fmt=gpu.FRB_RGBA32)
    assert np.array_equal(rgba[..., :3], ref) and (rgba[..., 3] == 255).all()


@pytest.mark.gpu
def test_gpu_fern_default_config_full_size(gpu):
    """The reference's default fern: 2000x1000, 10^7 points (Config::new(BarnsleyFern)), T = 16."""
    ref = ob.fern_get_image(FERN, 16, 2024)
    got = gpu.render_fern(FERN, threads=16, seed=2024)
    assert np.array_equal(got, ref)
    assert (got != 255).any() and (got == 255).any()


@pytest.mark.gpu
def test_gpu_fern_shards_tile_the_frame(gpu):
    cfg, threads, seed = GPU_CASES["zoomed_panned"]
    whole = gpu.render_fern(cfg, threads=threads, seed=seed)
    out = np.zeros_like(whole)
    for i in range(3):
        sh = gpu.Shard(i, 3, 16)
        part = gpu.render_fern(cfg, threads=threads, seed=seed, shard=sh)
        out[sh.image_rows(cfg.height)] = part
    assert np.array_equal(out, whole)


@pytest.mark.gpu
def test_get_image_dispatches_fern_like_the_reference(gpu):
    """get_image(config) with algo = BarnsleyFern plays the chaos game with fresh entropy (src/lib.rs:271, 428): two calls
    differ in detail, agree in distribution, and are not the all-black image of the per-pixel path."""
    cfg = FERN.with_(width=320, height=160, iterations=2_000_000)
    a = gpu.get_image(cfg).reshape(160, 320, 3).astype(np.int64)
    b = gpu.get_image(cfg).reshape(160, 320, 3).astype(np.int64)
    assert a.any() and not np.array_equal(a, b)
    A = a.reshape(20, 8, 40, 8, 3).mean((1, 3))
    B = b.reshape(20, 8, 40, 8, 3).mean((1, 3))
    assert np.abs(A - B).max() <= 24.0
    assert not gpu.render_shard(cfg, gpu.Shard(), gpu.FRB_RGB24).any()  # per-pixel path: get_recursive_pixel -> BLACK
