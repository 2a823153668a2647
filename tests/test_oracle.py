"""The CPU oracle against everything that can pin it (no GPU needed).

The reference has no tests, so the pins are: a by-hand orbit, the survey's independent restatement
(SHA-256 of counts / PPM, known-answer pixels, iteration totals — SURVEY.md Appendix C), the reference's screenshot,
the committed golden vectors, and a second, independent restatement of the spec in pure Python (tests/pyref.py)."""
import hashlib
import math
import os

import numpy as np
import pytest

import cases
import oracle_binding as ob
import pyref
from fractal_renderer_b200 import configs
from fractal_renderer_b200.calc import RGB, Algo, Config, Imaginary

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_kat_orbit_by_hand():
    # start = c = (1, 0), limit 65536: orbit 2, 5, 26, 677, 458330 -> returns ((458330, 0), 4)  
    (re, im), n = ob.recursive(100, complex(1.0, 0.0), complex(1.0, 0.0), 65536.0)
    assert (re, im, n) == (458330.0, 0.0, 4)
    # never escapes: c = 0 -> z stays 0, returns (previous, iterations)
    (re, im), n = ob.recursive(50, complex(0.0, 0.0), complex(0.0, 0.0), 65536.0)
    assert (re, im, n) == (0.0, 0.0, 50)
    # iterations = 0 returns the start untouched
    (re, im), n = ob.recursive(0, complex(0.3, -0.2), complex(9.0, 9.0), 2.0)
    assert (re, im, n) == (0.3, -0.2, 0)


def test_c1_known_answer_pixels_and_hashes():
    cfg = configs.c1()
    r = ob.render_rows(cfg, want_z=False)
    assert r["total"] == configs.TOTAL_ITERATIONS["C1"]
    counts = np.where(r["iters"] < cfg.iterations, r["iters"] + 1, cfg.iterations).astype("<u4")
    kat = {(0, 0): (4, (0, 0, 2)), (960, 540): (256, (42, 30, 0)), (1919, 1079): (4, (0, 0, 3)),
           (700, 300): (7, (2, 2, 18)), (1200, 540): (256, (0, 0, 0)), (400, 540): (256, (225, 159, 0)),
           (1000, 200): (8, (3, 3, 24)), (1365, 333): (15, (8, 8, 55))}
    for (x, y), (n, rgb) in kat.items():
        assert int(counts[y, x]) == n
        assert tuple(int(v) for v in r["rgb"][y, x]) == rgb
    assert hashlib.sha256(counts.tobytes()).hexdigest() == "b054f2cb1da4f85f828fe838a03bec82523105a616bce721753e68227bd69576"
    ppm = f"P6\n{cfg.width} {cfg.height}\n255\n".encode() + r["rgb"].tobytes()
    # colour hash depends on the platform log2 (glibc here and on the GPU box: same image)
    assert hashlib.sha256(ppm).hexdigest() == "e7778efc46ac5f1caa26d3989b6063c51507d976a03829615f3edf0d3db86784"


def test_c3_total_and_hash():
    cfg = configs.c3()
    r = ob.render_rows(cfg, want_z=False, want_rgb=False)
    assert r["total"] == configs.TOTAL_ITERATIONS["C3"]
    counts = np.where(r["iters"] < cfg.iterations, r["iters"] + 1, cfg.iterations).astype("<u4")
    assert hashlib.sha256(counts.tobytes()).hexdigest() == "3bb6c7294c7d8923c68d4207c812bd70c4be399dc7c0ceafb8c5fc5c1b1a2945"


def test_screenshot_fixture():
    """The reference's only artefact: 1000x1000 lossy AVIF of the C2 view (every 3rd pixel of the 3000^2 grid is
    coordinate-identical to the 1000^2 grid).  IoU of the interior mask and mean |delta| per channel."""
    from PIL import Image

    shot = np.asarray(Image.open(os.path.join(GOLDEN, "screenshot_mandelbrot_1000000x.png")).convert("RGB")).astype(int)
    cfg = configs.c2().with_(width=1000, height=1000)
    r = ob.render_rows(cfg, want_z=False)
    inter_o = r["iters"] == cfg.iterations
    inter_s = shot.sum(-1) < 6
    iou = (inter_o & inter_s).sum() / (inter_o | inter_s).sum()
    assert iou >= 0.9999
    mad = np.abs(r["rgb"].astype(int) - shot).mean(axis=(0, 1))
    assert (mad <= 1.0).all(), mad


def test_golden_vectors_frozen():
    g = np.load(os.path.join(GOLDEN, "small_cases.npz"))
    for name, cfg in cases.all_small_cases().items():
        r = ob.render_rows(cfg, threads=1)
        assert np.array_equal(r["iters"], g[name + "/iters"]), name
        assert np.array_equal(r["z"].view(np.uint64), g[name + "/zbits"]), name
        assert np.array_equal(r["rgb"], g[name + "/rgb"]), name
        assert r["total"] == int(g[name + "/total"]), name


@pytest.mark.parametrize("name", ["mandel_default", "mandel_iter0", "mandel_iter1", "mandel_no_smooth", "mandel_no_inside",
                                  "mandel_aniso_scale", "mandel_tall", "mandel_limit1p5", "mandel_limit1e155",
                                  "mandel_colors_255", "mandel_exposure_neg", "julia_default", "julia_zero", "fern_black",
                                  "mandel_deep_subulp", "mandel_stable_big", "mandel_iter15", "mandel_iter50", "mandel_iter300",
                                  "julia_iter301", "mandel_c2_view", "julia_c3_view"])
def test_oracle_vs_independent_python_restatement(name):
    """tests/pyref.py restates the spec with Python floats (IEEE binary64, no fusion); same bits."""
    cfg = cases.all_small_cases()[name]
    w, h = min(cfg.width, 24), min(cfg.height, 16)
    cfg = cfg.with_(width=w, height=h)
    r = ob.render_rows(cfg, threads=1)
    for y in range(h):
        for x in range(w):
            (zr, zi), n, rgb = pyref.get_recursive_pixel(cfg, x, y)
            assert n == int(r["iters"][y, x]), (name, x, y)
            got = r["z"][y, x]
            assert pyref.same_bits(zr, got[0]) and pyref.same_bits(zi, got[1]), (name, x, y)
            assert rgb == tuple(int(v) for v in r["rgb"][y, x]), (name, x, y)


@pytest.mark.parametrize("name", sorted(configs.examples_md()))
def test_both_restatements_agree_on_the_reference_documented_inputs(name):
    """examples.md:7-29 holds the only inputs the reference documents (14 command lines, no expected outputs).  The C oracle and
    the independent Python restatement must agree bit for bit on each of them AT ITS TRUE SIZE (the coordinate map depends on
    width and height), on a grid of sample pixels."""
    cfg = configs.examples_md()[name]
    ys = np.linspace(0, cfg.height - 1, 6).astype(np.uint32)
    xs = np.linspace(0, cfg.width - 1, 10).astype(int)
    r = ob.render_rows(cfg, rows=ys)
    for j, y in enumerate(ys):
        for x in xs:
            (zr, zi), n, rgb = pyref.get_recursive_pixel(cfg, int(x), int(y))
            assert n == int(r["iters"][j, x]), (name, x, y)
            got = r["z"][j, x]
            assert pyref.same_bits(zr, got[0]) and pyref.same_bits(zi, got[1]), (name, x, y)
            assert rgb == tuple(int(v) for v in r["rgb"][j, x]), (name, x, y)


def test_rgb_new_swaps_g_and_b():
    # RGB::new(r, b, g) (calc/src/lib.rs:129-131) and color_multiply's swap (133-139): with exposure giving
    # mult = 1 the output is (P.r, P.b, P.g) of the STORED fields.
    cfg = Config.cli(Algo.Mandelbrot, width=4, height=4, iterations=10, smooth=False, pos=Imaginary(100.0, 100.0),
                     primary_color=RGB(r=10, g=20, b=30))
    # every pixel escapes at i = 0 -> mult = 0/10*e = 0 -> black; use inside branch instead for a clean check
    cfg2 = cfg.with_(pos=Imaginary(0.0, 0.0), scale=Imaginary(1e9, 1e9), secondary_color=RGB(r=10, g=20, b=30),
                     stable_limit=1e9, iterations=1)
    px = ob.get_recursive_pixel(cfg2, 2, 2)
    # z after 1 iteration ~ c^2 + c with |c| ~ 1e-9 -> dist ~ 1e-18 -> all zero.  Use stable_limit large + far point:
    cfg3 = cfg2.with_(pos=Imaginary(1.0, 0.0), iterations=0)  # z = start = (1, ~0): dist = 1 -> secondary * 1
    assert ob.get_recursive_pixel(cfg3, 2, 2) == (10, 30, 20)
    assert px == (0, 0, 0)
    assert RGB.new(40, 40, 255) == RGB(r=40, g=255, b=40)


def test_iterations_zero_divides_by_zero_like_rust():
    # mult = 0/0*e = NaN -> `as u8` = 0 ; smooth adds (1 - nu) -> finite/0 = +-inf -> 255 or 0
    cfg = Config.cli(Algo.Mandelbrot, width=8, height=8, iterations=0, smooth=False)
    r = ob.render_rows(cfg)
    far = r["z"][..., 0] ** 2 + r["z"][..., 1] ** 2 > 2.0
    assert (r["rgb"][far] == 0).all()
    cfg = cfg.with_(smooth=True)
    r2 = ob.render_rows(cfg)
    assert set(np.unique(r2["rgb"][far])) <= {0, 255}


def test_shard_rows_cover_image_once():
    import fractal_renderer_b200 as f

    for h in (1, 15, 16, 17, 100, 1080, 3000):
        for count in (1, 2, 3, 8):
            for br in (1, 16, 64):
                rows = np.concatenate([f.Shard(i, count, br).image_rows(h) for i in range(count)])
                assert sorted(rows.tolist()) == list(range(h)), (h, count, br)
