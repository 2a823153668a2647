"""Parity cases shared by the golden generator, the oracle tests and the GPU parity tests.

Edge cases follow SURVEY.md Appendix A.4 (all derivable from calc/src/lib.rs:181-257)."""
from __future__ import annotations

from fractal_renderer_b200 import configs
from fractal_renderer_b200.calc import RGB, Algo, Config, Imaginary


def small(cfg: Config, w: int, h: int) -> Config:
    return cfg.with_(width=w, height=h)


def edge_cases() -> dict[str, Config]:
    m = Config.cli(Algo.Mandelbrot, width=96, height=64, iterations=200)
    j = Config.cli(Algo.Julia, width=96, height=64, iterations=200, julia_set=Imaginary(-0.7269, 0.1889))
    deep = configs.c4(64, 48)
    cases = {
        "mandel_default": m,
        "mandel_iter0": m.with_(iterations=0),
        "mandel_iter1": m.with_(iterations=1),
        "mandel_iter7": m.with_(iterations=7),          # < unroll
        "mandel_iter8": m.with_(iterations=8),          # == unroll
        "mandel_iter9": m.with_(iterations=9),
        # caps that are not a multiple of the deferred kernel's block (8 below 2048 iterations, 16 from there): a pixel that
        # runs into the cap is resolved from the start of its last full block, B + (cap mod B) steps short
        "mandel_iter15": m.with_(iterations=15),
        "mandel_iter50": m.with_(iterations=50),
        "mandel_iter203": m.with_(iterations=203),
        "mandel_iter300": m.with_(iterations=300),       # 4 blocks per trip from 256 on
        "mandel_iter2063": m.with_(iterations=2063, width=48, height=32),  # blocks of 16: cap mod 16 = 15
        "julia_iter301": j.with_(iterations=301),
        "mandel_no_smooth": m.with_(smooth=False),
        "mandel_no_inside": m.with_(inside=False),
        "mandel_aniso_scale": m.with_(scale=Imaginary(0.3, 0.7)),
        "mandel_tall": m.with_(width=40, height=150),
        "mandel_w1": m.with_(width=1, height=33),
        "mandel_h1": m.with_(width=77, height=1),
        "mandel_1x1": m.with_(width=1, height=1),
        "mandel_ragged": m.with_(width=131, height=67),  # not a multiple of 4, 32 or 64
        "mandel_limit2": m.with_(limit=2.0),
        "mandel_limit1p5": m.with_(limit=1.5),
        "mandel_limit_tiny": m.with_(limit=1e-3),
        "mandel_limit0": m.with_(limit=0.0),
        "mandel_limit1e150": m.with_(limit=1e150, iterations=40),
        "mandel_limit1e155": m.with_(limit=1e155, iterations=40),  # S = inf: orbit overflows, NaN never escapes
        "mandel_limit_inf": m.with_(limit=float("inf"), iterations=40),
        "mandel_limit_nan": m.with_(limit=float("nan"), iterations=24),
        "mandel_stable_big": m.with_(stable_limit=1e12),  # escaped pixels land in the `inside` branch
        "mandel_colors_255": m.with_(primary_color=RGB(255, 255, 255), secondary_color=RGB(255, 1, 128), exposure=9.0),
        "mandel_exposure_neg": m.with_(exposure=-3.0),
        "mandel_far_out": m.with_(pos=Imaginary(1e6, -1e6)),  # every pixel escapes at i = 0
        "mandel_huge_coords": m.with_(scale=Imaginary(1e-200, 1e-200), iterations=16),  # start overflows: FORM 7 territory
        "mandel_real_axis": m.with_(width=64, height=33, pos=Imaginary(-0.6, 0.0)),  # middle row has im0 == 0 exactly
        "mandel_deep_subulp": deep,
        "mandel_c2_view": small(configs.c2(), 80, 60),
        "julia_default": j,
        "julia_zero": j.with_(julia_set=Imaginary(0.0, 0.0), iterations=64),  # Config::new default; im(z) underflows
        "julia_real_c": j.with_(julia_set=Imaginary(-0.8, 0.0)),
        "julia_real_c_long": j.with_(julia_set=Imaginary(-0.5, 0.0), iterations=4000, width=48, height=32),  # im decays through subnormals
        "julia_tiny_im": j.with_(julia_set=Imaginary(-0.5, 1e-300)),
        "julia_c3_view": small(configs.c3(), 100, 100),
        "julia_no_smooth": j.with_(smooth=False, inside=False),
        # Near-miss territory of the sparse escape filter: start points with |z0|^2 between S/2 and S and a Julia constant that
        # cancels z0^2, so orbits hover around the escape radius.  limit = 2^20 with |c| = 0.49 limit^2 is the last input the
        # deferred kernel accepts (frb_api.cu: deferred_is_safe); at limit = 1e8 the filter's truncation slack would eat the
        # proof's margin, so that one must take the dense kernel.
        "julia_limit_2p20_bigc": j.with_(limit=1048576.0, julia_set=Imaginary(-0.49 * 1048576.0 ** 2, 1000.0),
                                         pos=Imaginary(0.7 * 1048576.0, 0.0), iterations=64),
        "julia_limit1e8_bigc": j.with_(limit=1e8, julia_set=Imaginary(-5.625e15, 1e7), pos=Imaginary(7.5e7, 0.0), iterations=64),
        "fern_black": Config.new(Algo.BarnsleyFern).with_(width=20, height=10, iterations=100),
    }
    return cases


def config_crops() -> dict[str, Config]:
    """The five BASELINE configs at sizes the oracle finishes instantly (same view, fewer pixels)."""
    return {
        "C1_small": small(configs.c1(), 192, 108),
        "C2_small": small(configs.c2(), 150, 150),
        "C2b_small": small(configs.c2b(), 120, 120),
        "C3_small": small(configs.c3(), 128, 128),
        "C4_small": configs.c4(96, 96),
        "C5_f0_small": small(configs.c5_frame(0), 96, 54),
        "C5_f120_small": small(configs.c5_frame(120), 96, 54),
        "C5_f239_small": small(configs.c5_frame(239), 96, 54),
    }


def all_small_cases() -> dict[str, Config]:
    d = dict(edge_cases())
    d.update(config_crops())
    return d
