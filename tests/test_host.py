"""Host-side mirror of the reference's value types and the frozen workload definitions (no GPU)."""
import struct

import pytest

from fractal_renderer_b200 import configs
from fractal_renderer_b200.calc import RGB, Algo, Config, Imaginary


def bits(v: float) -> int:
    return struct.unpack("<Q", struct.pack("<d", v))[0]


def test_algo_from_str():
    assert Algo.from_str("Mandelbrot") == Algo.Mandelbrot
    assert Algo.from_str("FERN") == Algo.BarnsleyFern and Algo.from_str("barnsleyfern") == Algo.BarnsleyFern
    assert Algo.from_str("julia") == Algo.Julia
    with pytest.raises(ValueError):
        Algo.from_str("newton")


def test_config_new_defaults():
    c = Config.new(Algo.Mandelbrot)
    assert (c.width, c.height, c.iterations, c.limit, c.stable_limit, c.exposure) == (2000, 1000, 50, 65536.0, 2.0, 2.0)
    assert c.scale == Imaginary(0.4, 0.4) and c.pos == Imaginary(0.0, 0.0) and c.inside and c.smooth
    assert c.primary_color == RGB(r=40, g=255, b=40) and c.secondary_color == RGB(r=240, g=0, b=170)
    f = Config.new(Algo.BarnsleyFern)
    assert f.iterations == 10_000_000 and f.primary_color == RGB(r=4, g=3, b=100)


def test_cli_defaults_differ_from_config_new():
    c = Config.cli(Algo.Mandelbrot)
    assert (c.width, c.height, c.exposure, c.pos.re) == (750, 500, 5.0, -0.6)
    assert Config.cli(Algo.Julia).pos.re == 0.0
    # the fern keeps -0.6 (only julia resets pos_x, src/lib.rs:66-71) and takes iterations and colours from Config::new(algo)
    f = Config.cli(Algo.BarnsleyFern)
    assert f.pos.re == -0.6 and f.iterations == 10_000_000 and (f.width, f.height, f.exposure) == (750, 500, 5.0)
    assert f.primary_color == Config.new(Algo.BarnsleyFern).primary_color == RGB(r=4, g=3, b=100)
    assert f.secondary_color == RGB(r=240, g=240, b=240)
    assert Config.cli(Algo.Julia).iterations == 50 and Config.cli(Algo.Mandelbrot).secondary_color == RGB(r=240, g=0, b=170)


def test_hex_colour_goes_through_rgb_new():
    assert RGB.from_hex("2828ff") == RGB.new(0x28, 0x28, 0xFF) == RGB(r=40, g=255, b=40)


def test_config_input_bit_patterns():
    # SURVEY.md Appendix B
    assert bits(configs.c1().pos.re) == 0xBFE3333333333333
    assert bits(configs.c1().scale.re) == 0x3FD999999999999A
    assert bits(configs.c2().pos.re) == 0xBFE7CBF02676BFD8 and bits(configs.c2().pos.im) == 0x3FC0DFA6607FDC29
    assert bits(configs.c3().julia_set.re) == 0xBFE742C3C9EECBFB and bits(configs.c3().julia_set.im) == 0x3FC82DE00D1B7176
    assert bits(configs.c4().pos.re) == 0xBFE7CBEE43D63CBE and bits(configs.c4().pos.im) == 0x3FC0DFABD5A9E9AE


def test_c5_scales():
    s = [configs.c5_scale(k) for k in range(configs.C5_FRAMES)]
    assert s[0] == 1.0 and abs(s[-1] / 1e12 - 1) < 1e-12 and all(b > a for a, b in zip(s, s[1:]))
    assert len(configs.c5()) == 240 and configs.c5_frame(7).iterations == 10000


def test_roundtrip_through_c_struct():
    c = configs.c3()
    assert Config.from_c(c.to_c()) == c
