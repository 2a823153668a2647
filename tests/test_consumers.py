"""Compiled, non-Python consumers of the C ABI: a plain-C stand-in for the reference's CLI binary (src/main.rs:4-23) against
include/frb.h, and a C++ program through include/fractal_renderer.hpp::get_image.  Built with gcc / g++ against the in-tree
libfrb.so.  Without a GPU they must compile, link, run and exit with the NO_DEVICE status (there is no CPU path); with one,
their bytes must equal the oracle's."""
import os
import subprocess

import numpy as np
import pytest

import oracle_binding as ob
from fractal_renderer_b200 import configs
from fractal_renderer_b200.calc import Algo, Config, Imaginary

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "fractal_renderer_b200")
BUILD = os.path.join(ROOT, "examples", "build")


def build(kind: str) -> str:
    os.makedirs(BUILD, exist_ok=True)
    if kind == "c":
        exe, cmd = os.path.join(BUILD, "frb_main"), ["gcc", "-std=c11", "-O2", "-Wall", "-Wextra", "-Werror", os.path.join(ROOT, "examples", "main.c")]
    else:
        exe, cmd = os.path.join(BUILD, "frb_main_cpp"), ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", os.path.join(ROOT, "examples", "main.cpp")]
    cmd += ["-I", os.path.join(ROOT, "include"), "-L", LIBDIR, "-lfrb", f"-Wl,-rpath,{LIBDIR}", "-o", exe]
    subprocess.check_call(cmd)
    return exe


def run(exe, *args):
    return subprocess.run([exe, *map(str, args)], capture_output=True, timeout=300)


@pytest.mark.parametrize("kind", ["c", "cpp"])
def test_consumers_compile_link_and_fail_loudly_without_a_device(kind):
    import fractal_renderer_b200 as f

    exe = build(kind)
    if f.device_count() > 0:
        pytest.skip("a GPU is visible: the -m gpu leg checks the bytes")
    r = run(exe, 64, 48, 50)
    assert r.returncode == 3, (r.returncode, r.stderr)
    assert r.stdout == b"" and b"no usable CUDA device" in r.stderr


def test_header_is_valid_c89_to_c17_and_cxx():
    """include/frb.h must compile as C (with and without _Static_assert) and as C++; the layout asserts are part of it."""
    src = '#include "frb.h"\nint main(void) { FrbConfig c; frb_config_default(&c, 0); return (int)sizeof c - 104; }\n'
    os.makedirs(BUILD, exist_ok=True)
    path = os.path.join(BUILD, "hdr_check.c")
    open(path, "w").write(src)
    for cc, std in (("gcc", "-std=c99"), ("gcc", "-std=c11"), ("gcc", "-std=c17"), ("g++", "-std=c++11"), ("g++", "-std=c++17")):
        lang = ["-x", "c++"] if cc == "g++" else []
        subprocess.check_call([cc, std, "-Wall", "-Wextra", "-Werror", "-pedantic", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), *lang, path])


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["c", "cpp"])
@pytest.mark.parametrize("algo", ["mandelbrot", "julia"])
def test_consumer_bytes_equal_the_oracle(gpu, kind, algo):
    exe = build(kind)
    w, h, it = 480, 270, 256
    r = run(exe, w, h, it, algo)
    assert r.returncode == 0, r.stderr
    got = np.frombuffer(r.stdout, np.uint8)
    assert got.size == w * h * 3
    if algo == "julia":
        cfg = Config.cli(Algo.Julia, width=w, height=h, iterations=it, julia_set=Imaginary(-0.7269, 0.1889))
    else:
        cfg = Config.cli(Algo.Mandelbrot, width=w, height=h, iterations=it)
    ref = ob.render_rows(cfg)["rgb"].reshape(-1)
    assert np.abs(got.astype(int) - ref.astype(int)).max() <= 1
