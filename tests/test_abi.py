"""The C-ABI boundary without a GPU: the library loads, exports every symbol include/frb.h declares, struct layouts
match, host-only helpers agree with the oracle, and compute entry points FAIL LOUDLY without a device."""
import ctypes
import os
import re

import numpy as np
import pytest

import oracle_binding as ob
import fractal_renderer_b200 as f
from fractal_renderer_b200 import _lib, configs
from fractal_renderer_b200.calc import Algo, Config, FrbConfig, FrbRGB

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "frb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(frb_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    L = _lib.load()
    syms = header_symbols()
    assert sorted(_lib.EXPORTS) == syms
    for s in syms:
        assert getattr(L, s) is not None, s
    assert L.frb_abi_version() == 1


def test_struct_layouts():
    assert ctypes.sizeof(FrbConfig) == 104
    assert FrbConfig.limit.offset == 16 and FrbConfig.pos.offset == 32 and FrbConfig.exposure.offset == 64
    assert FrbConfig.inside.offset == 72 and FrbConfig.primary_color.offset == 74 and FrbConfig.secondary_color.offset == 77
    assert FrbConfig.color_weight.offset == 80 and FrbConfig.julia_set.offset == 88
    assert ctypes.sizeof(FrbRGB) == 3
    assert ctypes.sizeof(_lib.FrbStats) == 56
    assert ctypes.sizeof(_lib.FrbShard) == 16
    assert ctypes.sizeof(_lib.FrbFp64Peak) == 48


@pytest.mark.parametrize("algo", [Algo.Mandelbrot, Algo.BarnsleyFern, Algo.Julia])
def test_config_default_matches_config_new(algo):
    a, b = FrbConfig(), FrbConfig()
    _lib.load().frb_config_default(ctypes.byref(a), int(algo))
    ob.lib().oracle_config_new(ctypes.byref(b), int(algo))
    assert bytes(a) == bytes(b)
    assert bytes(Config.new(algo).to_c()) == bytes(a)
    assert Config.from_c(a) == Config.new(algo)


def test_shard_helpers_match_python():
    L = _lib.load()
    for h in (0, 1, 15, 16, 17, 1080, 16384):
        for count in (1, 2, 5, 8):
            for br in (0, 1, 16, 100):
                tot = 0
                for i in range(count):
                    sh = f.Shard(i, count, br)
                    n = f.shard_rows(h, sh)
                    tot += n
                    rows = sh.image_rows(h)
                    assert len(rows) == n
                    for lr in (0, n // 2, n - 1):
                        if 0 <= lr < n:
                            assert f.shard_row_to_y(h, sh, lr) == int(rows[lr])
                assert tot == h
    bad = _lib.FrbShard(3, 2, 16, 0)
    assert L.frb_shard_rows(100, ctypes.byref(bad)) == 0


def test_no_device_fails_loudly():
    """On a box without a GPU every compute entry point must return FRB_ERR_NO_DEVICE — there is no CPU path."""
    if f.device_count() > 0:
        pytest.skip("a GPU is visible; the no-device contract is checked on CPU boxes")
    cfg = configs.c1().with_(width=8, height=8)
    with pytest.raises(f.FrbError) as e:
        f.get_image(cfg)
    assert e.value.code == _lib.FRB_ERR_NO_DEVICE
    with pytest.raises(f.FrbError):
        f.render_counts(cfg)
    with pytest.raises(f.FrbError):
        f.init(0)
    with pytest.raises(f.FrbError):
        f.measure_fp64_peak(0)
    with pytest.raises(f.FrbError) as e:
        f.pinned_empty(1024)          # pinned memory needs the driver: no quiet malloc in its place
    import ctypes
    buf = (ctypes.c_uint8 * 4096)()
    assert _lib.load().frb_host_register(buf, 4096) == _lib.FRB_ERR_NO_DEVICE


def test_product_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "fractal_renderer_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                src = open(os.path.join(dirpath, fn)).read()
                assert "oracle_binding" not in src and "calc_oracle" not in src and "libcalc_oracle" not in src, fn


def test_product_library_has_no_tuning_hooks_and_reads_no_environment():
    """frb_set_tuning only works in the -DFRB_TUNING build; the product library's frame path calls getenv nowhere (the one
    getenv left is RAYON_NUM_THREADS in frb_fern_default_threads, which mirrors rayon)."""
    L = _lib.load()
    assert L.frb_set_tuning(b"fast_shade", 0) == _lib.FRB_ERR_INVALID
    assert L.frb_set_tuning(b"reset", 0) == _lib.FRB_ERR_INVALID
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = open(os.path.join(root, "fractal_renderer_b200", "csrc", "frb_api.cu")).read()
    assert src.count("getenv(") == 1 and 'getenv("RAYON_NUM_THREADS")' in src
    for f in ("frb_server.cu", "frb_shared.cu", "frb_kernels.cuh", "frb_deferred.cuh", "frb_stream.cuh", "frb_tile.cuh"):
        assert "getenv" not in open(os.path.join(root, "fractal_renderer_b200", "csrc", f)).read(), f


def test_init_list_and_render_bands_fail_loudly_without_a_device():
    L = _lib.load()
    if L.frb_device_count() > 0:
        pytest.skip("a GPU is visible")
    arr = (ctypes.c_int32 * 2)(0, 0)
    assert L.frb_init_list(arr, 2) == _lib.FRB_ERR_NO_DEVICE
    cfg = Config.new(Algo.Mandelbrot).with_(width=32, height=32).to_c()
    buf = (ctypes.c_uint8 * (32 * 32 * 3))()
    sh = _lib.FrbShard(0, 2, 16, 0)
    assert L.frb_render_bands(ctypes.byref(cfg), ctypes.byref(sh), buf, len(buf), 3, 0) == _lib.FRB_ERR_NO_DEVICE
