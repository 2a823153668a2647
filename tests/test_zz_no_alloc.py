"""frb_render_static_async — the render entry point that allocates nothing — against frb_render_device_async.

Why the entry point exists: some hosts cannot call cudaMalloc at the moment they want a frame.  A plug-in inside a process
whose pool allocator already owns all of HBM gets an out-of-memory error (or starves its host); an interactive loop cannot
take the device-wide stall that cudaMalloc / cudaFree cause whenever the frame size changes; code called from inside another
library's callback may not re-enter the CUDA allocator at all.  For them the library keeps the frame's scratch (coordinate
tables, per-pixel records, work counter) in one fixed `static __device__` array (csrc/frb_noalloc.cu), at the price of a
frame-size ceiling and of calls that must not overlap on the GPU.

What is checked, through the C ABI: the image equals the allocating entry point's byte for byte (both are exact kernels and
share the shade code, so the tolerance is zero — the oracle comparison beside it uses the suite's usual +-1 LSB on colour) on
several calls IN A ROW with different scenes and sizes, i.e. with the previous call's leftovers in the shared array; the bytes
behind the image stay untouched; frames beyond the ceiling are refused with a message that names it.

This file sorts last on purpose (pytest -x stops at the first failure, and a failure here should not hide the rest of the
suite); it must also pass alone in a fresh process, where its first call is the one that does the per-device set-up.
"""
import ctypes

import numpy as np
import pytest

import oracle_binding as ob
from fractal_renderer_b200 import _lib, configs
from fractal_renderer_b200.calc import Algo, Config

MAX_PIXELS = 1 << 20      # the ceiling stated in include/frb.h for frb_render_static_async
MAX_EDGE_SUM = 16384


def _scenes():
    return [
        configs.c1().with_(width=640, height=360),                      # Mandelbrot, cap 256, inside + smooth
        configs.c3().with_(width=512, height=512),                      # Julia, cap 1000
        configs.c2().with_(width=300, height=200, iterations=1500),     # deep Mandelbrot view, inside off, smaller than the last
        configs.c1().with_(width=1024, height=1024, iterations=64),     # exactly the pixel ceiling
        Config.new(Algo.BarnsleyFern).with_(width=64, height=48),       # the per-pixel arm of the fern: all black
    ]


def _reference(frb, torch, cfg, fmt):
    n = cfg.width * cfg.height * fmt
    dev = torch.full((n + 64,), 0x5A, dtype=torch.uint8, device="cuda")
    frb.render_device_async(cfg, dev.data_ptr(), n, fmt, device=0)
    frb.sync_stats()
    torch.cuda.synchronize()
    return dev.cpu().numpy()


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", [3, 4])
def test_static_scratch_render_matches_the_allocating_entry_point_on_calls_in_a_row(gpu, fmt):
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    scenes = _scenes()
    refs = [_reference(gpu, torch, cfg, fmt) for cfg in scenes]
    # destinations are the caller's: allocated here, before the calls under test
    outs = [torch.full((r.size,), 0x5A, dtype=torch.uint8, device="cuda") for r in refs]
    torch.cuda.synchronize()
    # back-to-back on ONE stream (the legacy default stream), no synchronisation in between: stream order is what keeps
    # the calls from racing on the shared array; then the whole sequence a second time over the first pass's leftovers
    for _ in range(2):
        for cfg, out in zip(scenes, outs):
            out.fill_(0x5A)
            gpu.render_static_async(cfg, out.data_ptr(), cfg.width * cfg.height * fmt, fmt, device=0)
        torch.cuda.synchronize()
        for cfg, out, ref in zip(scenes, outs, refs):
            got = out.cpu().numpy()
            n = cfg.width * cfg.height * fmt
            assert np.array_equal(got[:n], ref[:n]), f"{cfg.algo.name} {cfg.width}x{cfg.height} fmt {fmt}"
            assert (got[n:] == 0x5A).all(), "bytes behind the image were written"


@pytest.mark.gpu
def test_static_scratch_render_twice_in_a_row_same_scene_and_against_the_oracle(gpu):
    torch = pytest.importorskip("torch")
    cfg = configs.c1().with_(width=320, height=180)
    n = cfg.width * cfg.height * 3
    ref = _reference(gpu, torch, cfg, 3)[:n]
    a = torch.zeros(n, dtype=torch.uint8, device="cuda")
    b = torch.zeros(n, dtype=torch.uint8, device="cuda")
    st = torch.cuda.Stream()
    torch.cuda.synchronize()
    gpu.render_static_async(cfg, a.data_ptr(), n, 3, device=0, stream=st.cuda_stream)   # first call
    gpu.render_static_async(cfg, b.data_ptr(), n, 3, device=0, stream=st.cuda_stream)   # second call, same stream: ordered
    st.synchronize()
    assert np.array_equal(a.cpu().numpy(), ref)
    assert np.array_equal(b.cpu().numpy(), ref)
    want = ob.render_rows(cfg)["rgb"].reshape(-1)
    assert np.abs(a.cpu().numpy().astype(int) - want.astype(int)).max() <= 1   # colour: +-1 LSB against the CPU oracle


@pytest.mark.gpu
def test_static_scratch_render_refuses_frames_beyond_its_ceiling(gpu):
    torch = pytest.importorskip("torch")
    out = torch.zeros(64, dtype=torch.uint8, device="cuda")
    for w, h in ((1025, 1024), (MAX_EDGE_SUM, 1), (1, MAX_PIXELS + 1)):
        with pytest.raises(gpu.FrbError) as e:
            gpu.render_static_async(configs.c1().with_(width=w, height=h), out.data_ptr(), 1 << 40, 3, device=0)
        assert e.value.code == _lib.FRB_ERR_TOO_LARGE
        assert str(MAX_PIXELS) in str(e.value) and str(MAX_EDGE_SUM) in str(e.value)
    # too small a destination, a misaligned one, an unknown format: refused before anything is launched
    cfg = configs.c1().with_(width=64, height=64)
    for ptr, length, fmt in ((out.data_ptr(), 63, 3), (out.data_ptr() + 4, 64 * 64 * 3, 3), (out.data_ptr(), 64 * 64 * 3, 5)):
        with pytest.raises(gpu.FrbError) as e:
            gpu.render_static_async(cfg, ptr, length, fmt, device=0)
        assert e.value.code == _lib.FRB_ERR_INVALID
    # an empty frame is fine and writes nothing
    gpu.render_static_async(configs.c1().with_(width=0, height=7), 0, 0, 3, device=0)
    torch.cuda.synchronize()
    assert (out.cpu().numpy() == 0).all()


def test_static_scratch_render_argument_checks_need_no_device():
    """The ceiling and the no-CPU-path contract, on a box without a GPU: too large is refused as such, an acceptable frame
    fails with FRB_ERR_NO_DEVICE instead of being computed somewhere else."""
    L = _lib.load()
    big = configs.c1().with_(width=2048, height=2048).to_c()
    assert L.frb_render_static_async(ctypes.byref(big), ctypes.c_void_p(0x1000), 1 << 40, 3, 0, None) == _lib.FRB_ERR_TOO_LARGE
    assert b"1048576" in L.frb_last_error()
    if L.frb_device_count() > 0:
        pytest.skip("a GPU is visible; the no-device half is checked on CPU boxes")
    ok = configs.c1().with_(width=64, height=64).to_c()
    assert L.frb_render_static_async(ctypes.byref(ok), ctypes.c_void_p(0x1000), 64 * 64 * 3, 3, 0, None) == _lib.FRB_ERR_NO_DEVICE
