"""Headless mirror of the reference GUI loop (src/gui.rs): state-update arithmetic and the latest-wins frame server.
The server's coalescing logic is exercised on CPU with an injected renderer; the GPU test renders real frames."""
import threading
import time

import numpy as np
import pytest

from fractal_renderer_b200 import configs
from fractal_renderer_b200.calc import Algo, Config, Imaginary
from fractal_renderer_b200.interactive import FrameServer, ViewState


def test_view_state_movement_as_written():
    v = ViewState(Config.new(Algo.Mandelbrot).with_(scale=Imaginary(0.5, 0.25)))
    v.apply_keys(0.1, left=True)
    assert v.config.pos.re == -(1.0 / 0.5) * 0.1 * 0.5
    v.apply_keys(0.1, right=True)  # the reference uses 1/scale.im for Right (src/gui.rs:293-294)
    assert v.config.pos.re == -(1.0 / 0.5) * 0.1 * 0.5 + (1.0 / 0.25) * 0.1 * 0.5
    v.apply_keys(0.2, up=True)     # ... and 1/scale.re for Up (src/gui.rs:295-297)
    assert v.config.pos.im == -(1.0 / 0.5) * 0.2 * 0.5


def test_view_state_zoom():
    v = ViewState()
    v.apply_scroll(0.5)  # |delta| < 1 ignored
    assert v.config.scale == Imaginary(0.4, 0.4)
    v.apply_scroll(40.0)
    assert v.config.scale.re == 0.4 * (1.0 + 40.0 / 80.0)
    s = v.config.scale.re
    v.apply_scroll(-90.0)  # log10(90/10 + 1)/2 = 0.5
    assert abs(v.config.scale.re - s * 0.5) < 1e-7
    v.apply_scroll(-1e9)   # clamped to a factor of 0
    assert v.config.scale.re == 0.0


def test_view_state_algo_switch_and_screenshot():
    v = ViewState(configs.c2())
    v.set_algo(Algo.Julia)
    assert v.config == Config.new(Algo.Julia)
    v.set_size(8, 4000)
    assert (v.config.width, v.config.height) == (16, 4000)
    shot = v.screenshot_config()
    assert (shot.width, shot.height) == (32, 8000) and shot.scale == v.config.scale


def test_frame_server_coalesces_latest_wins():
    """While a render is in flight further requests only set try_redraw; when the frame is taken the NEWEST state is
    rendered, intermediate ones never are (src/gui.rs:37-48, 110-117)."""
    rendered = []
    gate = threading.Event()

    def slow_renderer(cfg):
        rendered.append(cfg.iterations)
        gate.wait(5.0)
        return np.full((cfg.height, cfg.width, 4), cfg.iterations % 256, np.uint8)

    base = Config.new(Algo.Mandelbrot).with_(width=16, height=16)
    with FrameServer(renderer=slow_renderer) as srv:
        srv.request(base.with_(iterations=1))
        time.sleep(0.1)
        for it in (2, 3, 4, 5):          # arrive while frame 1 is rendering
            srv.request(base.with_(iterations=it))
        st = srv.stats()
        assert st["working"] == 1 and st["try_redraw"] == 1 and st["coalesced"] == 4
        assert srv.poll(0) is None
        gate.set()
        f1 = srv.poll(2000)
        assert f1 is not None and f1.shape == (16, 16, 4) and int(f1[0, 0, 0]) == 1
        f2 = srv.poll(2000)               # taking frame 1 re-requested the newest state (5), not 2, 3 or 4
        assert f2 is not None and int(f2[0, 0, 0]) == 5
        assert rendered == [1, 5]
        assert srv.poll(50) is None
        st = srv.stats()
        assert st["frames_rendered"] == 2 and st["working"] == 0 and st["try_redraw"] == 0


def test_frame_server_reports_render_failure_without_gpu():
    import fractal_renderer_b200 as f

    if f.device_count() > 0:
        pytest.skip("a GPU is visible")
    with FrameServer() as srv:          # default renderer = frb_render: no device -> error, never a CPU fallback
        srv.request(Config.new(Algo.Mandelbrot).with_(width=16, height=16))
        with pytest.raises(f.FrbError) as e:
            srv.poll(5000)
        assert e.value.code == -1


@pytest.mark.gpu
def test_frame_server_renders_rgba_on_gpu(gpu):
    cfg = configs.c1().with_(width=320, height=200)
    want = gpu.get_image_rgba(cfg)
    with FrameServer(device=0) as srv:
        view = ViewState(cfg)
        srv.request(view.config)
        frame = srv.poll(10000)
        assert frame is not None and np.array_equal(frame, want)
        view.apply_scroll(40.0)
        srv.request(view.config)
        frame2 = srv.poll(10000)
        assert frame2 is not None and np.array_equal(frame2, gpu.get_image_rgba(view.config))
        assert not np.array_equal(frame2, frame)


@pytest.mark.gpu
def test_frame_server_frames_are_streamed_into_pinned_memory(gpu):
    """The server's own frames are pinned, so frames large enough for it take the streaming finisher (3 launches: finisher,
    tables, iterate) — and resizing up and down between requests keeps working on the grow-only buffers."""
    big = configs.c1().with_(width=1280, height=720, iterations=600)
    small = configs.c3().with_(width=640, height=360)
    with FrameServer(device=0) as srv:
        for cfg in (big, small, big, big.with_(width=1920, height=1080)):
            srv.request(cfg)
            frame = srv.poll(10000)
            assert frame is not None and frame.shape[:2] == (cfg.height, cfg.width)
            assert np.array_equal(frame, gpu.get_image_rgba(cfg)), (cfg.width, cfg.height)


def test_frame_server_acquire_hands_out_frames_without_copying():
    """frb_server_acquire: the frame itself moves to the caller (src/gui.rs:110 `take()`), three buffers rotate, and a frame
    stays intact while the next ones are rendered and taken by copy."""
    def renderer(cfg):
        return np.full((cfg.height, cfg.width, 4), cfg.iterations % 256, np.uint8)

    base = Config.new(Algo.Mandelbrot).with_(width=32, height=8)
    with FrameServer(renderer=renderer) as srv:
        assert srv.poll(0, copy=False) is None
        srv.request(base.with_(iterations=7))
        a = srv.poll(2000, copy=False)
        assert a is not None and a.shape == (8, 32, 4) and (a == 7).all()
        srv.request(base.with_(iterations=9, width=16))
        b = srv.poll(2000)                      # by copy: does not disturb the acquired frame
        assert (b == 9).all() and b.shape == (8, 16, 4) and (a == 7).all()
        srv.request(base.with_(iterations=11))
        c = srv.poll(2000, copy=False)          # a's buffer goes back into circulation now
        assert (c == 11).all() and c.shape == (8, 32, 4)
        for it in range(20, 30):                # keep rotating: every frame is the one requested
            srv.request(base.with_(iterations=it))
            f = srv.poll(2000, copy=False)
            assert f is not None and (f == it).all()
        assert srv.stats()["frames_rendered"] == 13
