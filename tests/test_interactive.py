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


def _oracle_rgba_ok(frame, cfg, rows=None):
    """frame (H, W, 4) against the CPU oracle: RGB within +-1 LSB (BASELINE's tolerance), alpha 255."""
    import oracle_binding as ob

    ref = ob.render_rows(cfg, rows=rows, want_iters=False, want_z=False)["rgb"]
    got = frame if rows is None else frame[rows]
    assert frame.shape == (cfg.height, cfg.width, 4)
    assert (frame[..., 3] == 255).all()
    assert np.abs(got[..., :3].astype(int) - ref.astype(int)).max() <= 1


@pytest.mark.gpu
def test_frame_server_scripted_session_against_the_oracle(gpu):
    """A GUI session replayed through ViewState + FrameServer (src/gui.rs:287-341): arrow keys, scroll in and out, resize,
    algorithm switch, more iterations — every delivered frame is checked against the ORACLE, not against the library."""
    view = ViewState(Config.cli(Algo.Mandelbrot, width=320, height=200, iterations=120))
    script = [
        lambda v: None,
        lambda v: v.apply_keys(0.25, right=True),
        lambda v: v.apply_keys(0.25, down=True, left=True),
        lambda v: v.apply_scroll(40.0),
        lambda v: v.apply_scroll(120.0),
        lambda v: v.apply_scroll(-30.0),
        lambda v: v.set_size(417, 233),                      # ragged size
        lambda v: setattr(v, "config", v.config.with_(iterations=700)),   # past the tile kernel's cap: deferred kernel
        lambda v: v.apply_keys(0.1, up=True),
        lambda v: v.set_algo(Algo.Julia),                    # resets everything to Config::new(Julia): 2000x1000, c = 0
        lambda v: setattr(v, "config", v.config.with_(width=400, height=240, julia_set=Imaginary(-0.7269, 0.1889), iterations=300)),
        lambda v: v.apply_scroll(80.0),
        lambda v: v.set_size(8, 8),                          # clamped to 16x16
    ]
    with FrameServer(device=0) as srv:
        for step in script:
            step(view)
            srv.request(view.config)
            frame = srv.poll(20000)
            assert frame is not None
            cfg = view.config
            rows = None if cfg.height * cfg.width <= 200_000 else np.arange(3, cfg.height, 25, dtype=np.uint32)
            _oracle_rgba_ok(frame, cfg, rows)
        st = srv.stats()
        assert st["frames_rendered"] == len(script) and st["coalesced"] == 0


@pytest.mark.gpu
def test_frame_server_fern_frames(gpu, monkeypatch):
    """Algo = BarnsleyFern through the server: the reference's render thread calls get_image, which draws the fern
    (src/lib.rs:271-319) — not the all-black per-pixel arm.  The seed is entropy, so the check is structural: background =
    secondary colour, a few per cent of the pixels darkened, alpha 255.  The thread count is the reference's
    (rayon::current_num_threads: RAYON_NUM_THREADS, else the cores) and decides how much of the fern survives the saturating sum
    of the per-thread images (on a 200-core box almost nothing does), so it is pinned here."""
    monkeypatch.setenv("RAYON_NUM_THREADS", "8")
    view = ViewState()
    view.set_algo(Algo.BarnsleyFern)
    view.config = view.config.with_(width=400, height=200, iterations=200_000)
    with FrameServer(device=0) as srv:
        srv.request(view.config)
        frame = srv.poll(20000)
    assert frame is not None and frame.shape == (200, 400, 4) and (frame[..., 3] == 255).all()
    # every rayon thread's image starts as secondary_color and the T images are added with saturation (src/lib.rs:285-318): the
    # background is min(255, T * 240) per channel, the plotted points are darker
    px = frame[..., :3].reshape(-1, 3)
    colours, counts = np.unique(px, axis=0, return_counts=True)
    bg = colours[counts.argmax()]
    assert (bg >= 240).all(), bg
    assert 0.5 < counts.max() / px.shape[0] < 0.9995, counts.max() / px.shape[0]
    assert (px.astype(int).sum(1) < int(bg.astype(int).sum())).any()
    rgba = gpu.get_image_rgba(view.config)
    assert rgba.shape == frame.shape and not (rgba[..., :3] == 0).all()


@pytest.mark.gpu
def test_frame_server_latest_wins_under_the_gpu_renderer(gpu):
    """Requests fired faster than frames can be rendered: every delivered frame is one of the requested states (checked against
    the oracle), in request order, and the last delivered frame is the LAST requested state."""
    base = configs.c2().with_(width=640, height=360, iterations=3000)
    states = [base.with_(scale=base.scale * (1.0 + 0.01 * k)) for k in range(40)]
    delivered = []
    with FrameServer(device=0) as srv:
        for k, cfg in enumerate(states):
            srv.request(cfg)
            f = srv.poll(0)
            if f is not None:
                delivered.append(f)
        while True:
            f = srv.poll(3000)
            if f is None:
                break
            delivered.append(f)
        st = srv.stats()
    assert 1 <= len(delivered) <= len(states) and st["frames_rendered"] == len(delivered)
    rows = np.arange(5, base.height, 40, dtype=np.uint32)
    import oracle_binding as ob
    refs = [ob.render_rows(c, rows=rows, want_iters=False, want_z=False)["rgb"] for c in states]
    last = -1
    for f in delivered:
        match = [k for k in range(last + 1, len(states)) if np.abs(f[rows][..., :3].astype(int) - refs[k].astype(int)).max() <= 1]
        assert match, "a delivered frame is none of the (later) requested states"
        last = match[0]
    assert np.abs(delivered[-1][rows][..., :3].astype(int) - refs[-1].astype(int)).max() <= 1, "the last frame is not the last state"


@pytest.mark.gpu
def test_screenshot_thread_overlaps_the_render_thread(gpu):
    """The `S` key (src/gui.rs:319-328) renders the view at 2x on its own thread while the render thread keeps redrawing:
    both results must be the oracle's."""
    view = ViewState(Config.cli(Algo.Mandelbrot, width=300, height=180, iterations=400))
    shot_cfg = view.screenshot_config()
    result = {}

    def screenshot():
        result["img"] = gpu.get_image(shot_cfg).reshape(shot_cfg.height, shot_cfg.width, 3)

    with FrameServer(device=0) as srv:
        th = threading.Thread(target=screenshot)
        th.start()
        for k in range(6):
            view.apply_scroll(20.0)
            srv.request(view.config)
            frame = srv.poll(20000)
            assert frame is not None
            _oracle_rgba_ok(frame, view.config)
        th.join()
    import oracle_binding as ob
    ref = ob.render_rows(shot_cfg, want_iters=False, want_z=False)["rgb"]
    assert np.abs(result["img"].astype(int) - ref.astype(int)).max() <= 1


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
            _oracle_rgba_ok(frame, cfg, np.arange(1, cfg.height, 37, dtype=np.uint32))


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
