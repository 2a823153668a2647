"""frb_render(device = -1): one frame over a LIST of devices — interleaved 16-row bands, one host worker per shard, bands written
(streaming finisher, pinned destination) or copied (strided 2-D copies, slab by slab for large shards) into ONE caller-owned
image.  frb_init_list may name a device more than once, so the whole multi-shard path also runs — shard after shard — on a
single-GPU box; on a multi-GPU box the real devices are used as well (and a fresh-process test makes a device = -1 frame the
first thing a process ever renders).  The image must equal the single-device render byte for byte."""
import subprocess
import sys

import numpy as np
import pytest

import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    assert frb.device_count() >= 1, "no CUDA device visible: -m gpu tests need a B200"
    frb.init(0)
    yield frb
    frb.init(0)  # forget the device lists of these tests


CASES = {
    "C2_1500_streamed_shards": configs.c2().with_(width=1500, height=1200),
    "C2_ragged_height": configs.c2().with_(width=1111, height=1003, iterations=777),   # partial last band, odd row bytes
    "C1_tile": configs.c1(),
    "C3_julia": configs.c3().with_(width=2048, height=1100),
    "tiny": configs.c1().with_(width=50, height=40, iterations=100),
    "short": configs.c1().with_(width=300, height=16),                                 # one band: single-device path
}


def render_all(cfg, fmt, pinned):
    n = cfg.width * cfg.height * fmt
    buf = frb.pinned_empty(n) if pinned else np.empty(n, np.uint8)
    buf[:] = 0x5A
    frb.get_image_into(cfg, buf, fmt, device=-1)
    return buf.copy(), frb.last_stats()


@pytest.mark.parametrize("shards", [2, 3, 5])
@pytest.mark.parametrize("name", sorted(CASES))
def test_one_device_listed_several_times_tiles_to_the_single_device_image(gpu, name, shards):
    cfg = CASES[name]
    for fmt in (frb.FRB_RGB24, frb.FRB_RGBA32):
        ref = frb.render_shard(cfg, frb.Shard(), fmt).reshape(-1)
        ref_iters = frb.last_stats()["iterations"]
        frb.init_list([0] * shards)
        try:
            for pinned in (False, True):
                got, st = render_all(cfg, fmt, pinned)
                assert np.array_equal(got, ref), f"{name} fmt {fmt} shards {shards} pinned {pinned}: {np.count_nonzero(got != ref)} bytes differ"
                assert st["iterations"] == ref_iters
                assert st["devices"] == (shards if cfg.height > 16 else 1)
        finally:
            frb.init(0)


def test_large_shards_go_out_in_slabs(gpu):
    """>= 4 Mpx per shard: with a pageable destination the shards leave in row slabs under the iterations."""
    cfg = configs.c2().with_(width=3000, height=2900, iterations=300)
    ref = frb.render_shard(cfg, frb.Shard(), frb.FRB_RGB24).reshape(-1)
    frb.init_list([0, 0])
    try:
        got, st = render_all(cfg, frb.FRB_RGB24, pinned=False)
        assert np.array_equal(got, ref)
        got, st = render_all(cfg, frb.FRB_RGB24, pinned=True)
        assert np.array_equal(got, ref)
    finally:
        frb.init(0)


def test_render_bands_assembles_one_frame_shard_by_shard(gpu):
    """frb_render_bands: what each rank of a one-process-per-GPU job calls — its shard into its place in the shared frame."""
    for name in ("C2_ragged_height", "C3_julia", "C1_tile"):
        cfg = CASES[name]
        for fmt in (frb.FRB_RGB24, frb.FRB_RGBA32):
            ref = frb.render_shard(cfg, frb.Shard(), fmt).reshape(-1)
            n = cfg.width * cfg.height * fmt
            for pinned in (False, True):
                for count, band in ((4, 16), (3, 32)):
                    buf = frb.pinned_empty(n) if pinned else np.empty(n, np.uint8)
                    buf[:] = 0x33
                    for i in range(count):
                        frb.render_bands(cfg, frb.Shard(i, count, band), buf, fmt, 0)
                    assert np.array_equal(buf, ref), (name, fmt, pinned, count, band)


def test_two_threads_in_device_minus_one_renders_at_once(gpu):
    """The GUI's render thread and its screenshot thread (src/gui.rs:60, :325) may both be inside get_image on all devices.  The
    per-device workers take one job from one caller at a time: the second caller must neither overwrite the first one's job nor
    return with bands unrendered (it runs its shards on threads of its own).  One device listed three times on a 1-GPU box."""
    import threading

    scenes = [CASES["C2_ragged_height"], CASES["C3_julia"], CASES["C1_tile"], CASES["tiny"]]
    refs = [frb.render_shard(c, frb.Shard(), frb.FRB_RGB24).reshape(-1).copy() for c in scenes]
    frb.init_list([0, 0, 0])
    errors = []

    def caller(tid, rounds):
        try:
            for r in range(rounds):
                k = (tid + r) % len(scenes)
                cfg = scenes[k]
                buf = np.full(cfg.width * cfg.height * 3, 0x5A, np.uint8)
                frb.get_image_into(cfg, buf, frb.FRB_RGB24, device=-1)
                if not np.array_equal(buf, refs[k]):
                    errors.append(f"thread {tid} round {r} scene {k}: {np.count_nonzero(buf != refs[k])} bytes differ")
        except Exception as e:  # noqa: BLE001 - reported below
            errors.append(f"thread {tid}: {e!r}")

    def animation():
        try:
            frames = [scenes[3]] * 40
            out = frb.render_frames(frames, frb.FRB_RGB24, device=-1)
            n = refs[3].size
            for j in range(len(frames)):
                if not np.array_equal(out[j * n:(j + 1) * n], refs[3]):
                    errors.append(f"animation frame {j} differs")
        except Exception as e:  # noqa: BLE001
            errors.append(f"animation: {e!r}")

    try:
        threads = [threading.Thread(target=caller, args=(t, 12)) for t in range(3)] + [threading.Thread(target=animation)]
        for t in threads:
            t.start()
        for t in threads:
            t.join(timeout=300)
            assert not t.is_alive(), "a concurrent device = -1 render hung"
        assert not errors, errors[:5]
    finally:
        frb.init(0)


def test_real_devices_tile_to_the_single_device_image(gpu):
    if frb.device_count() < 2:
        pytest.skip("one device visible: the real multi-device run happens in bench.py's pre-flight at N > 1 and in scripts/inprocess_scale.py")
    for name, cfg in CASES.items():
        ref = frb.render_shard(cfg, frb.Shard(), frb.FRB_RGB24).reshape(-1)
        for pinned in (False, True):
            got, st = render_all(cfg, frb.FRB_RGB24, pinned)
            assert np.array_equal(got, ref), (name, pinned)
            assert st["devices"] == (frb.device_count() if cfg.height > 16 else 1)


def test_first_frames_of_a_fresh_process_over_all_devices():
    """The failure class of round 1's 8-GPU log: device = -1 frames into pinned memory as the FIRST frames of a process (C3-,
    C1-, C2-shaped, in that order), when no kernel has been loaded on any device yet.  One device listed twice on a 1-GPU box."""
    code = r'''
import sys, time
sys.path.insert(0, ".")
import numpy as np
import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs
n = frb.device_count()
assert n >= 1
if n == 1:
    frb.init_list([0, 0])
for name, cfg in (("C3", configs.c3().with_(width=2048, height=2048)), ("C1", configs.c1()), ("C2", configs.c2().with_(width=1500, height=1504))):
    buf = frb.pinned_empty(cfg.width * cfg.height * 3)
    t = time.perf_counter()
    frb.get_image_into(cfg, buf, frb.FRB_RGB24, device=-1)
    dt = time.perf_counter() - t
    st = frb.last_stats()
    assert st["devices"] == max(n, 2), (name, st)
    assert dt < 3.5, (name, dt)
    ref = frb.render_shard(cfg, frb.Shard(), frb.FRB_RGB24).reshape(-1)
    assert np.array_equal(buf, ref), name
print("ok")
'''
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout + r.stderr


def test_forced_kernel_variant_travels_to_the_device_workers(gpu):
    frb.init_list([0, 0, 0])
    try:
        frb.set_kernel_variant(frb.KERNEL_SIMPLE)
        cfg = configs.c1().with_(width=640, height=480, iterations=500)
        buf = np.empty(cfg.width * cfg.height * 3, np.uint8)
        frb.get_image_into(cfg, buf, frb.FRB_RGB24, device=-1)
        assert frb.last_stats()["kernel_variant"] == frb.KERNEL_SIMPLE
    finally:
        frb.set_kernel_variant(frb.KERNEL_AUTO)
        frb.init(0)
