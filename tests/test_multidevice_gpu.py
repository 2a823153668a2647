"""frb_render(device = -1): one frame over every device of the box — interleaved 16-row bands, one host worker per device,
bands written (streamed finisher, pinned destination) or copied (strided 2-D copies, slab by slab for large shards) into ONE
caller-owned image.  On a single-GPU box FRB_TUNE_FAKE_DEVICES lists the device several times, so the same code runs shard
after shard; on a multi-GPU box the real devices are used as well.  The image must equal the single-device render byte for byte."""
import os

import numpy as np
import pytest

import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def gpu():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    yield frb
    os.environ.pop("FRB_TUNE_FAKE_DEVICES", None)


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
    buf = torch.empty(n, dtype=torch.uint8).pin_memory().numpy() if pinned else np.empty(n, np.uint8)
    buf[:] = 0x5A
    frb.get_image_into(cfg, buf, fmt, device=-1)
    return buf.copy(), frb.last_stats()


@pytest.mark.parametrize("fake", [2, 3, 5])
@pytest.mark.parametrize("name", sorted(CASES))
def test_faked_devices_tile_to_the_single_device_image(gpu, name, fake):
    cfg = CASES[name]
    for fmt in (frb.FRB_RGB24, frb.FRB_RGBA32):
        os.environ.pop("FRB_TUNE_FAKE_DEVICES", None)
        ref = frb.render_shard(cfg, frb.Shard(), fmt).reshape(-1)
        ref_iters = frb.last_stats()["iterations"]
        os.environ["FRB_TUNE_FAKE_DEVICES"] = str(fake)
        try:
            for pinned in (False, True):
                got, st = render_all(cfg, fmt, pinned)
                assert np.array_equal(got, ref), f"{name} fmt {fmt} fake {fake} pinned {pinned}: {np.count_nonzero(got != ref)} bytes differ"
                assert st["iterations"] == ref_iters
                assert st["devices"] == (fake if cfg.height > 16 else 1)
        finally:
            os.environ.pop("FRB_TUNE_FAKE_DEVICES", None)


def test_faked_devices_large_shards_go_out_in_slabs(gpu):
    """>= 4 Mpx per shard with a pageable destination: iterate + shade + strided copy per row slab."""
    cfg = configs.c2().with_(width=3000, height=2900, iterations=300)
    ref = frb.render_shard(cfg, frb.Shard(), frb.FRB_RGB24).reshape(-1)
    os.environ["FRB_TUNE_FAKE_DEVICES"] = "2"
    try:
        got, st = render_all(cfg, frb.FRB_RGB24, pinned=False)
        assert np.array_equal(got, ref)
        got, st = render_all(cfg, frb.FRB_RGB24, pinned=True)
        assert np.array_equal(got, ref)
    finally:
        os.environ.pop("FRB_TUNE_FAKE_DEVICES", None)


def test_real_devices_tile_to_the_single_device_image(gpu):
    if frb.device_count() < 2:
        pytest.skip("one device")
    for name, cfg in CASES.items():
        ref = frb.render_shard(cfg, frb.Shard(), frb.FRB_RGB24).reshape(-1)
        for pinned in (False, True):
            got, st = render_all(cfg, frb.FRB_RGB24, pinned)
            assert np.array_equal(got, ref), (name, pinned)
            assert st["devices"] == (frb.device_count() if cfg.height > 16 else 1)


def test_forced_kernel_variant_travels_to_the_device_workers(gpu):
    os.environ["FRB_TUNE_FAKE_DEVICES"] = "3"
    try:
        frb.set_kernel_variant(frb.KERNEL_SIMPLE)
        cfg = configs.c1().with_(width=640, height=480, iterations=500)
        buf = np.empty(cfg.width * cfg.height * 3, np.uint8)
        frb.get_image_into(cfg, buf, frb.FRB_RGB24, device=-1)
        assert frb.last_stats()["kernel_variant"] == frb.KERNEL_SIMPLE
    finally:
        frb.set_kernel_variant(frb.KERNEL_AUTO)
        os.environ.pop("FRB_TUNE_FAKE_DEVICES", None)
