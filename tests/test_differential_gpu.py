"""Randomised differential tests on the GPU: whatever kernel the library picks (deferred, dense, 7-instruction, tile,
streamed, sliced) against the one-thread-per-pixel kernel in reference order (FRB_KERNEL_SIMPLE, itself pinned to the CPU
oracle by tests/test_parity_gpu.py), bit for bit on counts and positions and byte for byte on images, over seeded random
configurations.  Exists because hand-picked cases missed a whole class (caps that are not a multiple of the deferred
kernel's block); random caps, limits, sizes, shards and constants do not share the author's blind spots."""
import math

import numpy as np
import pytest

import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs
from fractal_renderer_b200.calc import RGB, Algo, Config, Imaginary

import oracle_binding as ob

pytestmark = pytest.mark.gpu

# FRB_DIFF_SOAK=k multiplies the number of seeds (a one-off soak: FRB_DIFF_SOAK=20 is ~5000 configurations in a few minutes)
import os
SOAK = max(1, int(os.environ.get("FRB_DIFF_SOAK", "1")))


@pytest.fixture(scope="module")
def gpu():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    yield frb
    frb.set_kernel_variant(frb.KERNEL_AUTO)


def random_config(rng: np.random.Generator, big: bool, huge: bool = False) -> Config:
    algo = Algo.Julia if rng.random() < 0.45 else Algo.Mandelbrot
    if huge:
        w, h = int(rng.integers(2049, 2700)), int(rng.integers(2049, 2500))
    elif big:
        w, h = int(rng.integers(300, 1400)), int(rng.integers(260, 900))
    else:
        w, h = int(rng.integers(1, 260)), int(rng.integers(1, 200))
    r = rng.random()
    if r < 0.25:
        it = int(rng.integers(0, 40))
    elif r < 0.75:
        it = int(rng.integers(40, 700))
    else:
        it = int(math.exp(rng.uniform(math.log(700), math.log(6000))))
    limit = float(rng.choice([2.0, 3.0, 7.9, 8.0, 10.0, 100.0, 65536.0, 65536.0, 65536.0, 1e10, 1e100, 1e150]))
    if rng.random() < 0.15:
        limit = float(math.exp(rng.uniform(math.log(0.5), math.log(1e6))))
    zoom = math.exp(rng.uniform(math.log(0.3), math.log(3e4)))
    if algo == Algo.Mandelbrot:
        # aim at the boundary so that orbit lengths are mixed: points near the main cardioid / period-2 bulb
        t = rng.uniform(0, 2 * math.pi)
        cr, ci = 0.5 * math.cos(t) - 0.25 * math.cos(2 * t), 0.5 * math.sin(t) - 0.25 * math.sin(2 * t)
        pos = Imaginary(cr + rng.normal(0, 0.02), ci + rng.normal(0, 0.02)) if rng.random() < 0.7 else Imaginary(rng.uniform(-2, 1), rng.uniform(-1.2, 1.2))
        julia = Imaginary(0.0, 0.0)
    else:
        pos = Imaginary(rng.normal(0, 0.4), rng.normal(0, 0.4))
        rad = rng.uniform(0, 1.1)
        t = rng.uniform(0, 2 * math.pi)
        julia = Imaginary(rad * math.cos(t), rad * math.sin(t))
        if rng.random() < 0.08:
            julia = Imaginary(float(rng.choice([1e3, -1e5, 3e9])), rng.normal())  # |c| beyond the sparse filter's precondition
        zoom = math.exp(rng.uniform(math.log(0.3), math.log(50)))
    return Config.cli(algo, width=w, height=h, iterations=it, julia_set=julia).with_(
        limit=limit, stable_limit=float(rng.choice([2.0, 2.0, 0.5, 4.0, 1e9])), pos=pos,
        scale=Imaginary(zoom, zoom * (1.0 if rng.random() < 0.8 else rng.uniform(0.5, 2))),
        exposure=float(rng.choice([2.0, 2.0, 1.0, 0.3, 11.0])), smooth=bool(rng.random() < 0.8), inside=bool(rng.random() < 0.8),
        primary_color=RGB(*(int(v) for v in rng.integers(0, 256, 3))), secondary_color=RGB(*(int(v) for v in rng.integers(0, 256, 3))))


def random_shard(rng: np.random.Generator, gpu):
    if rng.random() < 0.5:
        return gpu.Shard()
    count = int(rng.integers(1, 4))
    return gpu.Shard(int(rng.integers(0, count)), count, int(rng.choice([0, 1, 8, 16, 40])))


def check(gpu, cfg: Config, shard, fmt: int, tag: str, extra_destinations: bool = False) -> None:
    gpu.set_kernel_variant(gpu.KERNEL_SIMPLE)
    try:
        ni, nz = gpu.render_counts(cfg, shard)
        ref_img = gpu.render_shard(cfg, shard, fmt)
    finally:
        gpu.set_kernel_variant(gpu.KERNEL_AUTO)
    di, dz = gpu.render_counts(cfg, shard)
    v_raw = gpu.last_stats()["kernel_variant"]
    img = gpu.render_shard(cfg, shard, fmt)
    st = gpu.last_stats()
    what = f"{tag}: {cfg} shard=({shard.index},{shard.count},{shard.band_rows}) raw kernel {v_raw}, image kernel {st['kernel_variant']}"
    assert np.array_equal(di, ni), what
    assert np.array_equal(dz.view(np.uint64), nz.view(np.uint64)), what
    assert np.array_equal(img, ref_img), what
    if extra_destinations and ni.size:
        # pinned host buffer (streamed beside the iterations when the frame qualifies) and device buffer
        import torch
        rows = ni.shape[0]
        buf = torch.empty(rows * cfg.width * fmt, dtype=torch.uint8).pin_memory().numpy()
        buf[:] = 0x5A
        got = gpu.render_shard(cfg, shard, fmt, out=buf.reshape(rows, cfg.width, fmt))
        assert np.array_equal(got, ref_img), what + f" [pinned destination, kernel {gpu.last_stats()['kernel_variant']}]"
        dev = torch.full((buf.size + 64,), 0x5A, dtype=torch.uint8, device="cuda")
        gpu.render_device_async(cfg, dev.data_ptr(), buf.size, fmt, device=0, shard=shard)
        gpu.sync_stats()
        assert np.array_equal(dev[:buf.size].cpu().numpy().reshape(ref_img.shape), ref_img), what + " [device destination]"
        assert bool((dev[buf.size:] == 0x5A).all()), what + " [device destination overrun]"
    if cfg.iterations > 0 and ni.size:
        assert st["iterations"] == int((np.minimum(ni.astype(np.int64), cfg.iterations - 1) + 1).sum()), what


@pytest.mark.parametrize("seed", range(16 * SOAK))
def test_random_small_frames_every_kernel_agrees(gpu, seed):
    rng = np.random.default_rng(1000 + seed)
    for k in range(12):
        cfg = random_config(rng, big=False)
        check(gpu, cfg, random_shard(rng, gpu), gpu.FRB_RGB24 if rng.random() < 0.5 else gpu.FRB_RGBA32, f"seed {seed} case {k}")


@pytest.mark.parametrize("seed", range(10 * SOAK))
def test_random_large_frames_every_kernel_agrees(gpu, seed):
    """Sizes at which the sliced and tile-slab host paths and multi-CTA queues are in play."""
    rng = np.random.default_rng(2000 + seed)
    for k in range(4):
        cfg = random_config(rng, big=True)
        check(gpu, cfg, random_shard(rng, gpu), gpu.FRB_RGB24 if rng.random() < 0.5 else gpu.FRB_RGBA32, f"seed {seed} case {k}",
              extra_destinations=True)


def test_random_small_frames_against_the_cpu_oracle(gpu):
    """The same generator against the oracle itself (counts and positions bit-exact, colour within 1 LSB)."""
    rng = np.random.default_rng(77)
    for k in range(25):
        cfg = random_config(rng, big=False)
        if cfg.width * cfg.height * max(cfg.iterations, 1) > 400_000_000:
            continue
        ref = ob.render_rows(cfg, threads=1)
        di, dz = gpu.render_counts(cfg)
        assert np.array_equal(di.reshape(-1), ref["iters"].reshape(-1)), (k, cfg)
        assert np.array_equal(dz.view(np.uint64).reshape(-1), ref["z"].view(np.uint64).reshape(-1)), (k, cfg)
        img = gpu.render_shard(cfg, gpu.Shard(), gpu.FRB_RGB24)
        d = np.abs(img.astype(np.int16).reshape(-1) - ref["rgb"].astype(np.int16).reshape(-1))
        assert d.max(initial=0) <= 1, (k, cfg)


@pytest.mark.parametrize("seed", range(3 * SOAK))
def test_random_frames_above_4_megapixels_every_kernel_agrees(gpu, seed):
    """Above 4 Mpx a pageable destination goes out in row slabs (iterate + shade + copy per slab on two streams) and the
    tile kernel's grid spans hundreds of CTA rows; the pinned destination is streamed."""
    rng = np.random.default_rng(3000 + seed)
    for k in range(2):
        cfg = random_config(rng, big=True, huge=True)
        cfg = cfg.with_(iterations=min(cfg.iterations, 700))
        check(gpu, cfg, random_shard(rng, gpu) if k else gpu.Shard(), gpu.FRB_RGB24 if rng.random() < 0.5 else gpu.FRB_RGBA32,
              f"seed {seed} case {k}", extra_destinations=True)


@pytest.mark.parametrize("seed", range(4 * SOAK))
def test_random_frame_sequences_equal_frame_by_frame_renders(gpu, seed):
    """frb_render_frames (double-buffered, alternating streams; the zoom driver's call) on random frames of different sizes
    and kernels back to back, against one blocking reference-order render per frame."""
    rng = np.random.default_rng(4000 + seed)
    fmt = gpu.FRB_RGB24 if seed % 2 else gpu.FRB_RGBA32
    frames = [random_config(rng, big=bool(k % 3 == 0)) for k in range(7)]
    got = gpu.render_frames(frames, fmt, device=0)
    off = 0
    gpu.set_kernel_variant(gpu.KERNEL_SIMPLE)
    try:
        for k, cfg in enumerate(frames):
            ref = gpu.render_shard(cfg, gpu.Shard(), fmt).reshape(-1)
            assert np.array_equal(got[off:off + ref.size], ref), f"seed {seed} frame {k}: {cfg}"
            off += ref.size
    finally:
        gpu.set_kernel_variant(gpu.KERNEL_AUTO)
    assert off == got.size
