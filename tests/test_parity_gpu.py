"""Parity of the CUDA path (through the C ABI) with the CPU oracle.  Run on a B200: `pytest -m gpu`.

Bar: escape counts and final-z bits BIT-EXACT; 8-bit colour channels within +-1 LSB (only log2,
which the reference takes from the platform libm, can differ).  NaN final positions compare as a class (x86 and
CUDA produce different NaN sign/payload bits; the reference defines neither)."""
import os
import threading

import numpy as np
import pytest

import cases
import oracle_binding as ob
from fractal_renderer_b200 import configs

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
COLOUR_TOL = 1  # LSB, stated by BASELINE.json north_star


def z_equal(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """bitwise equality, NaN == NaN as a class"""
    same = a.view(np.uint64) == b.view(np.uint64)
    return same | (np.isnan(a) & np.isnan(b))


def assert_parity(frb, cfg, ref, shard=None, variant=None, what=""):
    frb.set_kernel_variant(variant if variant is not None else frb.KERNEL_AUTO)
    try:
        iters, z = frb.render_counts(cfg, shard=shard)
        stats = frb.last_stats()
        sh = shard if shard is not None else frb.Shard()
        rgb = frb.render_shard(cfg, sh, frb.FRB_RGB24)
        rgba = frb.render_shard(cfg, sh, frb.FRB_RGBA32)
    finally:
        frb.set_kernel_variant(frb.KERNEL_AUTO)
    assert np.array_equal(iters, ref["iters"]), f"{what}: {np.count_nonzero(iters != ref['iters'])} escape counts differ"
    zq = z_equal(z, ref["z"])
    assert zq.all(), f"{what}: {np.count_nonzero(~zq)} final-z components differ in bits"
    d = np.abs(rgb.astype(np.int16) - ref["rgb"].astype(np.int16))
    assert d.max(initial=0) <= COLOUR_TOL, f"{what}: colour off by {d.max()} LSB"
    assert np.array_equal(rgba[..., :3], rgb) and (rgba[..., 3] == 255).all(), f"{what}: RGBA32 != RGB24 + alpha"
    if cfg.algo != 1:
        assert stats["iterations"] == ref["total"], f"{what}: iteration total {stats['iterations']} != {ref['total']}"
    return int(np.count_nonzero(d)), stats


SMALL = cases.all_small_cases()


@pytest.mark.parametrize("name", sorted(SMALL))
def test_small_cases_vs_oracle_and_golden(gpu, name):
    cfg = SMALL[name]
    ref = ob.render_rows(cfg, threads=1)
    g = np.load(os.path.join(GOLDEN, "small_cases.npz"))
    assert np.array_equal(ref["iters"], g[name + "/iters"])  # the live oracle is the frozen oracle
    golden = dict(iters=g[name + "/iters"], z=g[name + "/zbits"].view(np.float64), rgb=g[name + "/rgb"], total=int(g[name + "/total"]))
    assert_parity(gpu, cfg, golden, what=name + "/auto")


@pytest.mark.parametrize("variant", ["defer6", "dense6", "stream7", "simple"])
@pytest.mark.parametrize("name", ["mandel_default", "mandel_iter0", "mandel_iter9", "mandel_limit1p5", "mandel_limit1e155",
                                  "mandel_limit_nan", "mandel_real_axis", "mandel_deep_subulp", "julia_default",
                                  "julia_c3_view", "C1_small", "C2_small", "C4_small"])
def test_every_kernel_variant(gpu, name, variant):
    """Forcing each kernel: STREAM7 and SIMPLE are exact for any input; DEFER6 and DENSE6 are downgraded where they are not."""
    cfg = SMALL[name]
    v = {"defer6": gpu.KERNEL_DEFER6, "dense6": gpu.KERNEL_DENSE6, "stream7": gpu.KERNEL_STREAM7, "simple": gpu.KERNEL_SIMPLE}[variant]
    ref = ob.render_rows(cfg, threads=1)
    _, stats = assert_parity(gpu, cfg, ref, variant=v, what=f"{name}/{variant}")
    # a forced kernel is downgraded (DEFER6 -> DENSE6 -> STREAM7) outside its proven domain, never run unsafely
    chain = {gpu.KERNEL_DEFER6: (gpu.KERNEL_DEFER6, gpu.KERNEL_DENSE6, gpu.KERNEL_STREAM7),
             gpu.KERNEL_DENSE6: (gpu.KERNEL_DENSE6, gpu.KERNEL_STREAM7)}.get(v, (v,))
    assert stats["kernel_variant"] in chain
    if name in ("mandel_default", "C2_small", "C4_small", "julia_default"):
        assert stats["kernel_variant"] == v


def test_auto_routes_exotic_inputs_to_form7(gpu):
    for name, expect in [("mandel_default", gpu.KERNEL_DEFER6), ("julia_default", gpu.KERNEL_DEFER6),
                         ("julia_zero", gpu.KERNEL_STREAM7), ("julia_real_c", gpu.KERNEL_STREAM7),
                         ("julia_tiny_im", gpu.KERNEL_STREAM7), ("mandel_limit1e155", gpu.KERNEL_STREAM7),
                         ("mandel_limit_nan", gpu.KERNEL_STREAM7), ("mandel_huge_coords", gpu.KERNEL_STREAM7),
                         ("mandel_real_axis", gpu.KERNEL_DEFER6), ("mandel_limit2", gpu.KERNEL_DENSE6),
                         ("mandel_limit1p5", gpu.KERNEL_DENSE6), ("mandel_iter7", gpu.KERNEL_DENSE6),
                         ("mandel_far_out", gpu.KERNEL_DEFER6), ("C4_small", gpu.KERNEL_DEFER6),
                         ("julia_limit_2p20_bigc", gpu.KERNEL_DEFER6), ("julia_limit1e8_bigc", gpu.KERNEL_DENSE6)]:
        gpu.render_counts(SMALL[name])
        assert gpu.last_stats()["kernel_variant"] == expect, name


def test_c1_full_frame(gpu):
    """C1 = the config that 'runs on the reference CPU path': every pixel of the 1920x1080 frame."""
    cfg = configs.c1()
    ref = ob.render_rows(cfg)
    assert ref["total"] == configs.TOTAL_ITERATIONS["C1"]
    ndiff, _ = assert_parity(gpu, cfg, ref, what="C1")
    assert ndiff <= cfg.width * cfg.height // 10000  # colours are expected identical; allow log2 ulp ties


def test_c3_full_frame(gpu):
    cfg = configs.c3()
    ref = ob.render_rows(cfg)
    assert ref["total"] == configs.TOTAL_ITERATIONS["C3"]
    assert_parity(gpu, cfg, ref, what="C3")


def test_c2_full_frame(gpu):
    """The north-star frame, all 9 M pixels (the oracle needs a few seconds on the box's cores)."""
    cfg = configs.c2()
    ref = ob.render_rows(cfg)
    assert ref["total"] == configs.TOTAL_ITERATIONS["C2"]
    assert_parity(gpu, cfg, ref, what="C2")


def test_c4_shards_vs_oracle(gpu):
    """C4 is 268 Mpx / 5.7e11 iterations: check three of 128 interleaved-band shards (128 rows each) bit for bit —
    the same partition the 8-GPU run uses (16-row bands)."""
    cfg = configs.c4()
    for idx in (0, 77, 127):
        sh = gpu.Shard(idx, 128, 16)
        rows = sh.image_rows(cfg.height)
        assert len(rows) == 128
        ref = ob.render_rows(cfg, rows=rows)
        assert_parity(gpu, cfg, ref, shard=sh, what=f"C4 shard {idx}/128")


def test_c4_whole_frame_total_and_the_eight_way_partition(gpu):
    """All 268 Mpx of C4.  (i) The survey's independent exact total (574 496 116 940 loop-body executions) is a
    size-independent checksum of every pixel: one whole-frame render must report it.  (ii) The partition the 8-GPU run actually
    uses — Shard(r, 8, 16), 2048 rows each — rendered shard by shard: the shards' totals add up to the same number, each equals the
    sum over its own counts, and on 32 sampled rows of EVERY shard (256 rows, spread over the frame) counts and final-z bits equal
    the oracle's."""
    torch = pytest.importorskip("torch")
    cfg = configs.c4()
    want_total = configs.TOTAL_ITERATIONS["C4"]
    assert want_total == 574_496_116_940
    n = cfg.width * cfg.height * 3
    dev = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
    gpu.render_device_async(cfg, dev.data_ptr(), n, gpu.FRB_RGB24, device=0)
    st = gpu.sync_stats()
    assert st["iterations"] == want_total and st["pixels"] == cfg.width * cfg.height
    whole_rows = dev[:n].view(cfg.height, cfg.width * 3)
    total = 0
    for r in range(8):
        sh = gpu.Shard(r, 8, 16)
        iters, z = gpu.render_counts(cfg, shard=sh)
        s = gpu.last_stats()
        assert iters.shape == (2048, cfg.width)
        assert s["iterations"] == int((np.minimum(iters, cfg.iterations - 1).astype(np.uint64) + 1).sum()), f"shard {r}: total != sum of its counts"
        total += s["iterations"]
        local = np.arange(3 + 5 * r, 2048, 64, dtype=np.int64)            # 32 local rows, a different phase per shard
        image_rows = sh.image_rows(cfg.height)[local].astype(np.uint32)
        ref = ob.render_rows(cfg, rows=image_rows)
        assert np.array_equal(iters[local], ref["iters"]), f"shard {r}/8: counts differ"
        assert z_equal(z[local], ref["z"]).all(), f"shard {r}/8: final positions differ"
        # and the whole-frame image holds the same rows, within the colour tolerance of the oracle
        got = whole_rows[torch.as_tensor(image_rows.astype(np.int64), device="cuda")].cpu().numpy().reshape(len(local), cfg.width, 3)
        assert np.abs(got.astype(int) - ref["rgb"].astype(int)).max() <= COLOUR_TOL, f"shard {r}/8: colour"
        del iters, z
    assert total == want_total


def test_c5_three_whole_frames_at_true_size(gpu):
    """Config 5, every row of the first, the middle and the last frame of the zoom (1920x1080 each) against the oracle."""
    for k in (0, 120, 239):
        cfg = configs.c5_frame(k)
        ref = ob.render_rows(cfg)
        assert_parity(gpu, cfg, ref, what=f"C5 frame {k}")


def test_streamed_frames_under_a_co_tenant_animation(gpu):
    """The streamed path's bounded device-side waits under real contention (src/gui.rs:60 + :325: render thread and screenshot
    thread on one device): while frb_render_frames keeps two expensive C5 frames in flight on device 0, a second thread renders
    a 1.8 Mpx frame into PINNED memory 60 times — the destination the streaming finisher writes while the iterate kernel runs.
    Every image must equal the quiet render; how often the finisher had to fall back (one more launch) is printed."""
    import time
    anim = [configs.c5_frame(150 + (k % 40)) for k in range(140)]
    cfg = configs.c2().with_(width=1500, height=1200)
    n = cfg.width * cfg.height * 3
    buf = gpu.pinned_empty(n)
    gpu.get_image_into(cfg, buf, gpu.FRB_RGB24, device=0)
    want = buf.copy()
    quiet_launches = gpu.last_stats()["kernel_launches"]
    errors = []

    def animate():
        try:
            gpu.render_frames(anim, gpu.FRB_RGB24, device=0)
        except Exception as e:  # noqa: BLE001 - reported below
            errors.append(repr(e))

    th = threading.Thread(target=animate)
    th.start()
    time.sleep(0.02)
    done = during = fallbacks = 0
    while done < 60:
        buf[:] = 0x5A
        gpu.get_image_into(cfg, buf, gpu.FRB_RGB24, device=0)
        assert np.array_equal(buf, want), f"frame {done} rendered beside the animation differs"
        fallbacks += gpu.last_stats()["kernel_launches"] != quiet_launches
        done += 1
        during += th.is_alive()
    th.join()
    assert not errors, errors
    print(f"co-tenant: {during} of {done} pinned frames completed while the animation ran, {fallbacks} took the fallback path")
    assert during >= 10, f"only {during} frames overlapped the animation"


def test_c5_frames(gpu):
    """Three frames of the zoom through frb_render_frames (the animation entry point), against the oracle."""
    ks = [0, 100, 239]
    cfgs = [configs.c5_frame(k).with_(width=480, height=270) for k in ks]
    out = gpu.render_frames(cfgs, gpu.FRB_RGB24, device=0)
    total = 0
    off = 0
    for cfg in cfgs:
        ref = ob.render_rows(cfg)
        n = cfg.width * cfg.height * 3
        got = out[off:off + n].reshape(cfg.height, cfg.width, 3)
        off += n
        assert np.abs(got.astype(int) - ref["rgb"].astype(int)).max() <= COLOUR_TOL
        total += ref["total"]
    assert gpu.last_stats()["iterations"] == total


def test_c5_frames_at_true_size(gpu):
    """Config 5 as BASELINE.json states it: 1920x1080 (the coordinate map divides by the height, so a shrunken frame is a
    different input at 1e12x), ten frames spread over k = 0 .. 239 through frb_render_frames — all devices, dynamic frame queue —
    against the oracle on every 40th row: counts and final positions bit-exact (frb_render_counts), image within +-1 LSB,
    and the animation's frames byte-identical to the same frames rendered one by one."""
    ks = [0, 13, 57, 101, 120, 160, 199, 222, 238, 239]
    cfgs = [configs.c5_frame(k) for k in ks]
    h, w = cfgs[0].height, cfgs[0].width
    out = gpu.render_frames(cfgs, gpu.FRB_RGB24, device=-1)
    st = gpu.last_stats()
    rows = np.arange(7, h, 40, dtype=np.uint32)
    n = w * h * 3
    total = 0
    for j, cfg in enumerate(cfgs):
        got = out[j * n:(j + 1) * n].reshape(h, w, 3)
        ref = ob.render_rows(cfg, rows=rows)
        iters, z = gpu.render_counts(cfg)
        total += gpu.last_stats()["iterations"]
        assert np.array_equal(iters[rows], ref["iters"]), f"frame {ks[j]}: counts differ"
        assert z_equal(z[rows], ref["z"]).all(), f"frame {ks[j]}: final positions differ"
        assert np.abs(got[rows].astype(int) - ref["rgb"].astype(int)).max() <= COLOUR_TOL, f"frame {ks[j]}"
        single = gpu.get_image(cfg).reshape(h, w, 3)
        assert np.array_equal(single, got), f"frame {ks[j]}: animation frame != single frame"
    assert st["iterations"] == total and st["pixels"] == len(ks) * w * h


def test_render_frames_leaves_room_for_other_callers(gpu):
    """frb_render_frames locks a device per frame, not per call: a get_image issued from another thread while an animation is
    running (the GUI's screenshot thread, src/gui.rs:325) completes long before the animation does."""
    import threading
    import time
    cfgs = [configs.c5_frame(150 + (k % 40)) for k in range(160)]
    small = configs.c1().with_(width=320, height=200)
    want = gpu.get_image(small)
    t_anim = {}

    def animate():
        t0 = time.perf_counter()
        gpu.render_frames(cfgs, gpu.FRB_RGB24, device=0)
        t_anim["s"] = time.perf_counter() - t0

    th = threading.Thread(target=animate)
    th.start()
    time.sleep(0.02)
    during = 0
    while th.is_alive():
        got = gpu.get_image(small)
        assert np.array_equal(got, want)
        if th.is_alive():
            during += 1   # this call started AND finished while the animation was still running
    th.join()
    assert during >= 3, f"{during} concurrent frames completed during an animation of {t_anim['s']:.3f} s"


def test_render_frames_strided_tiles_the_animation(gpu):
    """frb_render_frames_strided: what each rank of a one-process-per-GPU job calls; the ranks' frames land in one buffer."""
    cfgs = [configs.c5_frame(k).with_(width=320, height=180) for k in range(0, 240, 17)]
    whole = gpu.render_frames(cfgs, gpu.FRB_RGB24, device=0)
    out = np.zeros_like(whole)
    for r in range(3):
        gpu.render_frames(cfgs, gpu.FRB_RGB24, device=0, out=out, first=r, stride=3)
    assert np.array_equal(out, whole)


def test_full_size_properties_c2_variants_agree(gpu):
    """At full size, the three kernels must agree with each other bit for bit (SIMPLE is the reference order as
    written, on the GPU) and hit the exact iteration total."""
    cfg = configs.c2()
    res = {}
    for v in (gpu.KERNEL_DEFER6, gpu.KERNEL_DENSE6, gpu.KERNEL_STREAM7, gpu.KERNEL_SIMPLE):
        gpu.set_kernel_variant(v)
        try:
            res[v] = gpu.render_counts(cfg)
            assert gpu.last_stats()["iterations"] == configs.TOTAL_ITERATIONS["C2"]
        finally:
            gpu.set_kernel_variant(gpu.KERNEL_AUTO)
    base = res[gpu.KERNEL_SIMPLE]
    for v in (gpu.KERNEL_DEFER6, gpu.KERNEL_DENSE6, gpu.KERNEL_STREAM7):
        assert np.array_equal(res[v][0], base[0])
        assert z_equal(res[v][1], base[1]).all()


def test_shards_tile_the_frame(gpu):
    cfg = configs.c1().with_(width=333, height=250)
    full = gpu.get_image(cfg).reshape(cfg.height, cfg.width, 3)
    for count, br in [(2, 16), (3, 7), (8, 16), (8, 1)]:
        img = np.zeros_like(full)
        for i in range(count):
            sh = gpu.Shard(i, count, br)
            img[sh.image_rows(cfg.height)] = gpu.render_shard(cfg, sh)
        assert np.array_equal(img, full), (count, br)


def test_render_all_devices_matches_single(gpu):
    cfg = configs.c1().with_(width=640, height=403)
    one = gpu.get_image(cfg, device=0)
    allv = gpu.get_image(cfg, device=-1)
    assert np.array_equal(one, allv)


def test_idempotent_and_reentrant(gpu):
    """Two host threads rendering at once (the GUI's render thread and its screenshot thread may overlap:
    src/gui.rs:60 and :325) must both get the right frame."""
    a = configs.c1().with_(width=400, height=300)
    b = configs.c3().with_(width=256, height=256)
    ra, rb = gpu.get_image(a), gpu.get_image(b)
    out = {}

    def work(key, cfg, n):
        for _ in range(n):
            out[key] = gpu.get_image(cfg)

    ta = threading.Thread(target=work, args=("a", a, 6))
    tb = threading.Thread(target=work, args=("b", b, 6))
    ta.start(); tb.start(); ta.join(); tb.join()
    assert np.array_equal(out["a"], ra) and np.array_equal(out["b"], rb)


def test_invalid_arguments(gpu):
    cfg = configs.c1().with_(width=16, height=16)
    small = np.empty(10, np.uint8)
    with pytest.raises(gpu.FrbError) as e:
        gpu.get_image_into(cfg, small)
    assert e.value.code == -2
    with pytest.raises(gpu.FrbError):
        gpu.render_shard(cfg, gpu.Shard(3, 2, 16))
    with pytest.raises(gpu.FrbError):
        gpu.get_image(cfg, device=99)
    empty = gpu.get_image(cfg.with_(width=0))
    assert empty.size == 0


# ---- the lock-step tile kernel (csrc/frb_tile.cuh): image formats with short iteration caps -----------------------------------

@pytest.mark.parametrize("name", sorted(SMALL))
def test_tile_kernel_small_cases(gpu, name):
    """Forced on every small case (any limit, any iteration cap: it tests after every iteration, and takes the 7-instruction
    form where the 6-instruction one is not provably exact): colour against the oracle, iteration total exact."""
    cfg = SMALL[name]
    ref = ob.render_rows(cfg, threads=1)
    gpu.set_kernel_variant(gpu.KERNEL_TILE)
    try:
        rgb = gpu.render_shard(cfg, gpu.Shard(), gpu.FRB_RGB24)
        st = gpu.last_stats()
        rgba = gpu.render_shard(cfg, gpu.Shard(), gpu.FRB_RGBA32)
    finally:
        gpu.set_kernel_variant(gpu.KERNEL_AUTO)
    d = np.abs(rgb.astype(np.int16) - ref["rgb"].reshape(rgb.shape).astype(np.int16))
    assert d.max(initial=0) <= COLOUR_TOL, f"{name}: colour off by {d.max()} LSB"
    assert np.array_equal(rgba[..., :3], rgb) and (rgba[..., 3] == 255).all()
    if cfg.algo != 1 and cfg.width * cfg.height > 0:
        assert st["kernel_variant"] == gpu.KERNEL_TILE
        assert st["iterations"] == ref["total"], name
        assert st["kernel_launches"] == 2  # tables + the tile kernel: no records, no shade kernel


def test_tile_kernel_is_the_default_for_short_caps_and_matches_the_refill_kernel(gpu):
    """C1 (cap 256) takes the tile kernel by default; its image is byte-identical to the deferred kernel's (same arithmetic,
    same libdevice log2) and within 1 LSB of the oracle; longer caps keep the deferred kernel."""
    cfg = configs.c1()
    auto = gpu.render_shard(cfg, gpu.Shard(), gpu.FRB_RGB24)
    st = gpu.last_stats()
    assert st["kernel_variant"] == gpu.KERNEL_TILE
    gpu.set_kernel_variant(gpu.KERNEL_DEFER6)
    try:
        other = gpu.render_shard(cfg, gpu.Shard(), gpu.FRB_RGB24)
        st2 = gpu.last_stats()
    finally:
        gpu.set_kernel_variant(gpu.KERNEL_AUTO)
    assert st2["kernel_variant"] == gpu.KERNEL_DEFER6 and st2["iterations"] == st["iterations"] == 85_781_804
    assert np.array_equal(auto, other)
    ref = ob.render_rows(cfg)
    assert np.abs(auto.astype(np.int16) - ref["rgb"].reshape(auto.shape).astype(np.int16)).max() <= COLOUR_TOL
    gpu.render_shard(configs.c1().with_(iterations=385), gpu.Shard(), gpu.FRB_RGB24)
    assert gpu.last_stats()["kernel_variant"] == gpu.KERNEL_DEFER6
    # shards tile, ragged sizes
    cfg = configs.c1().with_(width=1003, height=517)
    whole = gpu.render_shard(cfg, gpu.Shard(), gpu.FRB_RGB24)
    out = np.zeros_like(whole)
    for i in range(3):
        sh = gpu.Shard(i, 3, 16)
        out[sh.image_rows(cfg.height)] = gpu.render_shard(cfg, sh, gpu.FRB_RGB24)
    assert np.array_equal(out, whole)


# ---- the colour fast path (table log2, corrected reciprocal division) against the libdevice path ------------------------------

FAST_SHADE_CASES = r"""
cases = [
    ("C1", configs.c1()),
    ("C1_it600", configs.c1().with_(iterations=600)),
    ("julia", configs.c3().with_(width=1200, height=900)),
    ("big_limit", configs.c1().with_(width=800, height=600, limit=1.0e4, iterations=300)),
    ("small_limit", configs.c1().with_(width=800, height=600, limit=1.2, stable_limit=1.2, iterations=200)),
    ("exposure", configs.c1().with_(width=800, height=600, exposure=37.5)),
]
"""


def test_fast_shade_matches_libdevice_shade(gpu):
    """Smooth colouring: the table log2 differs from libdevice's in the last bit or two, which moves a byte only when a
    channel sits within ~1e-13 of an integer — allow 1 LSB on at most a handful of bytes per frame (measured: none).
    The libdevice path is only selectable in the tuning build (frb_set_tuning), hence the subprocess."""
    from tune_run import run_with_tuning_build
    run_with_tuning_build(FAST_SHADE_CASES + r"""
for name, cfg in cases:
    frb.set_tuning("fast_shade", 1)
    fast = frb.render_shard(cfg, frb.Shard(), frb.FRB_RGB24)
    frb.set_tuning("fast_shade", 0)
    slow = frb.render_shard(cfg, frb.Shard(), frb.FRB_RGB24)
    d = np.abs(fast.astype(np.int16) - slow.astype(np.int16))
    assert d.max() <= 1, name
    assert int((d != 0).sum()) <= 4, f"{name}: {int((d != 0).sum())} bytes differ"
""")


def test_fast_shade_is_exact_without_smoothing(gpu):
    """smooth = false: no logarithm, and the corrected reciprocal division is correctly rounded, so the fast path must be
    byte-identical to the oracle's division — including the structured cases where count/iterations*exposure*channel lands
    exactly on an integer — and to the libdevice path (tuning build, subprocess)."""
    cases = ((50, 2.0), (255, 1.0), (100, 2.55), (256, 3.0), (1000, 10.0), (77, 0.7))
    for it, exposure in cases:
        cfg = configs.c1().with_(width=640, height=480, iterations=it, exposure=exposure, smooth=False)
        fast = gpu.render_shard(cfg, gpu.Shard(), gpu.FRB_RGBA32)
        ref = ob.render_rows(cfg, threads=1)
        assert np.array_equal(fast[..., :3].reshape(-1), ref["rgb"].reshape(-1)), (it, exposure)
    from tune_run import run_with_tuning_build
    run_with_tuning_build(f"cases = {cases!r}" + r"""
for it, exposure in cases:
    cfg = configs.c1().with_(width=640, height=480, iterations=it, exposure=exposure, smooth=False)
    frb.set_tuning("fast_shade", 1)
    fast = frb.render_shard(cfg, frb.Shard(), frb.FRB_RGBA32)
    frb.set_tuning("fast_shade", 0)
    slow = frb.render_shard(cfg, frb.Shard(), frb.FRB_RGBA32)
    assert np.array_equal(fast, slow), (it, exposure)
""")


def test_tile_kernel_shapes_agree(gpu):
    """The tile kernel's other shapes (tuning build: 8x4, 16x2, 32x1 pixels per warp, 1-8 warps per CTA) produce the bytes and
    the statistics of the product's shape on ragged sizes (width not a multiple of 4 or of the tile, few rows), both formats,
    sharded or not."""
    from tune_run import run_with_tuning_build
    run_with_tuning_build(r"""
frb.set_kernel_variant(frb.KERNEL_TILE)
cases = [configs.c1().with_(width=640, height=360), configs.c1().with_(width=333, height=77, iterations=100),
         configs.c3().with_(width=516, height=130, iterations=200), configs.c1().with_(width=8, height=3, iterations=60)]
for cfg in cases:
    for shard in (frb.Shard(), frb.Shard(1, 3, 16)):
        for fmt in (frb.FRB_RGB24, frb.FRB_RGBA32):
            ref = None
            for tw, wpc in ((8, 4), (8, 1), (8, 8), (16, 2), (32, 1), (32, 4)):
                frb.set_tuning("tile_w", tw); frb.set_tuning("tile_wpc", wpc)
                img = frb.render_shard(cfg, shard, fmt)
                st = frb.last_stats()
                assert st["kernel_variant"] == frb.KERNEL_TILE
                if ref is None:
                    ref, ref_it = img, st["iterations"]
                assert np.array_equal(img, ref), (cfg.width, cfg.height, tw, wpc, fmt)
                assert st["iterations"] == ref_it
""")


def test_tile_kernel_backs_off_on_scenes_with_poor_lane_utilisation(gpu):
    """Julia dust with a short cap: orbit lengths inside an 8x4 tile differ wildly, the lock-step kernel reports its lane
    utilisation, and the library keeps the lane-refill kernel for the following frames of that scene (same bytes either way)."""
    gpu.set_kernel_variant(gpu.KERNEL_AUTO)  # forgets earlier feedback
    cfg = configs.c3().with_(width=1024, height=768, iterations=300)
    first = gpu.render_shard(cfg, gpu.Shard(), gpu.FRB_RGB24)
    st = gpu.last_stats()
    assert st["kernel_variant"] == gpu.KERNEL_TILE
    util = st["iterations"] / st["lane_iterations"]
    assert util < 0.7, util
    second = gpu.render_shard(cfg.with_(scale=cfg.scale), gpu.Shard(), gpu.FRB_RGB24)
    assert gpu.last_stats()["kernel_variant"] == gpu.KERNEL_DEFER6
    assert np.array_equal(first, second)
    # another scene is not affected; C1 keeps the tile kernel (utilisation 0.93)
    gpu.render_shard(configs.c1(), gpu.Shard(), gpu.FRB_RGB24)
    st = gpu.last_stats()
    # (0.93 with 8x4 tiles; host destinations use 32x1 tiles, 96 contiguous bytes per tile row on the PCIe link: 0.884)
    assert st["kernel_variant"] == gpu.KERNEL_TILE and st["iterations"] / st["lane_iterations"] > 0.85
    gpu.render_shard(configs.c1(), gpu.Shard(), gpu.FRB_RGB24)
    assert gpu.last_stats()["kernel_variant"] == gpu.KERNEL_TILE
    gpu.set_kernel_variant(gpu.KERNEL_AUTO)
    gpu.render_shard(cfg, gpu.Shard(), gpu.FRB_RGB24)
    assert gpu.last_stats()["kernel_variant"] == gpu.KERNEL_TILE  # feedback forgotten
    gpu.set_kernel_variant(gpu.KERNEL_AUTO)


@pytest.mark.parametrize("cap", [9, 15, 50, 257, 300, 333, 511, 2049, 2063, 4095])
def test_deferred_kernel_caps_that_are_not_a_multiple_of_the_block(gpu, cap):
    """Regression: a pixel that runs into the iteration cap is recorded at the start of its last full block and needs
    B + (cap mod B) resolve steps; with B + 1 the last steps were dropped for every cap with cap mod B >= 2 (final z of
    bounded orbits wrong, escapes in the last iterations missed).  Counts and positions against the one-thread-per-pixel
    kernel, bit for bit, at a size where thousands of pixels reach the cap."""
    for base in (configs.c1(), configs.c3()):
        cfg = base.with_(width=512, height=384, iterations=cap)
        gpu.set_kernel_variant(gpu.KERNEL_SIMPLE)
        try:
            ni, nz = gpu.render_counts(cfg)
            gpu.set_kernel_variant(gpu.KERNEL_DEFER6)
            di, dz = gpu.render_counts(cfg)
            # below two blocks the library substitutes the dense kernel (frb_api.cu: make_plan)
            assert gpu.last_stats()["kernel_variant"] == (gpu.KERNEL_DEFER6 if cap >= 16 else gpu.KERNEL_DENSE6)
            img = gpu.render_shard(cfg, gpu.Shard(), gpu.FRB_RGB24)
            gpu.set_kernel_variant(gpu.KERNEL_SIMPLE)
            ref_img = gpu.render_shard(cfg, gpu.Shard(), gpu.FRB_RGB24)
        finally:
            gpu.set_kernel_variant(gpu.KERNEL_AUTO)
        if base.algo == configs.c1().algo:
            assert (ni == cap).sum() > 1000  # the Mandelbrot interior; the Julia set here is dust (a handful of pixels)
        assert np.array_equal(di, ni), cap
        assert np.array_equal(dz.view(np.uint64), nz.view(np.uint64)), cap
        assert np.array_equal(img, ref_img), cap
