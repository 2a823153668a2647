"""Zoom-sequence driver and lossless image writers."""
import os

import numpy as np
import pytest

from fractal_renderer_b200 import configs, image_io, zoom
from fractal_renderer_b200.calc import Imaginary


def test_zoom_sequence_geometric_and_reproduces_c5():
    seq = zoom.zoom_sequence(configs.c5_frame(0), 1.0, 1e12, 240)
    assert len(seq) == 240 and seq[0].scale.re == 1.0
    ratios = [seq[k + 1].scale.re / seq[k].scale.re for k in range(239)]
    assert max(ratios) / min(ratios) < 1 + 1e-9
    for k in (0, 1, 17, 120, 239):  # the frozen C5 definition (configs.c5_scale owns the exact bits)
        assert abs(seq[k].scale.re / configs.c5_scale(k) - 1.0) < 1e-12 and seq[k].pos == configs.DEEP_CENTRE


def test_zoom_sequence_moving_centre_stays_on_screen():
    base = configs.c1()
    a, b = Imaginary(-0.6, 0.0), Imaginary(-0.7436447860, 0.1318252536)
    seq = zoom.zoom_sequence(base, 0.4, 5e5, 50, a, b)
    assert seq[0].pos == a and abs(seq[-1].pos.re - b.re) < 1e-15 and abs(seq[-1].pos.im - b.im) < 1e-15
    # target's screen offset (distance * scale) shrinks monotonically
    off = [abs(c.pos.re - b.re) * c.scale.re for c in seq]
    assert all(y <= x + 1e-12 for x, y in zip(off, off[1:]))


def test_keyframes_and_supersample_and_magnification():
    ks = zoom.keyframe_configs()
    assert len(ks) == 5 and ks[-1].scale.re == 500000.0 and ks[-1].inside is False and ks[0].iterations == 5000
    assert zoom.supersampled(ks[0]).width == 1500
    assert abs(zoom.magnification(configs.c2()) - 1.25e6) < 1.0


def test_ppm_png_roundtrip(tmp_path):
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    p = os.path.join(tmp_path, "a.ppm")
    image_io.write_ppm(p, img.reshape(-1, 3), 53, 37)  # the Vec<RGB> layout get_image returns
    assert np.array_equal(image_io.read_ppm(p), img)
    from PIL import Image

    q = os.path.join(tmp_path, "a.png")
    image_io.write_png(q, img)
    assert np.array_equal(np.asarray(Image.open(q)), img)
    rgba = np.dstack([img, np.full((37, 53), 255, np.uint8)])
    image_io.write_png(q, rgba)
    assert np.array_equal(np.asarray(Image.open(q)), rgba)


@pytest.mark.gpu
def test_render_sequence_matches_single_frames(gpu):
    seq = zoom.zoom_sequence(configs.c1().with_(width=160, height=90, iterations=300), 0.4, 4000.0, 5,
                             Imaginary(-0.6, 0.0), Imaginary(-0.7436447860, 0.1318252536))
    frames = zoom.render_sequence(seq, device=-1)
    for cfg, fr in zip(seq, frames):
        assert np.array_equal(fr, gpu.get_image(cfg).reshape(cfg.height, cfg.width, 3))
