"""Zoom-sequence driver: key-framed (centre, scale) paths rendered as independent frames
through frb_render_frames (frame k on GPU k mod G).  Generalises BASELINE config 5 and the command lines of the
reference's examples.md:20-29 (the Wikipedia zoom gallery)."""
from __future__ import annotations

import math
from typing import Optional, Sequence

import numpy as np

from . import FRB_RGB24, render_frames
from .calc import Algo, Config, Imaginary

# examples.md:24-29 — (scale, x, y, iterations, exposure, inside) of the reference's walk through the Wikipedia gallery
WIKIPEDIA_ZOOM_KEYFRAMES = (
    (400.0, -0.7435669, 0.1314023, 5000, 10.0, True),
    (2000.0, -0.74364990, 0.13188204, 800, 5.0, True),
    (12000.0, -0.74364085, 0.13182733, 5000, 1.0, True),
    (100000.0, -0.743643135, 0.131825963, 2000, 3.0, False),
    (500000.0, -0.7436447860, 0.1318252536, 4000, 5.0, False),
)


def keyframe_configs(width: int = 750, height: int = 500) -> list[Config]:
    """The five stills of examples.md:24-29 as Configs (CLI defaults for everything the command lines do not set)."""
    return [Config.cli(Algo.Mandelbrot, width=width, height=height, scale=Imaginary(s, s), pos=Imaginary(x, y), iterations=it,
                       exposure=e, inside=inside) for (s, x, y, it, e, inside) in WIKIPEDIA_ZOOM_KEYFRAMES]


def zoom_sequence(base: Config, scale_from: float, scale_to: float, frames: int,
                  centre_from: Optional[Imaginary] = None, centre_to: Optional[Imaginary] = None) -> list[Config]:
    """`frames` configs whose scale runs geometrically from scale_from to scale_to (equal zoom factor per frame) while
    the centre moves linearly in *screen* units, i.e. proportionally to 1/scale, so the target stays put on screen.
    The scales are computed once here and travel as bits (pow is not correctly rounded across libms)."""
    if frames < 1:
        return []
    c0 = centre_from if centre_from is not None else base.pos
    c1 = centre_to if centre_to is not None else c0
    out = []
    for k in range(frames):
        t = k / (frames - 1) if frames > 1 else 0.0
        s = float(scale_from * (scale_to / scale_from) ** t)
        # weight of the start centre decays with the zoom: w = (1/s - 1/s1) / (1/s0 - 1/s1)
        if scale_to != scale_from:
            w = (1.0 / s - 1.0 / scale_to) / (1.0 / scale_from - 1.0 / scale_to)
        else:
            w = 1.0 - t
        pos = Imaginary(c1.re + (c0.re - c1.re) * w, c1.im + (c0.im - c1.im) * w)
        out.append(base.with_(scale=Imaginary(s, s), pos=pos))
    return out


def supersampled(config: Config, factor: int = 2) -> Config:
    """The GUI's screenshot path renders at 2x width and height (src/gui.rs:321-323); any integer factor here."""
    return config.with_(width=config.width * factor, height=config.height * factor)


def render_sequence(cfgs: Sequence[Config], fmt: int = FRB_RGB24, device: int = -1) -> list[np.ndarray]:
    """Renders the frames on the GPUs and returns them as (H, W, fmt) arrays (views into one buffer)."""
    flat = render_frames(list(cfgs), fmt, device)
    frames, off = [], 0
    for c in cfgs:
        n = c.width * c.height * fmt
        frames.append(flat[off:off + n].reshape(c.height, c.width, fmt))
        off += n
    return frames


def magnification(config: Config) -> float:
    """Zoom relative to the CLI's default view (scale 0.4): the README's "1,000,000x" is scale 5e5 / 0.4 = 1.25e6."""
    return math.sqrt(abs(config.scale.re * config.scale.im)) / 0.4
