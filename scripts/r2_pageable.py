"""End-to-end cost of frb_render by destination kind: pinned (reused), pageable (reused, touched), and what the Rust shim's
get_image does — a FRESH, untouched allocation per frame (Vec::with_capacity: rust/ffi.rs, reference src/lib.rs:253).
usage (GPU box): python scripts/r2_pageable.py [C2 C1 C3] [device]"""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs

which = [a for a in sys.argv[1:] if a.startswith("C")] or ["C2", "C1", "C3"]
device = int([a for a in sys.argv[1:] if not a.startswith("C")][0]) if any(not a.startswith("C") for a in sys.argv[1:]) else 0
frb.init(0)
cfgs = {"C1": configs.c1(), "C2": configs.c2(), "C3": configs.c3(), "C4": configs.c4()}
for name in which:
    cfg = cfgs[name]
    n = cfg.width * cfg.height * 3
    steps = 30 if name != "C4" else 3
    pinned = frb.pinned_empty(n)
    reused = np.empty(n, np.uint8)
    reused[:] = 0
    res = {}
    ref = None
    for kind in ("pinned", "pageable_reused", "pageable_fresh"):
        lat = []
        for k in range(steps + 3):
            if kind == "pinned":
                buf = pinned
            elif kind == "pageable_reused":
                buf = reused
            else:
                buf = np.empty(n, np.uint8)  # fresh mapping for large n (mmap threshold): untouched pages, like Vec::with_capacity
            t = time.perf_counter()
            frb.get_image_into(cfg, buf, frb.FRB_RGB24, device=device)
            lat.append((time.perf_counter() - t) * 1e3)
        if ref is None:
            ref = buf.copy()
        same = bool(np.array_equal(buf, ref))
        res[kind] = (float(np.median(lat[3:])), float(np.min(lat[3:])), same)
    st = frb.last_stats()
    print(name, "device", device, " ".join(f"{k}: median {v[0]:.3f} min {v[1]:.3f} ms same={v[2]}" for k, v in res.items()),
          "| device_ms", round(st["device_ms"], 3), "kernel", frb.KERNEL_NAMES[st["kernel_variant"]], flush=True)
