"""Tile kernel on C1: shape of a warp's tile and warps per CTA (tuning build) -> device time per frame (device destination)
and end-to-end time into a pinned host buffer (the kernel's stores go straight over PCIe there).
usage (GPU box): FRB_LIB_PATH=fractal_renderer_b200/libfrb_tune.so python scripts/r2_tile_shapes.py"""
import sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs
frb.init(1)
cfg = configs.c1()
n = cfg.width * cfg.height * 3
dev = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
pinned = frb.pinned_empty(n)
for tw, wpc in ((0, 4), (8, 1), (8, 2), (8, 4), (8, 8), (16, 2), (16, 4), (32, 2), (32, 4)):
    frb.set_tuning("tile_w", tw); frb.set_tuning("tile_wpc", wpc)
    it, dv, e2e = [], [], []
    for k in range(60):
        frb.render_device_async(cfg, dev.data_ptr(), n, 3)
        st = frb.sync_stats()
        it.append(st["iterate_ms"]); dv.append(st["device_ms"])
    util = st["iterations"] / st["lane_iterations"]
    for k in range(40):
        t = time.perf_counter(); frb.get_image_into(cfg, pinned, 3, 0); e2e.append((time.perf_counter() - t) * 1e3)
    print(f"tile {tw:2d}x{32//tw if tw else 0} wpc {wpc}: iterate(ev) median {np.median(it[10:])*1e3:.1f} us min {min(it[10:])*1e3:.1f}  device {np.median(dv[10:])*1e3:.1f} us  "
          f"util {util:.3f}  pinned e2e median {np.median(e2e[10:]):.3f} ms kernel", frb.KERNEL_NAMES[st["kernel_variant"]], flush=True)
