"""Tile kernel on C1: device time per frame against the CTAs resident per SM (capped through unused dynamic shared memory).
usage (GPU box): FRB_LIB_PATH=fractal_renderer_b200/libfrb_tune.so python scripts/r2_tile_occ.py"""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs
frb.init(1)
cfg = configs.c1()
n = cfg.width * cfg.height * 3
dev = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
for kb in (0, 24, 28, 32, 36, 44):
    frb.set_tuning("tile_smem_kb", kb)
    it, dv = [], []
    for k in range(60):
        frb.render_device_async(cfg, dev.data_ptr(), n, 3)
        st = frb.sync_stats()
        it.append(st["iterate_ms"]); dv.append(st["device_ms"])
    print(f"tile_smem_kb {kb:2d}: iterate(ev) median {np.median(it[10:])*1e3:.1f} us min {min(it[10:])*1e3:.1f}  device {np.median(dv[10:])*1e3:.1f} us kernel", frb.KERNEL_NAMES[st["kernel_variant"]], flush=True)
