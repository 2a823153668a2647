"""Deferred kernel CTA shape against the end of the frame: whole C2 frames and 1/8 shards, frames issued back to back (as bench.py's
device-resident arm does).  usage (GPU box): FRB_LIB_PATH=fractal_renderer_b200/libfrb_tune.so python scripts/r2_shape_tail.py"""
import sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs
frb.init(1)
cfg = configs.c2()
for ns in (8, 1):
    shard = frb.Shard(0, ns, 16) if ns > 1 else frb.Shard()
    n = shard.rows(cfg.height) * cfg.width * 3
    dev = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
    ref = None
    for shape, name in ((2, "768x1"), (1, "512x1"), (3, "1024x1"), (4, "512x2"), (0, "256x3")):
        for pool_max in (31, 0):
            frb.set_tuning("defer_shape", shape); frb.set_tuning("pool", 1 if pool_max else 0); frb.set_tuning("pool_max", pool_max)
            for _ in range(5):
                frb.render_device_async(cfg, dev.data_ptr(), n, 3, shard=shard)
            frb.sync_stats()
            reps = 40 if ns > 1 else 12
            torch.cuda.synchronize(); t0 = time.perf_counter()
            for _ in range(reps):
                frb.render_device_async(cfg, dev.data_ptr(), n, 3, shard=shard)
            st = frb.sync_stats(); torch.cuda.synchronize()
            ms = (time.perf_counter() - t0) * 1e3 / reps
            img = dev[:n].cpu().numpy()
            if ref is None:
                ref = img
            print(f"C2 1/{ns} shape {name:7s} pool_max {pool_max:2d}: {ms:.4f} ms/frame back to back  iterate(ev) {st['iterate_ms']:.4f}  same {bool(np.array_equal(img, ref))}", flush=True)
