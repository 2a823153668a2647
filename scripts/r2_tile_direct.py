"""C1 (tile kernel) into a pinned buffer: direct stores over PCIe vs slabs + copies.  Tuning build."""
import sys, time
import numpy as np
sys.path.insert(0, ".")
import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs
frb.init(1)
for name, cfg in (("C1", configs.c1()), ("C1 rgba", configs.c1()), ("C1 4k", configs.c1().with_(width=3840, height=2160))):
    fmt = 4 if "rgba" in name else 3
    buf = frb.pinned_empty(cfg.width * cfg.height * fmt)
    ref = None
    for direct in (0, 1):
        frb.set_tuning("tile_direct", direct)
        lat = []
        for k in range(60):
            t = time.perf_counter(); frb.get_image_into(cfg, buf, fmt, 0); lat.append((time.perf_counter() - t) * 1e3)
        if ref is None: ref = buf.copy()
        print(name, "direct" if direct else "slabs ", f"median {np.median(lat[10:]):.4f} min {min(lat[10:]):.4f} ms same={bool(np.array_equal(buf, ref))} launches {frb.last_stats()['kernel_launches']}", flush=True)
