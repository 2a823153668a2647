"""In-process multi-GPU: frb_render(device=-1) — the call a Rust `get_image` shim makes — on 1, 2, 4, 8 devices of one box.
Interleaved 16-row bands, one stream set per device, strided copies into ONE pinned host image; wall clock per blocking call.
usage (GPU box): python scripts/inprocess_scale.py [C2 C4 C1 C3]"""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs

which = sys.argv[1:] or ["C2", "C4"]
cfgs = {"C1": configs.c1(), "C2": configs.c2(), "C3": configs.c3(), "C4": configs.c4()}
ndev = frb.device_count()
out = {}
for name in which:
    cfg = cfgs[name]
    buf = frb.pinned_empty(cfg.width * cfg.height * 3)
    ref = None
    for n in [k for k in (1, 2, 4, 8) if k <= ndev]:
        frb.init_list(list(range(n)))
        steps = 3 if name == "C4" else 30
        for _ in range(2):
            frb.get_image_into(cfg, buf, frb.FRB_RGB24, device=-1)
        t0 = time.perf_counter()
        iters = 0
        for _ in range(steps):
            frb.get_image_into(cfg, buf, frb.FRB_RGB24, device=-1)
            iters += frb.last_stats()["iterations"]
        dt = (time.perf_counter() - t0) / steps
        st = frb.last_stats()
        if ref is None:
            ref = buf.copy()
            base = dt
        same = bool(np.array_equal(buf, ref))
        out[f"{name}_n{n}"] = {"ms_per_frame": round(dt * 1e3, 4), "g_iter_s": round(iters / steps / dt / 1e9, 1), "devices": st["devices"],
                               "efficiency": round(base / (n * dt), 4), "identical_to_n1": same, "kernel": st["kernel_variant"]}
        print(name, "n", n, out[f"{name}_n{n}"], flush=True)
print(json.dumps(out))
