"""Debug: globaltimer stamps of the streamed path (needs the -DFRB_STREAM_TRACE build: FRB_LIB_PATH=.../libfrb_trace.so)."""
import ctypes, sys
import numpy as np, torch
sys.path.insert(0, ".")
import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs
frb.init(1)
L = frb._lib.load()
wl = sys.argv[1] if len(sys.argv) > 1 else "C2"
cfg = {"C1": configs.c1(), "C2": configs.c2(), "C3": configs.c3(), "C4": configs.c4()}[wl]
n = cfg.width * cfg.height * 3
host = torch.empty(n, dtype=torch.uint8).pin_memory().numpy()
dev = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
for dest in ("device", "host"):
    for rep in range(3):
        if dest == "device":
            frb.render_device_async(cfg, dev.data_ptr(), n, 3); frb.sync_stats()
        else:
            frb.get_image_into(cfg, host, 3, 0)
        t = (ctypes.c_ulonglong * 16)()
        L.frb_debug_stream_times(0, t)
        t0 = t[8]
        f = lambda v: (v - t0) / 1e6
        print(f"{wl} {dest}: finisher start {f(t[3]):.3f} first slab go {f(t[4]):.3f} last slab go {f(t[7]):.3f} finisher end {f(t[5]):.3f} | iterate end {f(t[6]):.3f} ms", flush=True)
