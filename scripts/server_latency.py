"""Round trip of the headless frame server (frb_server_*: request -> render thread -> RGBA frame polled by the caller), the GUI
loop of src/gui.rs:37-117 without the window.  usage (GPU box): python scripts/server_latency.py"""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs
from fractal_renderer_b200.interactive import FrameServer, ViewState

frb.init()
views = {
    "C1 1920x1080 it256 (default view)": configs.c1(),
    "1920x1080 it1000 julia (C3 view)": configs.c3().with_(width=1920, height=1080),
    "1920x1080 it4000 (C2 view)": configs.c2().with_(width=1920, height=1080),
    "1920x1080 it10000 (C5 frame 120)": configs.c5_frame(120),
}
with FrameServer(device=0) as srv:
    for name, cfg in views.items():
        view = ViewState(cfg)
        for _ in range(3):
            srv.request(view.config); srv.poll(10000)
        n = 50
        for copy in (True, False):   # frb_server_poll into a fresh array / frb_server_acquire (the frame itself)
            lat = []
            for k in range(n):
                view.apply_scroll(1.0 if k % 2 else -1.0)   # a zoom step each frame, like a scroll-wheel drag
                t = time.perf_counter()
                srv.request(view.config)
                frame = srv.poll(10000, copy=copy)
                lat.append(time.perf_counter() - t)
                assert frame is not None
            lat = np.array(lat) * 1e3
            print(f"{name} [{'poll' if copy else 'acquire'}]: median {np.median(lat):.3f} ms  p95 {np.percentile(lat, 95):.3f} ms  -> {1e3 / np.median(lat):.0f} frames/s", flush=True)
