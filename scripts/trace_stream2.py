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
dev = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
buf = (ctypes.c_ulonglong * (3 * 8192))(); cnt = ctypes.c_uint()
for rep in range(3):
    L.frb_debug_stream_trace(buf, ctypes.byref(cnt), 1)
    frb.render_device_async(cfg, dev.data_ptr(), n, 3); frb.sync_stats()
    t = (ctypes.c_ulonglong * 16)(); L.frb_debug_stream_times(0, t); t0 = t[8]
    L.frb_debug_stream_trace(buf, ctypes.byref(cnt), 0)
    k = min(cnt.value, 8192)
    a = np.array(buf[:3 * k], dtype=np.uint64).reshape(k, 3)
    print(f"rep {rep}: {cnt.value} scans; iterate end {(t[6]-t0)/1e6:.3f} ms")
    step = max(1, k // 40)
    for row in a[::step]:
        print(f"   t={(int(row[0])-t0)/1e6:8.3f} ms  min={int(row[1]):>11d} (row {int(row[1])//cfg.width if row[1] < 2**32-1 else -1})  argmin warp {int(row[2])} (cta {int(row[2])//8})")
