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
for rep in range(2):
    frb.render_device_async(cfg, dev.data_ptr(), n, 3); frb.sync_stats()
t = (ctypes.c_ulonglong * 16)(); L.frb_debug_stream_times(0, t); t0 = t[8]
trips = (ctypes.c_uint * 16384)(); smid = (ctypes.c_uint * 16384)(); ex = (ctypes.c_ulonglong * 16384)()
L.frb_debug_warp_trace(trips, smid, ex)
nw = int(sys.argv[2]) if len(sys.argv) > 2 else 148 * 3 * 8
tr = np.array(trips[:nw]); sm = np.array(smid[:nw]); e = (np.array(ex[:nw], dtype=np.int64) - t0) / 1e6
print("warps", nw, "trips: min/median/max", tr.min(), np.median(tr), tr.max(), " exit ms: min/median/max", e.min(), np.median(e), e.max())
print("trip histogram:", np.histogram(tr, bins=10))
print("CTAs per SM histogram:", np.bincount(np.bincount(sm) // 8))
# within an SM: spread of trips
for s in (0, 1, 50, 147):
    m = sm == s
    print("SM", s, "warps", m.sum(), "trips", np.sort(tr[m]), "exit", np.round(np.sort(e[m]), 2))
w = tr.reshape(-1, 8)
print("per-warp-slot (warp index within CTA) mean trips:", w.mean(0))
