"""Timeline of one frame (or one shard of a frame cut in N) on one GPU, from the -DFRB_STREAM_TRACE build's globaltimer stamps:
where the time between the call and the finished frame goes — start-up, queue drained, last warp out, finisher done — and how
the warps' exit times are distributed (the tail).  usage (GPU box):
    FRB_LIB_PATH=fractal_renderer_b200/libfrb_trace.so python scripts/r2_timeline.py C2 8      # shard 0 of 8
"""
import ctypes
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs

frb.init(1)
L = frb._lib.load()
wl = sys.argv[1] if len(sys.argv) > 1 else "C2"
nshards = int(sys.argv[2]) if len(sys.argv) > 2 else 1
cfg = {"C1": configs.c1(), "C2": configs.c2(), "C3": configs.c3(), "C4": configs.c4()}[wl]
shard = frb.Shard(0, nshards, 16) if nshards > 1 else frb.Shard()
rows = shard.rows(cfg.height)
n = rows * cfg.width * 3
host = torch.empty(n, dtype=torch.uint8).pin_memory().numpy()
dev = torch.empty(n + 64, dtype=torch.uint8, device="cuda")


def stamps():
    t = (ctypes.c_ulonglong * 16)()
    L.frb_debug_stream_times(0, t)
    return list(t)


for dest in ("device", "host"):
    wall, devms, itms = [], [], []
    for rep in range(6):
        t0 = time.perf_counter()
        if dest == "device":
            frb.render_device_async(cfg, dev.data_ptr(), n, 3, shard=shard)
            st = frb.sync_stats()
        else:
            frb.render_shard(cfg, shard, 3, 0, out=host)
            st = frb.last_stats()
        wall.append((time.perf_counter() - t0) * 1e3)
        devms.append(st["device_ms"])
        itms.append(st["iterate_ms"])
    t = stamps()
    k0 = t[8]
    f = lambda v: (v - k0) / 1e6 if v not in (0, 2**64 - 1) else float("nan")
    print(f"{wl} 1/{nshards} {dest}: wall {np.median(wall[2:]):.3f} ms  device(ev) {np.median(devms[2:]):.3f}  iterate(ev) {np.median(itms[2:]):.3f} | "
          f"from K0: first iterate warp {f(t[13]):.3f}  queue drained {f(t[9]):.3f}  last warp out {f(t[6]):.3f}  "
          f"finisher: start {f(t[3]):.3f} first slab {f(t[4]):.3f} last slab {f(t[7]):.3f} end {f(t[5]):.3f}  "
          f"iterations {st['iterations']} lanes util {st['iterations'] / max(st['lane_iterations'], 1):.3f}", flush=True)
    trips = (ctypes.c_uint * 16384)()
    smid = (ctypes.c_uint * 16384)()
    ex = (ctypes.c_ulonglong * 16384)()
    L.frb_debug_warp_trace(trips, smid, ex)
    nw = 148 * 24
    e = (np.array(ex[:nw], dtype=np.int64) - k0) / 1e6
    e = e[(e > 0) & (e < 1e4)]
    if len(e):
        q = np.percentile(e, [0, 10, 25, 50, 75, 90, 99, 100])
        print(f"   warp exit times ms (p0 p10 p25 p50 p75 p90 p99 p100): {np.round(q, 3)}  mean {e.mean():.3f}", flush=True)
