"""Sweep of the end-of-frame remedies of the deferred kernel (orbit pool threshold x end-phase chunk size) on whole frames and on
1/8 shards, device-resident and streamed into pinned memory; every setting must produce the bytes of the reference setting.
usage (GPU box): FRB_LIB_PATH=fractal_renderer_b200/libfrb_trace.so python scripts/r2_pool_sweep.py [C2:1 C2:8 C3:1 C4:8]"""
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
trace = hasattr(L, "frb_debug_stream_times")
specs = [a for a in sys.argv[1:]] or ["C2:1", "C2:8", "C3:1", "C3:8", "C4:8"]
cfgs = {"C1": configs.c1(), "C2": configs.c2(), "C3": configs.c3(), "C4": configs.c4()}
settings = [(0, 32), (0, 8), (16, 32), (8, 8), (16, 8), (24, 8), (31, 8), (16, 4), (31, 4), (31, 16)]


def stamps():
    t = (ctypes.c_ulonglong * 16)()
    L.frb_debug_stream_times(0, t)
    return list(t)


for spec in specs:
    name, ns = spec.split(":")
    ns = int(ns)
    cfg = cfgs[name]
    shard = frb.Shard(0, ns, 16) if ns > 1 else frb.Shard()
    rows = shard.rows(cfg.height)
    n = rows * cfg.width * 3
    host = frb.pinned_empty(n)
    dev = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
    ref = None
    reps = 3 if name == "C4" else 12
    for pool_max, end_chunk in settings:
        frb.set_tuning("pool", 1 if pool_max else 0)
        frb.set_tuning("pool_max", pool_max)
        frb.set_tuning("end_chunk", end_chunk)
        out = {}
        for dest in ("device", "host"):
            wall, devms, itms = [], [], []
            for rep in range(reps):
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
            img = dev[:n].cpu().numpy() if dest == "device" else host.copy()
            if ref is None:
                ref = img
            same = bool(np.array_equal(img, ref))
            tail = ""
            if trace:
                t = stamps()
                tail = f" drained->last warp {(t[6] - t[9]) / 1e3:.0f} us"
            out[dest] = f"{dest}: wall {np.median(wall[2:]):.3f} dev {np.median(devms[2:]):.3f} iter {np.median(itms[2:]):.3f}{tail} same={same} util {st['iterations'] / max(st['lane_iterations'], 1):.3f}"
        print(f"{name} 1/{ns} pool_max {pool_max:2d} end_chunk {end_chunk:2d} | {out['device']} | {out['host']}", flush=True)
frb.set_tuning("reset", 0)
