"""Where does the tail of a frame go?  Per-SM time of the last warp exit relative to the queue-drained stamp (trace build)."""
import ctypes, sys
import numpy as np, torch
sys.path.insert(0, ".")
import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs
frb.init(1)
L = frb._lib.load()
cfg = configs.c2()
ns = int(sys.argv[1]) if len(sys.argv) > 1 else 8
shard = frb.Shard(0, ns, 16) if ns > 1 else frb.Shard()
n = shard.rows(cfg.height) * cfg.width * 3
dev = torch.empty(n + 64, dtype=torch.uint8, device="cuda")
for pool_max, end_chunk in ((0, 32), (0, 4), (16, 4), (31, 4)):
    frb.set_tuning("pool", 1 if pool_max else 0); frb.set_tuning("pool_max", pool_max); frb.set_tuning("end_chunk", end_chunk)
    for rep in range(3):
        frb.render_device_async(cfg, dev.data_ptr(), n, 3, shard=shard); st = frb.sync_stats()
    t = (ctypes.c_ulonglong * 16)(); L.frb_debug_stream_times(0, t)
    trips = (ctypes.c_uint * 16384)(); smid = (ctypes.c_uint * 16384)(); ex = (ctypes.c_ulonglong * 16384)()
    L.frb_debug_warp_trace(trips, smid, ex)
    nw = 148 * 24
    e = (np.array(ex[:nw], dtype=np.int64) - t[9]) / 1e3   # us after the queue ran dry
    sm = np.array(smid[:nw])
    w = np.arange(nw) % 24
    per_sm = np.array([e[sm == s].max() for s in np.unique(sm)])
    per_smsp = np.array([e[(sm == s) & (w % 4 == q)].max() for s in np.unique(sm) for q in range(4)])
    print(f"pool {pool_max} chunk {end_chunk}: device_ms {st['device_ms']:.3f}  last exit per SM (us after drain): p0 {per_sm.min():.0f} p25 {np.percentile(per_sm,25):.0f} p50 {np.median(per_sm):.0f} "
          f"p75 {np.percentile(per_sm,75):.0f} p90 {np.percentile(per_sm,90):.0f} max {per_sm.max():.0f} | per SMSP: p50 {np.median(per_smsp):.0f} p90 {np.percentile(per_smsp,90):.0f} max {per_smsp.max():.0f} "
          f"| warps: p50 {np.median(e):.0f} mean {e.mean():.0f}", flush=True)
