"""Generates tests/golden/small_cases.npz from the CPU oracle (oracle/calc_oracle.c).

The reference has no tests or golden vectors for this path and cannot be built here (Rust toolchain absent), so
these vectors are the oracle's own output, frozen: they protect against silent drift of the oracle and let the GPU
parity tests check against committed data as well as against the live oracle.  Per case: `iters` (u32), `zbits`
(u64 view of the final position, 2 per pixel), `rgb` (u8), `total` (loop-body executions)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cases  # noqa: E402
import oracle_binding as ob  # noqa: E402

DST = os.path.join(HERE, "small_cases.npz")

if __name__ == "__main__":
    out = {}
    for name, cfg in cases.all_small_cases().items():
        r = ob.render_rows(cfg, threads=1)
        out[name + "/iters"] = r["iters"]
        out[name + "/zbits"] = r["z"].view(np.uint64)
        out[name + "/rgb"] = r["rgb"]
        out[name + "/total"] = np.uint64(r["total"])
    np.savez_compressed(DST, **out)
    print("wrote", DST, os.path.getsize(DST), "bytes,", len(cases.all_small_cases()), "cases")
