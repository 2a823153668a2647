"""Runs a test body against the TUNING build of the library (libfrb_tune.so, -DFRB_TUNING) in a fresh process: the product
library has no tuning hooks (frb_set_tuning rejects every knob, and it reads no environment variable), so the few tests that
must force a rarely taken path — the libdevice colour path, a finisher that never sees its frame — go through this."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

PRELUDE = r"""
import sys
sys.path.insert(0, {root!r})
sys.path.insert(0, {tests!r})
import numpy as np
import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs
assert frb._lib.LIB_PATH.endswith("libfrb_tune.so")
frb.init(0)
"""


def run_with_tuning_build(body: str, timeout: int = 300) -> str:
    lib = os.path.join(ROOT, "fractal_renderer_b200", "libfrb_tune.so")
    assert os.path.exists(lib), "libfrb_tune.so not built (make -C fractal_renderer_b200/csrc)"
    env = dict(os.environ, FRB_LIB_PATH=lib)
    code = PRELUDE.format(root=ROOT, tests=os.path.join(ROOT, "tests")) + body + '\nprint("tuning-run ok")\n'
    r = subprocess.run([sys.executable, "-c", code], cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0 and "tuning-run ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
    return r.stdout
