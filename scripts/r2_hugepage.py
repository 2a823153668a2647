"""Fresh pageable destinations the way a Rust Vec::with_capacity gets them — malloc() above the mmap threshold: untouched pages, no
huge-page hint from the allocator (numpy gives its own MADV_HUGEPAGE to arrays >= 4 MB) — against the staging pipeline's knobs
(tuning build): huge-page hint, populate ahead of the copies, copy threads.
usage (GPU box): FRB_LIB_PATH=fractal_renderer_b200/libfrb_tune.so python scripts/r2_hugepage.py"""
import ctypes, sys, time
import numpy as np
sys.path.insert(0, ".")
import fractal_renderer_b200 as frb
from fractal_renderer_b200 import configs
libc = ctypes.CDLL("libc.so.6")
libc.malloc.restype = ctypes.c_void_p; libc.malloc.argtypes = [ctypes.c_size_t]; libc.free.argtypes = [ctypes.c_void_p]
libc.mallopt(-3, 1 << 20)  # M_MMAP_THRESHOLD fixed at 1 MB: every frame is a fresh mapping
frb.init(1)
L = frb._lib.load()
for name, cfg in (("C1", configs.c1()), ("C3", configs.c3()), ("C2", configs.c2())):
    n = cfg.width * cfg.height * 3
    c = cfg.to_c() if hasattr(cfg, "to_c") else None
    for hp, pop, ct in ((0, 0, 3), (1, 0, 3), (0, 1, 3), (1, 1, 3), (1, 1, 6), (0, 1, 6)):
        frb.set_tuning("hugepage", hp); frb.set_tuning("populate", pop); frb.set_tuning("copy_threads", ct)
        lat = []
        for k in range(25):
            p = libc.malloc(n)
            buf = np.ctypeslib.as_array((ctypes.c_uint8 * n).from_address(p))
            t = time.perf_counter(); frb.get_image_into(cfg, buf, 3, 0); lat.append((time.perf_counter() - t) * 1e3)
            del buf; libc.free(p)
        print(name, "hugepage", hp, "populate", pop, "copy_threads", ct, f"fresh malloc median {np.median(lat[5:]):.3f} min {min(lat[5:]):.3f} ms", flush=True)
