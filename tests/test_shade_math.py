"""CPU checks of the arithmetic the colour fast path relies on (csrc/frb_kernels.cuh: shade_log2, shade_pixel), restated with
exact rational arithmetic — no GPU, no oracle: the committed log2 table is what the generator produces, the table algorithm
meets its stated error bound, and the corrected-reciprocal division is correctly rounded."""
import math
import os
import re
import subprocess
import sys
from fractions import Fraction

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INC = os.path.join(ROOT, "fractal_renderer_b200", "csrc", "frb_log2_table.inc")


def load_table():
    rows = []
    for line in open(INC):
        m = re.match(r"\s*\{(\S+), (\S+)\}", line)
        if m:
            rows.append((float.fromhex(m.group(1)), float.fromhex(m.group(2))))
    return np.array(rows)


def test_committed_table_is_what_the_generator_writes(tmp_path):
    out = tmp_path / "table.inc"
    subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "gen_log2_table.py"), str(out)], check=True)
    assert open(out).read() == open(INC).read()


def test_table_entries_are_consistent():
    t = load_table()
    assert t.shape == (128, 2)
    for j, (inv_c, log2_c) in enumerate(t):
        c = Fraction(2 * (128 + j) + 1, 256)
        assert inv_c == float(1 / c)                      # RN(1/c), Fraction -> float rounds correctly
        assert abs(log2_c + math.log2(inv_c)) < 3e-16     # -log2 of the ROUNDED reciprocal
        # |r| = |m * inv_c - 1| < 2^-8 over the whole interval the entry serves
        lo, hi = Fraction(128 + j, 128), Fraction(129 + j, 128)
        assert max(abs(lo * Fraction(inv_c) - 1), abs(hi * Fraction(inv_c) - 1)) < Fraction(1, 256) + Fraction(1, 2**50)


def shade_log2_model(x: np.ndarray, t: np.ndarray) -> np.ndarray:
    """shade_log2 as written, with the fma of the reduction done in long double (the polynomial's own fmas only matter below 1e-18)."""
    bits = x.view(np.uint64)
    hi = (bits >> np.uint64(32)).astype(np.uint32)
    m = ((bits & np.uint64(0x000FFFFFFFFFFFFF)) | np.uint64(0x3FF0000000000000)).view(np.float64)
    j = (hi >> 13) & 127
    r = (m.astype(np.longdouble) * t[j, 0].astype(np.longdouble) - 1).astype(np.float64)
    e = (hi >> 20).astype(np.float64) - 1023.0
    a = [1.4426950408889634, -0.7213475204444817, 0.48089834696298783, -0.36067376022224085, 0.28853900817779266, -0.24044917348149392]
    q = np.full_like(r, a[5])
    for k in (4, 3, 2, 1, 0):
        q = r * q + a[k]
    return r * q + (e + t[j, 1])


def test_table_log2_meets_its_error_bound():
    t = load_table()
    rng = np.random.default_rng(7)
    x = np.concatenate([
        np.exp(rng.uniform(np.log(2.0), np.log(1e300), 400_000)),   # dist: [2, huge)
        rng.uniform(0.25, 256.0, 400_000),                          # log_zn: [0.25, 256)
        2.0 ** np.arange(-2, 1000, 7), np.nextafter(2.0 ** np.arange(1, 1000, 7), 0), np.nextafter(2.0 ** np.arange(1, 1000, 7), np.inf),
    ])
    got = shade_log2_model(x, t).astype(np.longdouble)
    ref = np.log2(x.astype(np.longdouble))
    err = np.abs(got - ref) / np.maximum(1, np.abs(ref))
    assert float(err.max()) < 3e-16, float(err.max())


def rn(fr: Fraction) -> float:
    return float(fr)  # correctly rounded


def fast_div(a: float, n: int) -> float:
    """shade_pixel: q = a * rN; q = fma(fma(-q, N, a), rN, q) with rN = RN(1/N)."""
    r_n = rn(Fraction(1, n))
    q = rn(Fraction(a) * Fraction(r_n))
    rem = rn(Fraction(a) - Fraction(q) * n)
    return rn(Fraction(rem) * Fraction(r_n) + Fraction(q))


def test_corrected_reciprocal_division_is_correctly_rounded():
    rng = np.random.default_rng(11)
    cases = []
    for n in (1, 2, 3, 7, 50, 77, 100, 255, 256, 1000, 4000, 10000, 65535, 2**31 - 1, 2**32 - 1):
        cases += [(float(k), n) for k in range(0, min(n, 300) + 2)]            # non-smooth: integer counts
        cases += [(float(rng.uniform(-3, n + 3)), n) for _ in range(300)]      # smooth: count + 1 - nu
        cases += [(float(np.nextafter(float(k), np.inf)), n) for k in range(1, 40)]
    for _ in range(3000):
        n = int(rng.integers(1, 2**32))
        cases.append((float(rng.uniform(-10, n + 10)), n))
    for a, n in cases:
        assert fast_div(a, n) == rn(Fraction(a) / n), (a, n)
