"""CPU check of the sorted engine's per-pixel logic (siril_b200/csrc/sorted.cuh, compiled for the
host by tests/helpers) against the oracle.  A stack either gets decided (FAST_DONE: rejection
counts must be identical and the mean within 1e-6 relative — in practice 1 ulp) or is handed to
the exact engine (FAST_FALLBACK), which test_exact_host.py covers; the fallback rate on
well-behaved stacks must stay small, otherwise the fast path is not doing the work."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "helpers", "libsorted_host.so")
SRC = os.path.join(HERE, "helpers", "sorted_host.cu")
HDRS = [os.path.join(HERE, "..", "siril_b200", "csrc", h) for h in ("sorted.cuh", "exact.cuh")]

FAST = [O.NO_REJEC, O.PERCENTILE, O.SIGMA, O.MAD, O.SIGMEDIAN, O.WINSORIZED, O.LINEARFIT, O.GESDT]
NAMES = {O.NO_REJEC: "none", O.PERCENTILE: "percentile", O.SIGMA: "sigma", O.WINSORIZED: "winsorized",
         O.LINEARFIT: "linearfit", O.GESDT: "gesdt", O.MAD: "mad", O.SIGMEDIAN: "sigmedian"}
SIGS = {O.NO_REJEC: (3, 3), O.PERCENTILE: (0.2, 0.3), O.SIGMA: (3, 3), O.WINSORIZED: (3, 3), O.LINEARFIT: (5, 5),
        O.GESDT: (0.3, 0.05), O.MAD: (3, 3), O.SIGMEDIAN: (3, 3)}


@pytest.fixture(scope="module")
def sh():
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(p) for p in [SRC] + HDRS):
        subprocess.check_call(["nvcc", "-O2", "-std=c++17", "-fmad=false", "--extended-lambda", "--expt-relaxed-constexpr",
                               "-Wno-deprecated-gpu-targets", "-diag-suppress", "20013,20011,20014,20012",
                               "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-shared", "-o", SO, SRC],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    L = C.CDLL(SO)
    L.sorted_host_mean_and_reject.restype = C.c_int
    L.sorted_host_mean_and_reject.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p,
                                              C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double), C.c_int]
    L.sorted_host_div_by_int.restype = C.c_double
    L.sorted_host_div_by_int.argtypes = [C.c_double, C.c_int]
    L.sorted_host_median.restype = C.c_double
    L.sorted_host_median.argtypes = [C.c_int, C.c_void_p, C.c_int]
    return L


def make_stack(rng, n, dtype, kind):
    base = 0.2 + 0.01 * rng.standard_normal(n)
    if kind == "clean":
        pass
    elif kind == "hot":
        k = min(int(rng.integers(0, 4)), n)
        if k:
            base[rng.choice(n, size=k, replace=False)] += rng.uniform(0.05, 0.6, size=k)
        if rng.random() < 0.2:
            base[rng.integers(n)] -= 0.08
    elif kind == "outliers":
        k = rng.integers(1, max(2, n // 4))
        idx = rng.choice(n, size=k, replace=False)
        base[idx] += rng.uniform(-0.15, 0.6, size=k)
    elif kind == "zeros":
        k = rng.integers(1, max(n, 2))
        base[rng.choice(n, size=min(k, n), replace=False)] = 0.0
    elif kind == "ties":
        base = np.round(base, 2)
    elif kind == "wild":
        base = rng.uniform(0, 1, n) ** 3
    elif kind == "negative":
        base = base - 0.2
    base = np.clip(base, -1, 1) if kind == "negative" else np.clip(base, 0, 1)
    if dtype == np.uint16:
        return np.round(np.clip(base, 0, 1) * 65535).astype(np.uint16)
    return base.astype(np.float32)


def run_fast(sh, st, algo, sig, crit=None, weights=None, use_s0=1):
    rej = (C.c_int * 2)()
    mean = C.c_double()
    cr = None if crit is None else np.ascontiguousarray(crit, dtype=np.float32)
    w = None if weights is None else np.ascontiguousarray(weights, dtype=np.float64)
    status = sh.sorted_host_mean_and_reject(int(st.dtype == np.float32), st.ctypes.data_as(C.c_void_p), len(st), algo,
                                            sig[0], sig[1], None if cr is None else cr.ctypes.data_as(C.c_void_p),
                                            None if w is None else w.ctypes.data_as(C.c_void_p), 0, rej, C.byref(mean), use_s0)
    return status, mean.value, rej[0], rej[1]


def check(sh, orc, st, algo, sig, crit=None, weights=None):
    s_a, m_a, lo_a, hi_a = run_fast(sh, st, algo, sig, crit, weights, use_s0=0)
    status, m, lo, hi = run_fast(sh, st, algo, sig, crit, weights, use_s0=1)
    if s_a == 0 and status == 0:   # own sums or phase A's pre-accumulated ones: same decisions
        assert (lo, hi) == (lo_a, hi_a) and abs(m - m_a) <= 1e-12 * max(abs(m), 1e-30)
    if status != 0:
        return False
    om, olo, ohi = orc.mean_and_reject(st, algo, sig, crit=crit, weights=weights)
    assert (lo, hi) == (olo, ohi), (NAMES[algo], st.dtype, len(st), (lo, hi), (olo, ohi), st.tolist())
    assert abs(m - om) <= 1e-6 * max(abs(om), 1e-30), (NAMES[algo], m, om, st.tolist())
    assert abs(np.float32(m) - np.float32(om)) <= np.spacing(np.float32(abs(om))), (NAMES[algo], m, om)
    return True


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("algo", FAST, ids=[NAMES[a] for a in FAST])
def test_sorted_engine_matches_oracle(sh, orc, algo, dtype):
    rng = np.random.default_rng(977 + algo)
    decided = {}
    for kind in ("clean", "hot", "outliers", "zeros", "ties", "wild", "negative"):
        if kind == "negative" and dtype == np.uint16:
            continue
        done = tot = 0
        for n in (2, 3, 4, 5, 7, 8, 9, 10, 16, 17, 33, 64, 100, 150):
            if algo == O.GESDT and n < 3:
                continue
            for _ in range(40):
                st = make_stack(rng, n, dtype, kind)
                crit = O.gesdt_critical_values(n, *SIGS[algo]) if algo == O.GESDT else None
                w = rng.uniform(0.5, 1.5, n) if rng.random() < 0.3 else None
                done += check(sh, orc, st, algo, SIGS[algo], crit, w)
                tot += 1
        decided[kind] = done / tot
    # well-behaved stacks are decided by the fast path (u16 winsorized is the exact engine's)
    if not (algo == O.WINSORIZED and dtype == np.uint16):
        assert decided["clean"] > 0.9 and decided["hot"] > 0.9, decided


@pytest.mark.parametrize("algo", [O.SIGMA, O.WINSORIZED, O.GESDT, O.LINEARFIT], ids=["sigma", "winsorized", "gesdt", "linearfit"])
def test_fallback_rate_realistic(sh, orc, algo):
    """deep stacks as in the benchmark configs: nearly every pixel is decided by the fast path"""
    rng = np.random.default_rng(5 + algo)
    n = 100
    crit = O.gesdt_critical_values(n, *SIGS[algo]) if algo == O.GESDT else None
    done = 0
    T = 1500
    for _ in range(T):
        st = make_stack(rng, n, np.float32, "hot")
        done += check(sh, orc, st, algo, SIGS[algo], crit)
    assert done / T > 0.97, done / T


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
def test_median_sorted(sh, orc, dtype):
    rng = np.random.default_rng(3)
    for n in list(range(1, 20)) + [50, 101, 200]:
        for kind in ("clean", "zeros", "ties"):
            st = make_stack(rng, n, dtype, kind)
            m = sh.sorted_host_median(int(dtype == np.float32), st.ctypes.data_as(C.c_void_p), n)
            om, _ = orc.quickmedian(st)
            assert m == om, (n, kind, m, om)


def test_div_by_int(sh):
    """the Markstein division used for GESDT's window mean is the correctly rounded quotient"""
    rng = np.random.default_rng(11)
    for _ in range(20000):
        n = int(rng.integers(1, 2049))
        a = float(np.float32(rng.uniform(0, 1))) * int(rng.integers(1, 4 * n)) if rng.random() < 0.5 else float(rng.uniform(-1e3, 1e3))
        assert sh.sorted_host_div_by_int(a, n) == a / n, (a, n)
