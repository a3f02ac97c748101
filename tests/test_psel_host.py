"""CPU check of the selection engine's per-pixel logic (siril_b200/csrc/psel.cuh: the group-level
planning and phase B, compiled for the host by tests/helpers/psel_host.cu, which restates the
kernel's streaming passes over plain arrays) against the oracle.  A stack is either decided
(FAST_DONE: rejection counts identical, mean within 1e-6 relative — in practice 1 ulp — and the
median bit-equal) or handed to the exact engine (FAST_FALLBACK, covered by test_exact_host.py);
on well-behaved deep stacks the hand-over rate must be tiny, otherwise the engine is not doing the
work."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle as O
from test_sorted_host import make_stack

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "helpers", "libpsel_host.so")
SRC = os.path.join(HERE, "helpers", "psel_host.cu")
HDRS = [os.path.join(HERE, "..", "siril_b200", "csrc", h) for h in ("psel.cuh", "sorted.cuh", "exact.cuh")]

ALGOS = [O.NO_REJEC, O.PERCENTILE, O.SIGMA]
NAMES = {O.NO_REJEC: "none", O.PERCENTILE: "percentile", O.SIGMA: "sigma"}
SIGS = {O.NO_REJEC: (3, 3), O.PERCENTILE: (0.2, 0.3), O.SIGMA: (3, 3)}


@pytest.fixture(scope="module")
def ph():
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(p) for p in [SRC] + HDRS):
        subprocess.check_call(["nvcc", "-O2", "-std=c++17", "-fmad=false", "--extended-lambda", "--expt-relaxed-constexpr",
                               "-Wno-deprecated-gpu-targets", "-diag-suppress", "20013,20011,20014,20012",
                               "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-shared", "-o", SO, SRC],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    L = C.CDLL(SO)
    L.psel_host_pixel.restype = C.c_int
    L.psel_host_pixel.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p,
                                  C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_int]
    return L


def run(ph, st, T, is_mean, algo, sig, weights=None, lockstep=0):
    rej = (C.c_int * 2)()
    res = C.c_double()
    diag = (C.c_int * 6)()
    w = None if weights is None else np.ascontiguousarray(weights, dtype=np.float64)
    status = ph.psel_host_pixel(int(st.dtype == np.float32), st.ctypes.data_as(C.c_void_p), len(st), T, int(is_mean), algo,
                                sig[0], sig[1], None if w is None else w.ctypes.data_as(C.c_void_p), rej, C.byref(res), diag, lockstep)
    return status, res.value, rej[0], rej[1], list(diag)


def check_mean(ph, orc, st, T, algo, sig, weights=None):
    status, m, lo, hi, diag = run(ph, st, T, True, algo, sig, weights)
    if status != 0:
        return False, diag
    om, olo, ohi = orc.mean_and_reject(st, algo, sig, weights=weights)
    assert (lo, hi) == (olo, ohi), (NAMES[algo], st.dtype, len(st), T, (lo, hi), (olo, ohi), diag, st.tolist())
    assert abs(m - om) <= 1e-6 * max(abs(om), 1e-30), (NAMES[algo], m, om, diag, st.tolist())
    assert abs(np.float32(m) - np.float32(om)) <= np.spacing(np.float32(abs(om))), (NAMES[algo], m, om)
    return True, diag


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("algo", ALGOS, ids=[NAMES[a] for a in ALGOS])
def test_selection_engine_matches_oracle(ph, orc, algo, dtype):
    rng = np.random.default_rng(1977 + algo)
    decided = {}
    for kind in ("clean", "hot", "outliers", "zeros", "ties", "wild", "negative"):
        if kind == "negative" and dtype == np.uint16:
            continue
        done = tot = 0
        for n in (2, 3, 5, 8, 11, 12, 13, 17, 33, 64, 100, 150, 300):
            for _ in range(25):
                st = make_stack(rng, n, dtype, kind)
                T = int(rng.choice([1, 2, 4, 8]))
                w = rng.uniform(0.5, 1.5, n) if rng.random() < 0.2 else None
                ok, _ = check_mean(ph, orc, st, T, algo, SIGS[algo], w)
                if n >= 64:
                    done += ok; tot += 1
        decided[kind] = done / tot
    assert decided["clean"] > 0.97 and decided["hot"] > 0.97, decided


@pytest.mark.parametrize("n,T", [(100, 2), (200, 4), (300, 4), (500, 8)])
def test_handover_rate_realistic(ph, orc, n, T):
    """deep stacks as in the benchmark configs (sigma clip 3/3): the engine decides nearly all of
    them, with three streaming passes"""
    rng = np.random.default_rng(50 + n)
    done = 0
    K = 1500
    passes = []
    for _ in range(K):
        st = make_stack(rng, n, np.float32, "hot")
        ok, diag = check_mean(ph, orc, st, T, O.SIGMA, (3, 3))
        done += ok
        passes.append(diag[1])
    assert done / K > 0.995, done / K
    assert np.mean(passes) < 3.5, np.mean(passes)   # sums + count + extract, plus the occasional re-count or moved window


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
def test_median_mode(ph, orc, dtype):
    """median stacking: all N samples, zeros included; bit-equal to quickmedian when decided"""
    rng = np.random.default_rng(7)
    done = tot = 0
    for n in list(range(2, 20)) + [50, 64, 101, 200, 300]:
        for kind in ("clean", "hot", "zeros", "ties", "wild"):
            for _ in range(8):
                st = make_stack(rng, n, dtype, kind)
                T = int(rng.choice([1, 2, 4, 8]))
                status, m, _, _, diag = run(ph, st, T, False, O.NO_REJEC, (3, 3))
                if n >= 64 and kind in ("clean", "hot"):
                    done += status == 0; tot += 1
                if status == 0:
                    om, _ = orc.quickmedian(st)
                    assert m == om, (n, kind, T, m, om, diag)
    assert done / tot > 0.97, done / tot


def test_lockstep_with_other_groups(ph):
    """on the device the groups of a warp run every pass together: a group that is done must come
    through the passes other groups still need (re-counts, moved windows) unchanged"""
    rng = np.random.default_rng(12)
    for n in (64, 100, 300):
        for kind in ("clean", "hot", "zeros", "outliers"):
            for _ in range(40):
                st = make_stack(rng, n, np.float32, kind)
                for is_mean, algo in ((True, O.SIGMA), (True, O.PERCENTILE), (False, O.NO_REJEC)):
                    a = run(ph, st, 4, is_mean, algo, SIGS[algo])
                    b = run(ph, st, 4, is_mean, algo, SIGS[algo], lockstep=1)
                    assert a[:4] == b[:4], (n, kind, algo, a, b)


def test_weird_samples_are_handed_over(ph):
    st = (0.2 + 0.01 * np.random.default_rng(1).standard_normal(100)).astype(np.float32)
    for bad in (np.nan, np.inf, -np.inf):
        s2 = st.copy(); s2[17] = bad
        status, *_ = run(ph, s2, 4, True, O.SIGMA, (3, 3))
        assert status != 0


def test_ill_conditioned_sums_are_recentred(ph, orc):
    """a spread tiny beside the level: the plain sums are re-accumulated about the mean"""
    rng = np.random.default_rng(3)
    for _ in range(50):
        st = (0.9 + 2e-5 * rng.standard_normal(300)).astype(np.float32)
        ok, diag = check_mean(ph, orc, st, 4, O.SIGMA, (3, 3))
        assert diag[1] >= 4 or not ok
