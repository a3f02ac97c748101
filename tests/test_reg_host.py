"""CPU check of the register engine's per-pixel logic (siril_b200/csrc/reg.cuh, compiled for the
host by tests/helpers with one lane per pixel) against the oracle.  A stack is either decided
(FAST_DONE: rejection counts identical, mean within 1e-6 relative; median identical) or handed to
the exact engine (FAST_FALLBACK); on well-behaved deep stacks nearly everything must be decided."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle as O
from test_sorted_host import make_stack

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "helpers", "libreg_host.so")
SRC = os.path.join(HERE, "helpers", "reg_host.cu")
HDRS = [os.path.join(HERE, "..", "siril_b200", "csrc", h) for h in ("reg.cuh", "sorted.cuh", "exact.cuh")]

ALGOS = [O.NO_REJEC, O.PERCENTILE, O.SIGMA, O.WINSORIZED]
NAMES = {O.NO_REJEC: "none", O.PERCENTILE: "percentile", O.SIGMA: "sigma", O.WINSORIZED: "winsorized"}
SIGS = {O.NO_REJEC: (3, 3), O.PERCENTILE: (0.2, 0.3), O.SIGMA: (3, 3), O.WINSORIZED: (3, 3)}


@pytest.fixture(scope="module")
def rh():
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(p) for p in [SRC] + HDRS):
        subprocess.check_call(["nvcc", "-O2", "-std=c++17", "-fmad=false", "--extended-lambda", "--expt-relaxed-constexpr",
                               "-Wno-deprecated-gpu-targets", "-diag-suppress", "20013,20011,20014,20012",
                               "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-shared", "-o", SO, SRC],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    L = C.CDLL(SO)
    L.reg_host_stack.restype = C.c_int
    L.reg_host_stack.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.POINTER(C.c_int),
                                 C.POINTER(C.c_double)]
    return L


@pytest.fixture(scope="module")
def orc():
    return O.Oracle("port")


def run_reg(rh, st, is_mean, algo, sig):
    rej = (C.c_int * 2)()
    res = C.c_double()
    status = rh.reg_host_stack(st.ctypes.data_as(C.c_void_p), len(st), is_mean, algo, sig[0], sig[1], rej, C.byref(res))
    return status, res.value, rej[0], rej[1]


def check(rh, orc, st, algo, sig):
    status, m, lo, hi = run_reg(rh, st, 1, algo, sig)
    if status != 0:
        return False
    om, olo, ohi = orc.mean_and_reject(st, algo, sig)
    assert (lo, hi) == (olo, ohi), (NAMES[algo], len(st), (lo, hi), (olo, ohi), st.tolist())
    assert abs(m - om) <= 1e-6 * max(abs(om), 1e-30), (NAMES[algo], m, om, st.tolist())
    return True


@pytest.mark.parametrize("algo", ALGOS, ids=[NAMES[a] for a in ALGOS])
def test_reg_engine_matches_oracle(rh, orc, algo):
    rng = np.random.default_rng(4242 + algo)
    decided = {}
    for kind in ("clean", "hot", "outliers", "ties", "wild", "negative"):
        done = tot = 0
        for n in (32, 33, 47, 64, 100, 101, 150, 300, 301, 500):
            for _ in range(30):
                st = make_stack(rng, n, np.float32, kind)
                done += check(rh, orc, st, algo, SIGS[algo])
                tot += 1
        decided[kind] = done / tot
    assert decided["clean"] > 0.97 and decided["hot"] > 0.97, decided


@pytest.mark.parametrize("algo", [O.SIGMA, O.WINSORIZED], ids=["sigma", "winsorized"])
@pytest.mark.parametrize("n", [100, 300])
def test_reg_fallback_rate_realistic(rh, orc, algo, n):
    """stacks as in the benchmark configs: nearly every pixel is decided in registers"""
    rng = np.random.default_rng(50 + algo + n)
    done = 0
    T = 1200
    for _ in range(T):
        st = make_stack(rng, n, np.float32, "hot")
        done += check(rh, orc, st, algo, SIGS[algo])
    assert done / T > 0.995, done / T


def test_reg_median(rh, orc):
    rng = np.random.default_rng(31)
    done = tot = 0
    for n in (32, 33, 64, 100, 101, 200, 301, 500):
        for kind in ("clean", "hot", "zeros", "ties", "wild"):
            for _ in range(20):
                st = make_stack(rng, n, np.float32, kind)
                status, m, _, _ = run_reg(rh, st, 0, O.NO_REJEC, (3, 3))
                tot += 1
                if status == 0:
                    om, _ = orc.quickmedian(st)
                    assert m == om, (n, kind, m, om)
                    done += kind in ("clean", "hot")
    assert done >= 0.97 * (8 * 2 * 20), done
