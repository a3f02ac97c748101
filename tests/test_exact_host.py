"""CPU check of the product's thread-per-pixel device functions (siril_b200/csrc/exact.cuh
compiled for the host by tests/helpers) against the oracle: every rejection algorithm, both
sample types, weights.  Bit-exact (same IEEE operations in the same order)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "helpers", "libexact_host.so")
SRC = os.path.join(HERE, "helpers", "exact_host.cu")
HDR = os.path.join(HERE, "..", "siril_b200", "csrc", "exact.cuh")


@pytest.fixture(scope="module")
def xh():
    if not os.path.exists(SO) or os.path.getmtime(SO) < max(os.path.getmtime(SRC), os.path.getmtime(HDR)):
        subprocess.check_call(["nvcc", "-O2", "-std=c++17", "-fmad=false", "--extended-lambda", "--expt-relaxed-constexpr",
                               "-Wno-deprecated-gpu-targets", "-diag-suppress", "20013,20011,20014,20012",
                               "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-shared", "-o", SO, SRC],
                              stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    L = C.CDLL(SO)
    L.exact_host_mean_and_reject.restype = C.c_double
    L.exact_host_mean_and_reject.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p,
                                             C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    return L


ALGOS = list(range(8))
NAMES = ["none", "percentile", "sigma", "mad", "sigmedian", "winsorized", "linearfit", "gesdt"]
SIGS = {0: (3, 3), 1: (0.2, 0.3), 2: (3, 3), 3: (3, 3), 4: (3, 3), 5: (3, 3), 6: (5, 5), 7: (0.3, 0.05)}


def make_stack(rng, n, dtype, kind):
    base = 0.2 + 0.01 * rng.standard_normal(n)
    if kind == "outliers":
        k = rng.integers(1, max(2, n // 4))
        idx = rng.choice(n, size=k, replace=False)
        base[idx] += rng.uniform(-0.15, 0.6, size=k)
    elif kind == "zeros":
        k = rng.integers(1, n)
        base[rng.choice(n, size=k, replace=False)] = 0.0
    elif kind == "ties":
        base = np.round(base, 2)
    elif kind == "wild":
        base = rng.uniform(0, 1, n) ** 3
    base = np.clip(base, 0, 1)
    if dtype == np.uint16:
        return np.round(base * 65535).astype(np.uint16)
    return base.astype(np.float32)


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("algo", ALGOS, ids=NAMES)
def test_exact_engine_matches_oracle(xh, orc, algo, dtype):
    rng = np.random.default_rng(4321 + algo)
    for n in [2, 3, 4, 5, 6, 8, 9, 10, 12, 15, 16, 17, 20, 24, 31, 33, 50, 64, 100, 129, 300]:
        if algo == O.GESDT and n < 3:
            continue
        crit = O.gesdt_critical_values(n, *SIGS[algo]) if algo == O.GESDT else None
        for kind in ["clean", "outliers", "zeros", "ties", "wild"]:
            for rep in range(4):
                st = make_stack(rng, n, dtype, kind)
                for w in (None, rng.uniform(0.5, 1.5, n)):
                    exp = orc.mean_and_reject(st, algo, SIGS[algo], crit=crit, weights=w)
                    rej = (C.c_int * 2)()
                    got = xh.exact_host_mean_and_reject(int(dtype == np.float32), st.ctypes.data_as(C.c_void_p), n, algo,
                                                        SIGS[algo][0], SIGS[algo][1],
                                                        crit.ctypes.data_as(C.c_void_p) if crit is not None else None,
                                                        w.ctypes.data_as(C.c_void_p) if w is not None else None, 0, rej)
                    got = (got, rej[0], rej[1])
                    if np.isnan(got[0]) and np.isnan(exp[0]):
                        assert got[1:] == exp[1:]
                    else:
                        assert got == exp, (n, kind, rep, w is not None, got, exp, st.tolist())
