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
    L.exact_host_mean_and_reject_planes.restype = C.c_double
    L.exact_host_mean_and_reject_planes.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p,
                                                    C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int)]
    return L


def mask_ramp(v, feather_dist):
    """the mask normalization of median_and_mean.c:513-521 + blending.c:34-50 in float32 numpy (test-side twin)"""
    f32 = np.float32
    v = np.asarray(v, dtype=f32)
    t = v * (f32(1) / f32(feather_dist))
    r = (t * f32(1000)).astype(np.int32).astype(f32) * (f32(1) / f32(1000))
    ramp = r * r * r * (f32(6) * r * r - f32(15) * r + f32(10))
    return np.where(v == 0, f32(0), np.where(v > f32(feather_dist), f32(1), ramp)).astype(f32)


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


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("algo", ALGOS, ids=NAMES)
def test_exact_engine_masks_and_drizzle(xh, algo, dtype):
    """row n3: feather masks and drizzle weights as per-sample weights (median_and_mean.c:1043-1070) and the
    drizzle-null removal (rejection_float.c:116-126), against the reference build's own mean_and_reject"""
    if not O.have_reference():
        pytest.skip("oracle/_ref not built")
    ref = O.Oracle("reference")
    rng = np.random.default_rng(777 + algo)
    fd = 25
    for n in [2, 3, 5, 8, 13, 20, 33, 64]:
        if algo == O.GESDT and n < 3:
            continue
        crit = O.gesdt_critical_values(n, *SIGS[algo]) if algo == O.GESDT else None
        W = 24
        stacks = np.stack([make_stack(rng, n, dtype, kind) for kind in ["clean", "outliers", "zeros", "ties", "wild", "outliers"] * 4], axis=1)
        frames = stacks.reshape(n, 1, 1, W)
        masks = (rng.random((n, 1, W)) * 60).astype(np.float32)
        masks[rng.random(masks.shape) < 0.25] = 0
        dz = rng.random((n, 1, 1, W)).astype(np.float32)
        dz[rng.random(dz.shape) < 0.2] = 0
        wts = rng.uniform(0.5, 1.5, (1, n))
        for use_m, use_d, use_w in [(1, 0, 0), (0, 1, 0), (1, 1, 0), (1, 1, 1), (0, 1, 1)]:
            exp = ref.stack(frames, rejection=algo, sig=SIGS[algo], crit=crit, create_rejmaps=True,
                            masks=masks if use_m else None, feather_dist=fd, drizzle_weights=dz if use_d else None,
                            weights=wts if use_w else None, use_32bit_output=True, nthreads=1)
            rm = mask_ramp(masks, fd)
            for x in range(W):
                st = np.ascontiguousarray(frames[:, 0, 0, x])
                ms = np.ascontiguousarray(rm[:, 0, x]) if use_m else None
                ds = np.ascontiguousarray(dz[:, 0, 0, x]) if use_d else None
                rej = (C.c_int * 2)()
                got = xh.exact_host_mean_and_reject_planes(
                    int(dtype == np.float32), st.ctypes.data_as(C.c_void_p), n, algo, SIGS[algo][0], SIGS[algo][1],
                    crit.ctypes.data_as(C.c_void_p) if crit is not None else None,
                    wts.ctypes.data_as(C.c_void_p) if use_w else None,
                    ms.ctypes.data_as(C.c_void_p) if ms is not None else None,
                    ds.ctypes.data_as(C.c_void_p) if ds is not None else None, rej)
                gotf = np.float32(got) if dtype == np.float32 else np.float32(got) * np.float32(.000015259022)
                e = exp["out"][0, 0, x]
                assert (np.isnan(gotf) and np.isnan(e)) or gotf == e, (n, x, use_m, use_d, use_w, got, e)
                assert (rej[0], rej[1]) == (exp["rej_low"][0, 0, x], exp["rej_high"][0, 0, x])
