"""The Siril-side glue (integration/siril/src/gpu/gpu_stack.c), compiled against the reference's own
stacking.h and a miniature sequence store (integration/harness), driven the way main_stack drives the
CPU path: a `struct stacking_args` is filled, `do_normalization` is called, then the method through its
`stack_method` pointer (stacking.c:98,104).  Results are compared with the oracle running the same
pipeline on the CPU.  The harness library is built here by integration/build_harness.py (it needs the
reference's header); on the GPU box the prebuilt library is used."""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "integration", "_build", "libgpu_stack_harness.so")
ST_OK, ST_GENERIC_ERROR, ST_CANCEL = 0, -1, -9
NO_WEIGHT, NBSTARS_WEIGHT, WFWHM_WEIGHT, NOISE_WEIGHT = 0, 1, 2, 3


class Params(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("N", "C", "H", "W", "is_float", "is_mean", "rejection")] + [("sig", C.c_float * 2)] + \
               [(n, C.c_int) for n in ("normalize", "lite_norm", "weighting", "ref_image", "reglayer", "equalizeRGB")] + \
               [("dx", C.c_void_p), ("dy", C.c_void_p)] + \
               [(n, C.c_int) for n in ("output_norm", "use_32bit_output", "create_rejmaps", "merge_rejmaps", "cancel_after", "runs")] + \
               [("wfwhm", C.c_void_p), ("nbstars", C.c_void_p)]


@pytest.fixture(scope="module")
def harness():
    if not os.path.exists(SO):
        if not os.path.isdir("/root/reference"):
            pytest.skip("harness library not prebuilt and the reference's header is not on this machine")
        import subprocess, sys
        subprocess.check_call([sys.executable, os.path.join(ROOT, "integration", "build_harness.py")], stdout=subprocess.DEVNULL)
    L = C.CDLL(SO)
    L.harness_stack.restype = C.c_int
    L.harness_stack.argtypes = [C.POINTER(Params), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_char_p, C.c_int, C.POINTER(C.c_long)]
    return L


def synth(rng, N, Cn, H, W, dtype):
    sky = 0.1 * (1 + 0.1 * rng.standard_normal((N, 1, 1, 1)))
    gain = 1 + 0.1 * rng.standard_normal((N, 1, 1, 1))
    yy, xx = np.mgrid[0:H, 0:W]
    sig = 0.2 * np.exp(-((yy - H / 2) ** 2 + (xx - W / 2) ** 2) / (0.1 * H * W))
    fr = gain * (sky + sig[None, None]) + 0.01 * rng.standard_normal((N, Cn, H, W))
    m = rng.random(fr.shape) < 0.01
    fr[m] += rng.uniform(0.2, 0.6, m.sum())
    fr = np.clip(fr, 1e-4, 1.0)
    fr[rng.random(fr.shape) < 0.002] = 0.0
    return np.round(fr * 65535).astype(np.uint16) if dtype == np.uint16 else fr.astype(np.float32)


def run(L, fr, *, is_mean=True, rejection=O.NO_REJEC, sig=(3, 3), normalize=O.NO_NORM, lite_norm=False, weighting=NO_WEIGHT,
        ref_image=0, reglayer=-1, equalizeRGB=False, dx=None, dy=None, output_norm=False, use_32bit_output=True,
        create_rejmaps=False, merge_rejmaps=False, cancel_after=-1, runs=1, wfwhm=None, nbstars=None):
    N, Cn, H, W = fr.shape
    p = Params(N=N, C=Cn, H=H, W=W, is_float=int(fr.dtype == np.float32), is_mean=int(is_mean), rejection=rejection,
               normalize=normalize, lite_norm=int(lite_norm), weighting=weighting, ref_image=ref_image, reglayer=reglayer,
               equalizeRGB=int(equalizeRGB), output_norm=int(output_norm), use_32bit_output=int(use_32bit_output),
               create_rejmaps=int(create_rejmaps), merge_rejmaps=int(merge_rejmaps), cancel_after=cancel_after, runs=runs)
    p.sig[0], p.sig[1] = sig
    keep = []
    if dx is not None:
        dxa, dya = np.ascontiguousarray(dx, dtype=np.float64), np.ascontiguousarray(dy, dtype=np.float64)
        keep += [dxa, dya]
        p.dx, p.dy = dxa.ctypes.data, dya.ctypes.data
    if wfwhm is not None:
        keep.append(np.ascontiguousarray(wfwhm, dtype=np.float32))
        p.wfwhm = keep[-1].ctypes.data
    if nbstars is not None:
        keep.append(np.ascontiguousarray(nbstars, dtype=np.int32))
        p.nbstars = keep[-1].ctypes.data
    out = np.zeros((Cn, H, W), dtype=np.float32 if use_32bit_output else np.uint16)
    lo = np.zeros((Cn, H, W), dtype=np.uint16)
    hi = np.zeros((Cn, H, W), dtype=np.uint16)
    coeff = np.zeros((3, Cn, N))
    log = C.create_string_buffer(8192)
    info = (C.c_long * 4)()
    fr = np.ascontiguousarray(fr)
    rv = L.harness_stack(C.byref(p), fr.ctypes.data, out.ctypes.data, lo.ctypes.data, hi.ctypes.data, coeff.ctypes.data, log, 8192, info)
    return dict(rv=rv, out=out, rej_low=lo, rej_high=hi, offset=coeff[0], mul=coeff[1], scale=coeff[2],
                log=log.value.decode(errors="replace"), open=info[0], reads=info[1], cached=info[2], polls=info[3])


def oracle_pipeline(orc, fr, normalize, lite, ref_image, reglayer, equalize):
    """do_normalization on the CPU: statistics -> estimators -> factors; also the noise estimates"""
    N, Cn = fr.shape[:2]
    off = np.zeros((Cn, N)); mul = np.ones((Cn, N)); sc = np.ones((Cn, N)); bg = np.zeros((Cn, N))
    for f in range(N):
        for l in range(Cn):
            st = orc.image_stats(fr[f, l], lite)
            bg[l, f] = st["bgnoise"]
            loc = st["median"] if lite else st["location"]
            if normalize in (O.ADDITIVE_SCALING, O.MULTIPLICATIVE_SCALING):
                sc[l, f] = 1.5 * st["mad"] if lite else st["scale"]
            if normalize in (O.ADDITIVE, O.ADDITIVE_SCALING):
                off[l, f] = loc
            else:
                mul[l, f] = loc
    return orc.compute_factors(normalize, off, mul, sc, ref_image, reglayer=reglayer, equalize_rgb=equalize) + (bg,)


def compare(g, o, what):
    for k in ("rej_low", "rej_high"):
        if o[k] is not None:
            assert np.array_equal(g[k], o[k]), (what, k)
    a, b = g["out"], o["out"]
    if a.dtype == np.uint16:
        assert np.array_equal(a, b), what
    else:
        rel = np.abs(a - b) / np.maximum(np.abs(b), 1e-30)
        rel[a == b] = 0
        assert rel.max() <= 1e-6, (what, float(rel.max()))


def log_percentages(irej, N, H, W):
    tot = float(W) * float(H) * float(N)
    return ["Pixel rejection in channel #%d: %.3f%% - %.3f%%" % (c, irej[c][0] / tot * 100.0, irej[c][1] / tot * 100.0) for c in range(len(irej))]


CASES = [
    dict(rejection=O.SIGMA, sig=(3, 3), dtype=np.float32, Cn=1),
    dict(rejection=O.WINSORIZED, sig=(3, 3), dtype=np.float32, Cn=3, normalize=O.ADDITIVE_SCALING, weighting=NOISE_WEIGHT,
         ref_image=2, equalizeRGB=True, reglayer=1, create_rejmaps=True),
    dict(rejection=O.GESDT, sig=(0.3, 0.05), dtype=np.float32, Cn=1, create_rejmaps=True, merge_rejmaps=True),
    dict(rejection=O.LINEARFIT, sig=(5, 5), dtype=np.uint16, Cn=1, normalize=O.MULTIPLICATIVE, use_32bit_output=False, create_rejmaps=True),
    dict(rejection=O.SIGMA, sig=(3, 3), dtype=np.uint16, Cn=3, normalize=O.MULTIPLICATIVE_SCALING, lite_norm=True, weighting=NOISE_WEIGHT,
         shifts=True, create_rejmaps=True),
    dict(rejection=O.PERCENTILE, sig=(0.2, 0.3), dtype=np.float32, Cn=1, normalize=O.ADDITIVE, output_norm=True, shifts=True),
    dict(is_mean=False, dtype=np.float32, Cn=3, normalize=O.ADDITIVE_SCALING, ref_image=1),
    dict(rejection=O.SIGMA, sig=(3, 3), dtype=np.float32, Cn=1, weighting=WFWHM_WEIGHT, create_rejmaps=True),
    dict(rejection=O.WINSORIZED, sig=(3, 3), dtype=np.uint16, Cn=3, weighting=NBSTARS_WEIGHT, shifts=True),
    dict(is_mean=False, dtype=np.uint16, Cn=1, use_32bit_output=False, shifts=True),
]


@pytest.mark.parametrize("max_rows", [0, 12], ids=["one_block", "row_blocks"])
@pytest.mark.parametrize("case", range(len(CASES)))
def test_method_pointer_matches_oracle(harness, orc, case, max_rows, monkeypatch):
    cs = dict(CASES[case])
    if max_rows:
        monkeypatch.setenv("SIRIL_GPU_MAX_ROWS", str(max_rows))
    rng = np.random.default_rng(300 + case)
    N, H, W = 14, 40, 72
    dtype, Cn, shifts = cs.pop("dtype"), cs.pop("Cn"), cs.pop("shifts", False)
    fr = synth(rng, N, Cn, H, W, dtype)
    fr *= (1 + 0.2 * rng.random((N, 1, 1, 1))).astype(fr.dtype) if dtype == np.float32 else 1
    kw = dict(cs)
    reg_w = None
    if cs.get("weighting") in (WFWHM_WEIGHT, NBSTARS_WEIGHT):   # registration data of the frames (median_and_mean.c:1137-1230)
        from oracle.oracle import Oracle
        fw, ns = rng.uniform(1.5, 6.0, N).astype(np.float32), rng.integers(20, 3000, N).astype(np.int32)
        kw.update(wfwhm=fw, nbstars=ns)
        R = Oracle("reference").lib
        R.ref_reg_weights.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        reg_w = np.zeros((Cn, N))
        assert R.ref_reg_weights(int(cs["weighting"] == NBSTARS_WEIGHT), Cn, N, fw.ctypes.data, ns.ctypes.data, reg_w.ctypes.data) == 0
    if shifts:
        kw["reglayer"] = kw.get("reglayer", 0) if kw.get("reglayer", -1) >= 0 else 0
        kw.update(dx=rng.integers(-6, 7, N) + rng.uniform(-0.4, 0.4, N), dy=rng.integers(-9, 10, N) + rng.uniform(-0.4, 0.4, N))
    g = run(harness, fr, **kw)
    assert g["rv"] == ST_OK, g["log"]
    assert g["open"] == 0                      # every frame opened was closed (median_and_mean.c:1784-1786)
    is_mean = cs.get("is_mean", True)
    nrm = cs.get("normalize", O.NO_NORM)
    okw = dict(is_mean=is_mean, rejection=cs.get("rejection", O.NO_REJEC), sig=cs.get("sig", (3, 3)),
               output_norm=cs.get("output_norm", False), use_32bit_output=cs.get("use_32bit_output", True),
               create_rejmaps=bool(is_mean and cs.get("create_rejmaps")), merge_rejmaps=cs.get("merge_rejmaps", False))
    if nrm != O.NO_NORM:
        reglayer = kw.get("reglayer", -1)
        eo, em, es, bg = oracle_pipeline(orc, fr, nrm, cs.get("lite_norm", False), cs.get("ref_image", 0), reglayer, cs.get("equalizeRGB", False))
        for got, exp in ((g["offset"], eo), (g["mul"], em), (g["scale"], es)):
            np.testing.assert_allclose(got, exp, rtol=1e-12, atol=0)
        # the stack itself is compared with the glue's own coefficients on both sides
        okw.update(normalize=nrm, offset=g["offset"], mul=g["mul"], scale=g["scale"])
        assert g["cached"] == N * Cn           # add_stats_to_seq: one record per frame and layer
        if cs.get("weighting") == NOISE_WEIGHT and is_mean:
            okw["weights"] = orc.noise_weights(g["scale"], bg)
    if reg_w is not None:
        okw["weights"] = reg_w
    if shifts:
        okw.update(shiftx=np.rint(kw["dx"]).astype(np.int32), shifty=np.rint(kw["dy"]).astype(np.int32))
    o = orc.stack(fr, **okw)
    compare(g, o, str(CASES[case]))
    if is_mean:
        for line in log_percentages(o["irej"][:Cn], N, H, W):
            assert line in g["log"], (line, g["log"])


def test_second_run_is_served_by_the_statistics_cache(harness):
    rng = np.random.default_rng(5)
    fr = synth(rng, 8, 1, 32, 48, np.float32)
    one = run(harness, fr, rejection=O.SIGMA, normalize=O.ADDITIVE_SCALING)
    two = run(harness, fr, rejection=O.SIGMA, normalize=O.ADDITIVE_SCALING, runs=2)
    assert one["rv"] == two["rv"] == ST_OK
    # reads of the LAST run: the first run loads every frame for its statistics and again for the stack,
    # the second only for the stack
    assert two["reads"] == one["reads"] - 8
    assert np.array_equal(one["out"], two["out"])
    assert np.array_equal(one["scale"], two["scale"])


def test_cancel_and_errors(harness):
    rng = np.random.default_rng(6)
    fr = synth(rng, 6, 1, 32, 48, np.float32)
    g = run(harness, fr, rejection=O.SIGMA, cancel_after=0)
    assert g["rv"] == ST_CANCEL and g["open"] == 0 and "cancel" in g["log"].lower()
    g = run(harness, fr, rejection=O.SIGMA, normalize=O.ADDITIVE, cancel_after=3)   # during the statistics
    assert g["rv"] == ST_CANCEL and g["open"] == 0
    g = run(harness, fr[:1], rejection=O.SIGMA)                                      # median_and_mean.c:1279-1285
    assert g["rv"] == ST_GENERIC_ERROR and "at least two" in g["log"]
    g = run(harness, fr[:2], rejection=O.GESDT, sig=(0.3, 0.05))
    assert g["rv"] == ST_GENERIC_ERROR and "three frames" in g["log"]
    g = run(harness, fr, rejection=O.SIGMA, weighting=NOISE_WEIGHT)                  # command.c:11700-11703
    assert g["rv"] == ST_GENERIC_ERROR and "requires normalization" in g["log"]
    g = run(harness, fr, rejection=O.SIGMA, use_32bit_output=False)                  # float input, 16-bit output
    assert g["rv"] == ST_GENERIC_ERROR
