"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/sgpu.h declares,
its host-side scalar helpers match the oracle, and compute fails loudly without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_exported():
    from siril_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "sgpu.h")).read()
    declared = set(re.findall(r"\b(sgpu_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"sgpu_last_error"} - set(_lib.SIGNATURES)
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    L = _lib.lib()
    for name in declared:
        assert hasattr(L, name), name
    assert L.sgpu_abi_version() == 5


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from siril_b200 import _lib
    with pytest.raises(_lib.SgpuError):
        _lib.Context(0)


def test_product_does_not_touch_oracle():
    """the product path must never import / link / call the oracle"""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "siril_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")) or fn == "Makefile":
                txt = open(os.path.join(dirpath, fn), errors="ignore").read()
                for needle in ("liborc", "libsiril_ref", "from oracle", "import oracle", "oracle/"):
                    assert needle not in txt, (fn, needle)


def test_planner_matches_oracle(orc):
    from siril_b200 import _lib
    L = _lib.lib()
    rng = np.random.default_rng(4)
    for _ in range(200):
        W, H = int(rng.integers(16, 9000)), int(rng.integers(16, 7000))
        Cn = int(rng.choice([1, 3]))
        thr = int(rng.integers(1, 33))
        mr = int(rng.integers(1, 3 * H * Cn))
        rv_o, blocks_o, largest_o = orc.planner(mr, W, H, Cn, thr)
        blocks = (_lib.ImageBlock * 4096)()
        naxes = (C.c_long * 3)(W, H, Cn)
        largest, nb = C.c_long(0), C.c_int(0)
        rv = L.sgpu_compute_parallel_blocks(blocks, 4096, mr, naxes, thr, C.byref(largest), C.byref(nb))
        assert (rv == 0) == (rv_o == 0)
        if rv == 0:
            got = np.array([[b.channel, b.start_row, b.end_row, b.height] for b in blocks[:nb.value]])
            assert np.array_equal(got, blocks_o) and largest.value == largest_o


def test_factors_and_weights_match_oracle(orc):
    from siril_b200 import _lib
    from oracle import oracle as O
    L = _lib.lib()
    rng = np.random.default_rng(6)
    for Cn in (1, 3):
        N = 11
        off, mul, sc = (rng.uniform(0.05, 0.2, (Cn, N)) for _ in range(3))
        sc[0, 3] = 0.0
        mul[Cn - 1, 5] = 0.0
        for nrm in (O.ADDITIVE, O.MULTIPLICATIVE, O.ADDITIVE_SCALING, O.MULTIPLICATIVE_SCALING):
            for eq in (False, True):
                exp = orc.compute_factors(nrm, off, mul, sc, 2, reglayer=-1 if Cn == 1 else 1, equalize_rgb=eq and Cn == 3)
                a, b, c = (np.array(x, copy=True) for x in (off, mul, sc))
                rv = L.sgpu_compute_factors(nrm, Cn, N, 2, -1 if Cn == 1 else 1, int(eq and Cn == 3),
                                            a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), c.ctypes.data_as(C.c_void_p))
                assert rv == 0
                for x, y in zip((a, b, c), exp):
                    assert np.array_equal(x, y)
        s = rng.uniform(0.8, 1.2, (Cn, N)); bg = rng.uniform(1e-3, 2e-3, (Cn, N))
        w = np.zeros((Cn, N))
        assert L.sgpu_noise_weights(Cn, N, s.ctypes.data_as(C.c_void_p), bg.ctypes.data_as(C.c_void_p), w.ctypes.data_as(C.c_void_p)) == 0
        assert np.array_equal(w, orc.noise_weights(s, bg))


def test_registration_weights_match_the_reference(ref):
    """WFWHM / NBSTARS / NBSTACK weights (median_and_mean.c:1137-1230, :150-159): host scalar code of the C ABI against the
    reference's own functions (oracle/_ref, ref_reg_weights), bit for bit"""
    from siril_b200 import _lib
    L = _lib.lib()
    R = ref.lib
    R.ref_reg_weights.restype = C.c_int
    R.ref_reg_weights.argtypes = [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    rng = np.random.default_rng(8)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    for N, Cn in [(2, 1), (7, 3), (40, 1), (300, 3)]:
        for case in range(4):
            fw = rng.uniform(1.5, 6.0, N).astype(np.float32)
            ns = rng.integers(5, 4000, N).astype(np.int32)
            if case == 1:
                fw[rng.integers(0, N)] = 0          # a frame without a measurement gets weight 0
                ns[:] = ns[0]                       # all alike: weights 1
            if case == 2:
                fw[:] = 0                           # nothing measured: the reference fails (norm == 0)
            if case == 3:
                fw[:] = fw[0]                       # all alike: 1 / 0 in invdenom, the reference's inf / NaN weights
            for nbstars in (0, 1):
                got, exp = np.full((Cn, N), -1.0), np.full((Cn, N), -1.0)
                rv = (L.sgpu_nbstars_weights(Cn, N, p(ns), p(got)) if nbstars else L.sgpu_wfwhm_weights(Cn, N, p(fw), p(got)))
                rr = R.ref_reg_weights(nbstars, Cn, N, p(fw), p(ns), p(exp))
                assert rv == rr, (N, Cn, case, nbstars, rv, rr)
                if rv == 0:
                    assert np.array_equal(got, exp, equal_nan=True), (N, Cn, case, nbstars)
            cnt = rng.integers(1, 50, N).astype(np.int32)
            got = np.zeros((Cn, N))
            assert L.sgpu_nbstack_weights(Cn, N, p(cnt), p(got)) == 0
            assert np.array_equal(got, np.broadcast_to(cnt.astype(np.float64), (Cn, N)))


def test_siril_glue_builds_against_the_reference_header_and_keeps_its_symbols():
    """integration/siril/src/gpu/gpu_stack.c compiles against the reference's stacking.h; the harness
    library carries the reference's own symbol names around it (no compute without a GPU)"""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    so = os.path.join(root, "integration", "_build", "libgpu_stack_harness.so")
    if os.path.isdir("/root/reference"):
        subprocess.check_call([sys.executable, os.path.join(root, "integration", "build_harness.py")], stdout=subprocess.DEVNULL)
    elif not os.path.exists(so):
        pytest.skip("neither the reference's header nor a prebuilt harness library here")
    H = C.CDLL(so)
    for name in ("stack_mean_with_rejection", "stack_median", "do_normalization", "gpu_stack_mean_or_median",
                 "gpu_do_normalization", "harness_stack"):
        assert hasattr(H, name), name


def test_seq_statistics_records_match_the_reference(ref):
    """the .seq file's `M` records (io/seqfile.c:838-853 written, :544-611 read): same text as the reference's own fprintf,
    same numbers back as its sscanf — %lg's six digits included"""
    from siril_b200 import _lib
    L, R = _lib.lib(), ref.lib
    R.ref_seq_stats_write.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_char_p, C.c_size_t]
    R.ref_seq_stats_read.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.c_void_p]
    rng = np.random.default_rng(12)
    for k in range(200):
        v = np.concatenate([[float(rng.integers(1, 1 << 33)), float(rng.integers(0, 1 << 33))],
                            rng.standard_normal(12) * 10.0 ** rng.integers(-8, 8, 12)])
        if k % 7 == 0:
            v[2:] = np.round(v[2:])
        if k % 11 == 0:
            v[5] = v[7] = 0.0
        st = _lib.ImStats(*v.tolist())
        layer, image = int(rng.integers(0, 3)), int(rng.integers(0, 100000))
        mine, theirs = C.create_string_buffer(1024), C.create_string_buffer(1024)
        assert L.sgpu_imstats_seq_record(layer, image, C.byref(st), mine, 1024) == 0
        assert R.ref_seq_stats_write(layer, image, v.ctypes.data, theirs, 1024) == 0
        assert mine.value == theirs.value and mine.value.startswith(b"M%d-%d " % (layer, image))
        back, gl, gi = _lib.ImStats(), C.c_int(), C.c_int()
        assert L.sgpu_imstats_from_seq_record(mine.value, C.byref(gl), C.byref(gi), C.byref(back)) == 0
        exp, ei = np.zeros(14), C.c_int()
        assert R.ref_seq_stats_read(theirs.value, C.byref(ei), exp.ctypes.data) == 15
        assert (gl.value, gi.value) == (layer, image) and ei.value == image
        assert [getattr(back, n) for n, _ in _lib.ImStats._fields_] == exp.tolist()
    last = mine.value
    assert L.sgpu_imstats_seq_record(0, 1, C.byref(st), mine, 10) != 0            # buffer too small
    assert L.sgpu_imstats_from_seq_record(b"R0 1 2 3", C.byref(gl), C.byref(gi), C.byref(back)) == -2
    assert L.sgpu_imstats_from_seq_record(b"M0-3 12 11 0.5", C.byref(gl), C.byref(gi), C.byref(back)) == -2
    assert L.sgpu_imstats_from_seq_record(b"M*-3 " + last.split(b" ", 1)[1], C.byref(gl), C.byref(gi), C.byref(back)) == 0 and gl.value == 0
