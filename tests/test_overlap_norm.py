"""Overlap normalization (normalization.c:296-355, 420-456, 549-593, 875-906): the C ABI against
the reference's own lines compiled into oracle/_ref.  The host-side pieces (rectangles, linear systems) run here
without a GPU; the pair statistics are the GPU test at the end."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle as O


def _ref_lib(ref):
    R = ref.lib
    R.ref_compute_overlap.restype = C.c_size_t
    R.ref_compute_overlap.argtypes = [C.c_int] * 4 + [C.c_double] * 4 + [C.c_void_p, C.c_void_p]
    R.ref_overlap_pair.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
    R.ref_overlap_solve.argtypes = [C.c_int] * 3 + [C.c_void_p] * 6
    return R


def test_compute_overlap_matches_the_reference(ref):
    from siril_b200 import _lib
    L, R = _lib.lib(), _ref_lib(ref)
    rng = np.random.default_rng(1)
    hits = 0
    for _ in range(3000):
        rxi, ryi, rxj, ryj = (int(v) for v in rng.integers(8, 200, 4))
        dxi, dyi, dxj, dyj = (float(v) for v in rng.uniform(-150, 150, 4))
        if rng.random() < 0.2:
            dxj, dyj = dxi + float(rng.integers(-3, 4)) + 0.5, dyi + float(rng.integers(-3, 4)) - 0.5   # rounding ties
        ai, aj = _lib.Rect(), _lib.Rect()
        ei, ej = np.zeros(4, dtype=np.int32), np.zeros(4, dtype=np.int32)
        n = L.sgpu_compute_overlap(rxi, ryi, rxj, ryj, dxi, dyi, dxj, dyj, C.byref(ai), C.byref(aj))
        m = R.ref_compute_overlap(rxi, ryi, rxj, ryj, dxi, dyi, dxj, dyj, ei.ctypes.data, ej.ctypes.data)
        assert n == m
        if n:
            hits += 1
            assert [ai.x, ai.y, ai.w, ai.h] == ei.tolist() and [aj.x, aj.y, aj.w, aj.h] == ej.tolist()
    assert hits > 500


def _system(rng, nf, ref_index, sparse, unit_scale=False):
    """a consistent mosaic: frame i sees  true * s_i + o_i ; pair statistics over random overlaps"""
    Nij = rng.integers(50, 5000, (nf, nf))
    Nij = np.triu(Nij, 1)
    if sparse:      # a chain plus a few extra overlaps: most pairs do not overlap at all
        keep = np.zeros((nf, nf), dtype=bool)
        for i in range(nf - 1):
            keep[i, i + 1] = True
        keep |= np.triu(rng.random((nf, nf)) < 0.15, 1)
        Nij = Nij * keep
    Nij = (Nij + Nij.T).astype(np.uint64)
    s = np.ones(nf) if unit_scale else rng.uniform(0.8, 1.25, nf)
    o = rng.uniform(-0.02, 0.02, nf)
    level = rng.uniform(0.05, 0.2, (nf, nf)); level = np.triu(level, 1); level = level + level.T      # sky of the common area
    width = rng.uniform(0.005, 0.02, (nf, nf)); width = np.triu(width, 1); width = width + width.T
    Mij = (level * s[:, None] + o[:, None]) * (Nij > 0)
    Sij = (width * s[:, None]) * (Nij > 0)
    return Nij, Mij, Sij, s, o


@pytest.mark.parametrize("normalize", [O.ADDITIVE, O.MULTIPLICATIVE, O.ADDITIVE_SCALING, O.MULTIPLICATIVE_SCALING],
                         ids=["add", "mul", "addscale", "mulscale"])
def test_overlap_solve_matches_the_reference(ref, normalize):
    from siril_b200 import _lib
    L, R = _lib.lib(), _ref_lib(ref)
    rng = np.random.default_rng(40 + normalize)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    for nf, sparse in [(2, False), (3, False), (6, False), (12, True), (40, True)]:
        ref_index = int(rng.integers(0, nf))
        Nij, Mij, Sij, s, o = _system(rng, nf, ref_index, sparse, unit_scale=normalize == O.ADDITIVE)
        got = [np.zeros(nf), np.ones(nf), np.ones(nf)]
        exp = [np.zeros(nf), np.ones(nf), np.ones(nf)]
        m1, m2 = Mij.copy(), Mij.copy()
        assert L.sgpu_overlap_solve(normalize, nf, ref_index, p(Nij), p(m1), p(Sij), *[p(a) for a in got]) == 0
        assert R.ref_overlap_solve(normalize, nf, ref_index, p(Nij), p(m2), p(Sij), *[p(a) for a in exp]) == 0
        for g, e in zip(got, exp):
            np.testing.assert_allclose(g, e, rtol=1e-12, atol=1e-15)
        off, mul, sca = got
        assert off[ref_index] == 0 and mul[ref_index] == 1 and sca[ref_index] == 1
        # the system is consistent by construction, so the solution is the exact levelling of the mosaic: an
        # independent check of the LU solve (the reference's is GSL's, absent here)
        if normalize in (O.ADDITIVE_SCALING, O.MULTIPLICATIVE_SCALING):
            np.testing.assert_allclose(sca, s[ref_index] / s, rtol=1e-9)
        if normalize == O.ADDITIVE:
            np.testing.assert_allclose(off, o - o[ref_index], rtol=0, atol=1e-9)     # stack value = px * scale - offset
        if normalize == O.ADDITIVE_SCALING:
            np.testing.assert_allclose(off, o * sca - o[ref_index], rtol=0, atol=1e-9)


@pytest.mark.gpu
@pytest.mark.parametrize("on_device", [1, 0], ids=["device", "host"])
@pytest.mark.parametrize("lite", [0, 1], ids=["ikss", "lite"])
@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
def test_overlap_pair_statistics_match_the_reference(ref, dtype, lite, on_device):
    torch = pytest.importorskip("torch")
    from siril_b200 import _lib
    L, R = _lib.lib(), _ref_lib(ref)
    ctx = _lib.Context(0)
    rng = np.random.default_rng(17)
    for rxi, ryi, rxj, ryj, dxi, dyi, dxj, dyj in [(300, 200, 300, 200, 0, 0, 41.3, -77.6), (257, 190, 310, 222, 12.2, 5.5, -100.7, 60.1),
                                                      (128, 96, 128, 96, 0, 0, 100.0, 0.0), (400, 300, 400, 300, 3, 4, 3.4, 3.6)]:
        def frame(rx, ry, lvl):
            yy, xx = np.mgrid[0:ry, 0:rx]
            f = lvl + 0.05 * np.sin(xx / 37.0) * np.cos(yy / 23.0) + 0.01 * rng.standard_normal((ry, rx))
            f[rng.random(f.shape) < 0.01] += 0.4
            f = np.clip(f, 1e-4, 1)
            f[rng.random(f.shape) < 0.03] = 0          # null pixels: dropped from BOTH sides of the pair
            f[: ry // 10] = 0                           # a black border, as registered frames have
            return np.round(f * 65535).astype(np.uint16) if dtype == np.uint16 else f.astype(np.float32)
        fi, fj = frame(rxi, ryi, 0.12), frame(rxj, ryj, 0.17)
        ai, aj = _lib.Rect(), _lib.Rect()
        n = L.sgpu_compute_overlap(rxi, ryi, rxj, ryj, dxi, dyi, dxj, dyj, C.byref(ai), C.byref(aj))
        assert n > 0
        # what readfits_partial_all_layers hands the reference: the rectangle's rows in memory (bottom-up) order
        reg = lambda f, a, ry: np.ascontiguousarray(f[ry - a.y - a.h: ry - a.y, a.x: a.x + a.w])
        exp = np.zeros(9)
        ri, rj = reg(fi, ai, ryi), reg(fj, aj, ryj)
        assert R.ref_overlap_pair(int(dtype == np.float32), ri.ctypes.data, rj.ctypes.data, n, lite, exp.ctypes.data) == 0
        if on_device:
            dev = lambda f: torch.from_numpy(f.view(np.int16) if f.dtype == np.uint16 else f).cuda()
            ti, tj = dev(fi), dev(fj)
            pi, pj = ti.data_ptr(), tj.data_ptr()
        else:
            pi, pj = fi.ctypes.data, fj.ctypes.data
        out = _lib.OverlapStats()
        ctx.check(L.sgpu_overlap_stats_pair(ctx.h, pi, rxi, ryi, pj, rxj, ryj, on_device, 1 if dtype == np.float32 else 0,
                                            C.byref(ai), C.byref(aj), lite, C.byref(out)))
        got = [out.Nij, out.medij, out.medji, out.madij, out.madji, out.locij, out.locji, out.scaij, out.scaji]
        assert got[0] == exp[0] and exp[0] > 1000
        assert got[1:7] == [np.float32(v) for v in exp[1:7]], (got, exp)       # histogram medians: bit-equal floats
        np.testing.assert_allclose(got[7:], exp[7:], rtol=3e-7)                  # bwmv double sums, then a float
        if lite:
            assert got[5:] == [0, 0, 0, 0]


@pytest.mark.gpu
def test_overlap_pair_with_too_few_common_pixels(ref):
    torch = pytest.importorskip("torch")
    from siril_b200 import _lib
    L = _lib.lib()
    ctx = _lib.Context(0)
    fi = np.zeros((20, 30), dtype=np.float32); fj = np.zeros((20, 30), dtype=np.float32)
    fi[3, 4:7] = 0.5; fj[3, 4:7] = 0.25          # three common pixels: "at least 3" means more than 3 (normalization.c:565)
    ai = _lib.Rect(0, 0, 30, 20)
    out = _lib.OverlapStats()
    ctx.check(L.sgpu_overlap_stats_pair(ctx.h, fi.ctypes.data, 30, 20, fj.ctypes.data, 30, 20, 0, 1, C.byref(ai), C.byref(ai), 0, C.byref(out)))
    assert out.Nij == 0 and out.medij == 0 and out.scaij == 0
    bad = _lib.Rect(5, 5, 30, 20)
    assert L.sgpu_overlap_stats_pair(ctx.h, fi.ctypes.data, 30, 20, fj.ctypes.data, 30, 20, 0, 1, C.byref(bad), C.byref(bad), 0, C.byref(out)) != 0


@pytest.mark.gpu
@pytest.mark.parametrize("normalize", [O.ADDITIVE_SCALING, O.MULTIPLICATIVE], ids=["addscale", "mul"])
def test_do_normalization_by_overlaps_levels_a_mosaic(ref, normalize):
    """do_normalization with args->overlap_norm (normalization.c:64-67): four panels of one sky, each with its own gain
    and pedestal, shifted against each other; coefficients against the reference's own lines driven pair by pair"""
    torch = pytest.importorskip("torch")
    import siril_b200 as sb
    from siril_b200 import _lib
    L, R = _lib.lib(), _ref_lib(ref)
    rng = np.random.default_rng(3)
    N, Cn, H, W = 4, 2, 120, 160
    yy, xx = np.mgrid[0:H + 100, 0:W + 100]
    sky = 0.1 + 0.05 * np.sin(xx / 29.0) * np.cos(yy / 31.0)
    dx, dy = np.array([0.0, 37.4, -12.6, 40.2]), np.array([0.0, -8.3, 33.7, 30.1])
    gain, ped = np.array([1.0, 1.15, 0.9, 1.05]), np.array([0.0, 0.01, -0.004, 0.02]) * (normalize != O.MULTIPLICATIVE)
    fr = np.zeros((N, Cn, H, W), dtype=np.float32)
    for f in range(N):      # frame f sees the sky displaced by its registration shift (x to the right, y up in display terms)
        ox, oy = 50 + int(round(dx[f])), 50 - int(round(dy[f]))   # x_j = x_i - (dx_j - dx_i), y likewise top-down (normalization.c:427-447)
        for l in range(Cn):
            view = sky[oy:oy + H, ox:ox + W][::-1] * (1 + 0.1 * l)
            fr[f, l] = (gain[f] * view + ped[f] + 0.004 * rng.standard_normal((H, W))).astype(np.float32)
    fr[rng.random(fr.shape) < 0.01] = 0
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=torch.from_numpy(fr).cuda(), normalize=normalize,
                           overlap_norm=True, reg_dx=dx, reg_dy=dy, ref_image=1)
    assert sb.do_normalization(args) == 0
    # the same through the reference: rectangles, per-pair statistics, solves
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    for l in range(Cn):
        Nij = np.zeros((N, N), dtype=np.uint64); Mij = np.zeros((N, N)); Sij = np.zeros((N, N))
        for i in range(N):
            for j in range(i + 1, N):
                ei, ej = np.zeros(4, dtype=np.int32), np.zeros(4, dtype=np.int32)
                n = R.ref_compute_overlap(W, H, W, H, dx[i], dy[i], dx[j], dy[j], ei.ctypes.data, ej.ctypes.data)
                assert n > 0
                reg = lambda f, a: np.ascontiguousarray(f[H - a[1] - a[3]: H - a[1], a[0]: a[0] + a[2]])
                o = np.zeros(9)
                ri, rj = reg(fr[i, l], ei), reg(fr[j, l], ej)
                R.ref_overlap_pair(1, ri.ctypes.data, rj.ctypes.data, n, 0, o.ctypes.data)
                Nij[i, j] = Nij[j, i] = int(o[0]); Mij[i, j], Mij[j, i], Sij[i, j], Sij[j, i] = o[5], o[6], o[7], o[8]
        exp = [np.zeros(N), np.ones(N), np.ones(N)]
        assert R.ref_overlap_solve(normalize, N, 1, p(Nij), p(Mij), p(Sij), *[p(a) for a in exp]) == 0
        for got, e in zip((args.coeff.offset[l], args.coeff.mul[l], args.coeff.scale[l]), exp):
            np.testing.assert_allclose(got, e, rtol=2e-6, atol=1e-9)   # scaij is a float of double sums (3e-7), then a solve
        if normalize == O.MULTIPLICATIVE:     # and the mosaic is levelled: mul_f = gain_ref / gain_f
            np.testing.assert_allclose(args.coeff.mul[l], gain[1] / gain, rtol=5e-3)
