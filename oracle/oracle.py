"""TEST INFRASTRUCTURE ONLY — ctypes front-end for the CPU oracle.

Two interchangeable back-ends with the same Python surface:
  Oracle("port")      -> oracle/liborc.so           (our plain-C restatement, stack_oracle.c)
  Oracle("reference") -> oracle/_ref/libsiril_ref.so (the reference's own sources, build_ref.py)
  Oracle("reference_fast") -> oracle/_ref/libsiril_ref_fast.so (same sources, -fopenmp incl. the
                        `omp simd` reductions; used only as the TIMED CPU baseline)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product (siril_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

# enum values: reference src/core/settings.h:34-53
NO_REJEC, PERCENTILE, SIGMA, MAD, SIGMEDIAN, WINSORIZED, LINEARFIT, GESDT = range(8)
NO_NORM, ADDITIVE, MULTIPLICATIVE, ADDITIVE_SCALING, MULTIPLICATIVE_SCALING = range(5)

STAT_FIELDS = ("total", "ngoodpix", "mean", "median", "sigma", "avgDev", "mad", "sqrtbwmv",
               "location", "scale", "min", "max", "normValue", "bgnoise")


class _StackParams(C.Structure):
    _fields_ = [
        ("is_float", C.c_int), ("N", C.c_int), ("C", C.c_int), ("H", C.c_int), ("W", C.c_int),
        ("is_mean", C.c_int), ("rejtype", C.c_int),
        ("sig", C.c_float * 2),
        ("crit", C.c_void_p),
        ("normalize", C.c_int),
        ("scale", C.c_void_p), ("offset", C.c_void_p), ("mul", C.c_void_p),
        ("weights", C.c_void_p),
        ("shiftx", C.c_void_p), ("shifty", C.c_void_p),
        ("output_norm", C.c_int), ("use_32bit_output", C.c_int),
        ("create_rejmaps", C.c_int), ("merge_rejmaps", C.c_int),
        ("nthreads", C.c_int),
        ("max_rows", C.c_long),
        ("source_bitpix", C.c_int),
        ("mask", C.c_void_p), ("drizz", C.c_void_p),
        ("feather_dist", C.c_int), ("drizzle", C.c_int),
    ]


def gesdt_critical_values(nb_frames, sig0, sig1):
    """Critical values of the generalized ESD test as the reference computes them
    (median_and_mean.c:1479-1486): the probability `1 - sig[1] / (2 * size)` is evaluated in
    FLOAT, the t quantile (GSL gsl_cdf_tdist_Pinv in the reference; scipy here — GSL is an
    unpinned system dependency of the reference) is rounded to float, and the rest is float."""
    from scipy.stats import t as tdist
    f32 = np.float32
    n_out = int(np.floor(f32(nb_frames) * f32(sig0)))     # int * float is a FLOAT product in the reference (median_and_mean.c:1479)
    out = np.empty(max(n_out, 0), dtype=np.float32)
    size = nb_frames
    for j in range(n_out):
        p = f32(1) - f32(sig1) / f32(2 * size)
        td = f32(tdist.ppf(float(p), size - 2))
        num = f32(size - 1) * td
        den = np.sqrt(f32(size)) * np.sqrt(f32(size - 2) + td * td)
        out[j] = f32(num / den)
        size -= 1
    return out


def ensure_built(kind="port"):
    if kind == "port":
        so = os.path.join(HERE, "liborc.so")
        src = os.path.join(HERE, "stack_oracle.c")
        if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
            subprocess.check_call(["make", "-C", HERE, "liborc.so"], stdout=subprocess.DEVNULL)
        return so
    so = os.path.join(HERE, "_ref", "libsiril_ref_fast.so" if kind == "reference_fast" else "libsiril_ref.so")
    if not os.path.exists(so) and os.path.isdir("/root/reference"):
        subprocess.check_call(["python3", os.path.join(HERE, "build_ref.py")], stdout=subprocess.DEVNULL)
    return so


def have_reference():
    return os.path.exists(os.path.join(HERE, "_ref", "libsiril_ref.so")) or os.path.isdir("/root/reference")


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Oracle:
    def __init__(self, kind="port"):
        self.kind = kind
        self.lib = C.CDLL(ensure_built(kind))
        self.px = "orc_" if kind == "port" else "ref_"
        L = self.lib
        f = getattr(L, self.px + "mean_and_reject")
        f.restype = C.c_double
        f.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p,
                      C.c_void_p, C.c_int, C.POINTER(C.c_int)]
        self._mar = f
        f = getattr(L, self.px + "stack")
        f.restype = C.c_int
        f.argtypes = [C.POINTER(_StackParams), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        self._stack = f
        f = getattr(L, self.px + "image_stats")
        f.restype = C.c_int
        f.argtypes = [C.c_int, C.c_void_p, C.c_long, C.c_long, C.c_int, C.POINTER(C.c_double)]
        self._stats = f
        f = getattr(L, self.px + "compute_factors")
        f.restype = None
        f.argtypes = [C.c_int] * 6 + [C.c_void_p] * 3
        self._factors = f
        f = getattr(L, self.px + "noise_weights")
        f.restype = None
        f.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        self._nw = f
        f = getattr(L, self.px + "planner")
        f.restype = C.c_int
        f.argtypes = [C.c_long, C.c_long, C.c_long, C.c_long, C.c_int, C.c_void_p, C.c_int,
                      C.POINTER(C.c_long), C.POINTER(C.c_int)]
        self._planner = f
        for nm in ("stack_sum", "stack_minmax"):
            f = getattr(L, self.px + nm)
            f.restype = C.c_int
            f.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        self._sum = getattr(L, self.px + "stack_sum")
        self._minmax = getattr(L, self.px + "stack_minmax")
        f = getattr(L, self.px + "norm_to_0_1")
        f.restype = None
        f.argtypes = [C.c_void_p, C.c_long, C.c_long, C.c_long]
        self._n01 = f
        # selection helpers (names differ between the two libs)
        if kind == "port":
            self._qmf = L.orc_quickmedian_float
            self._qm = L.orc_quickmedian
            self._hm = L.orc_histogram_median
            self._hmf = L.orc_histogram_median_float
            self._hm.argtypes = [C.c_void_p, C.c_size_t]
            self._hmf.argtypes = [C.c_void_p, C.c_size_t]
            self._hmf.restype = C.c_float
        else:
            self._qmf = L.quickmedian_float
            self._qm = L.quickmedian
            self._hm = L.histogram_median
            self._hmf = L.histogram_median_float
            self._hm.argtypes = [C.c_void_p, C.c_size_t, C.c_int]
            self._hmf.argtypes = [C.c_void_p, C.c_size_t, C.c_int]
            self._hmf.restype = C.c_double
        for f in (self._qmf, self._qm):
            f.restype = C.c_double
            f.argtypes = [C.c_void_p, C.c_size_t]
        self._hm.restype = C.c_double

    # -- per-stack -----------------------------------------------------------------------
    def mean_and_reject(self, stack, rejection, sig=(3.0, 3.0), crit=None, weights=None, layer=0):
        stack = np.ascontiguousarray(stack)
        is_float = stack.dtype == np.float32
        assert is_float or stack.dtype == np.uint16
        rej = (C.c_int * 2)()
        w = None if weights is None else np.ascontiguousarray(weights, dtype=np.float64)
        cr = None if crit is None else np.ascontiguousarray(crit, dtype=np.float32)
        m = self._mar(int(is_float), _ptr(stack), len(stack), rejection, sig[0], sig[1],
                      _ptr(cr), _ptr(w), layer, rej)
        return m, rej[0], rej[1]

    def quickmedian(self, a):
        a = np.array(a, copy=True)
        f = self._qmf if a.dtype == np.float32 else self._qm
        return f(_ptr(a), len(a)), a

    def histogram_median(self, a):
        a = np.array(a, copy=True)
        if a.dtype == np.float32:
            r = self._hmf(_ptr(a), len(a)) if self.kind == "port" else self._hmf(_ptr(a), len(a), 1)
            return float(np.float32(r))
        return self._hm(_ptr(a), len(a)) if self.kind == "port" else self._hm(_ptr(a), len(a), 1)

    # -- whole stack ---------------------------------------------------------------------
    def stack(self, frames, *, is_mean=True, rejection=NO_REJEC, sig=(3.0, 3.0), crit=None,
              normalize=NO_NORM, scale=None, offset=None, mul=None, weights=None,
              shiftx=None, shifty=None, output_norm=False, use_32bit_output=True,
              create_rejmaps=False, merge_rejmaps=False, nthreads=None, max_rows=0, source_bitpix=0,
              masks=None, feather_dist=0, drizzle_weights=None):
        """frames: ndarray [N][C][H][W] float32|uint16 (memory / FITS row order).
        Returns dict(out, rej_low, rej_high, irej[3][2])."""
        frames = np.ascontiguousarray(frames)
        N, Cn, H, W = frames.shape
        p = _StackParams()
        p.is_float = int(frames.dtype == np.float32)
        p.N, p.C, p.H, p.W = N, Cn, H, W
        p.is_mean, p.rejtype = int(is_mean), rejection
        p.sig[0], p.sig[1] = sig
        keep = []

        def arr(a, dt):
            if a is None:
                return None
            a = np.ascontiguousarray(a, dtype=dt)
            keep.append(a)
            return _ptr(a)
        if rejection == GESDT and crit is None:
            crit = gesdt_critical_values(N, sig[0], sig[1])
        p.crit = arr(crit, np.float32)
        p.normalize = normalize
        p.scale, p.offset, p.mul = arr(scale, np.float64), arr(offset, np.float64), arr(mul, np.float64)
        p.weights = arr(weights, np.float64)
        p.shiftx, p.shifty = arr(shiftx, np.int32), arr(shifty, np.int32)
        p.output_norm, p.use_32bit_output = int(output_norm), int(use_32bit_output)
        p.create_rejmaps, p.merge_rejmaps = int(create_rejmaps), int(merge_rejmaps)
        # (the reference's planner is made for images far taller than the thread count: with more threads than
        # rows it plans blocks past the image; the result does not depend on the thread count, so cap it)
        p.nthreads = max(1, min(nthreads or os.cpu_count() or 1, H))
        p.max_rows = max_rows
        p.source_bitpix = source_bitpix
        # feather masks [N][H][W] (distances, before the ramp) / drizzle weights [N][C][H][W] (median_and_mean.c:483-542)
        p.mask, p.feather_dist = arr(masks, np.float32), int(feather_dist if masks is not None else 0)
        p.drizz, p.drizzle = arr(drizzle_weights, np.float32), int(drizzle_weights is not None)
        out = np.zeros((Cn, H, W), dtype=np.float32 if use_32bit_output else np.uint16)
        rl = np.zeros((Cn, H, W), dtype=np.uint16) if create_rejmaps else None
        rh = np.zeros((Cn, H, W), dtype=np.uint16) if (create_rejmaps and not merge_rejmaps) else None
        irej = np.zeros(6, dtype=np.uint64)
        rv = self._stack(C.byref(p), _ptr(frames), _ptr(out), _ptr(rl), _ptr(rh), _ptr(irej))
        if rv:
            raise RuntimeError(f"oracle stack failed: {rv}")
        return dict(out=out, rej_low=rl, rej_high=rh, irej=irej.reshape(3, 2))

    # -- normalization -------------------------------------------------------------------
    def stack_simple(self, frames, op, shiftx=None, shifty=None, output_32bits=None):
        """sum.c / stackminmax.c on frames [N][C][H][W]; op in ('sum', 'max', 'min').
        Returns the [C][H][W] result, or None when the reference reports an error (all-zero sum)."""
        frames = np.ascontiguousarray(frames)
        N, Cn, H, W = frames.shape
        is_float = frames.dtype == np.float32
        sx = None if shiftx is None else np.ascontiguousarray(shiftx, dtype=np.int32)
        sy = None if shifty is None else np.ascontiguousarray(shifty, dtype=np.int32)
        if op == "sum":
            o32 = is_float if output_32bits is None else bool(output_32bits)
            out = np.zeros((Cn, H, W), dtype=np.float32 if o32 else np.uint16)
            rv = self._sum(int(is_float), _ptr(frames), N, Cn, H, W, _ptr(sx), _ptr(sy), int(o32), _ptr(out))
        else:
            out = np.zeros((Cn, H, W), dtype=frames.dtype)
            rv = self._minmax(int(is_float), _ptr(frames), N, Cn, H, W, _ptr(sx), _ptr(sy), int(op == "max"), _ptr(out))
        return None if rv else out

    def image_stats(self, plane, lite=False):
        plane = np.ascontiguousarray(plane)
        ny, nx = plane.shape
        out = (C.c_double * 14)()
        rv = self._stats(int(plane.dtype == np.float32), _ptr(plane), nx, ny, int(lite), out)
        if rv:
            raise RuntimeError(f"oracle image_stats failed: {rv}")
        return dict(zip(STAT_FIELDS, list(out)))

    def compute_factors(self, normalize, offset, mul, scale, ref_index, reglayer=-1, equalize_rgb=False):
        """arrays [C][N] of estimators; returns (offset, mul, scale) factor arrays"""
        offset, mul, scale = (np.array(a, dtype=np.float64, copy=True) for a in (offset, mul, scale))
        nl, n = offset.shape
        self._factors(normalize, nl, n, ref_index, reglayer, int(equalize_rgb), _ptr(offset), _ptr(mul), _ptr(scale))
        return offset, mul, scale

    def noise_weights(self, scale, bgnoise):
        scale = np.ascontiguousarray(scale, dtype=np.float64)
        bgnoise = np.ascontiguousarray(bgnoise, dtype=np.float64)
        nl, n = scale.shape
        w = np.zeros((nl, n), dtype=np.float64)
        self._nw(nl, n, _ptr(scale), _ptr(bgnoise), _ptr(w))
        return w

    def planner(self, max_rows, W, H, Cn, nb_threads):
        blocks = np.zeros((4096, 4), dtype=np.int64)
        largest, nb = C.c_long(0), C.c_int(0)
        rv = self._planner(max_rows, W, H, Cn, nb_threads, _ptr(blocks), 4096, C.byref(largest), C.byref(nb))
        return rv, blocks[:nb.value].copy(), largest.value

    def norm_to_0_1(self, img):
        img = np.array(img, dtype=np.float32, copy=True)
        Cn, H, W = img.shape
        self._n01(_ptr(img), W, H, Cn)
        return img
