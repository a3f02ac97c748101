"""Host-side mirror of the reference's stacking interface for the GPU path.

Names, argument meaning and error behaviour follow the reference:
  struct stacking_args / norm_coeff        src/stacking/stacking.h:55-117
  main_stack                               src/stacking/stacking.c:76-108
  stack_mean_with_rejection / stack_median src/stacking/median_and_mean.c:1103-1109
  do_normalization                         src/stacking/normalization.c:46-78
All computation happens in libsgpu.so (CUDA); this file only marshals arguments.
"""
import os
import ctypes as C
from dataclasses import dataclass, field
from typing import Any, Callable, Optional

import numpy as np

from . import _lib
from .enums import *  # noqa: F401,F403
from .enums import (ADDITIVE, ADDITIVE_SCALING, GESDT, MULTIPLICATIVE, MULTIPLICATIVE_SCALING, NO_NORM, NO_REJEC,
                    NO_WEIGHT, NOISE_WEIGHT, ST_GENERIC_ERROR, ST_OK)


def gesdt_critical_values(nb_frames, sig0, sig1):
    """args->critical_value as stack_mean_or_median fills it (median_and_mean.c:1479-1486):
    t = tdist_Pinv(1 - sig[1] / (2 * size), size - 2) with the probability evaluated in FLOAT,
    then float arithmetic.  The reference uses GSL's quantile; this harness uses scipy's."""
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


@dataclass
class NormCoeff:
    """norm_coeff (stacking.h:55-62): [layer][frame] doubles"""
    offset: Optional[np.ndarray] = None
    mul: Optional[np.ndarray] = None
    scale: Optional[np.ndarray] = None


@dataclass
class StackResult:
    """what the reference leaves in args->result / rejmap_low / rejmap_high / the log counters"""
    data: Any = None          # [C][H][W] float32 or uint16 (torch CUDA tensor or numpy array)
    rejmap_low: Any = None
    rejmap_high: Any = None
    irej: Optional[np.ndarray] = None   # [3][2] uint64: low/high rejections per channel (:1766-1772)


@dataclass
class StackingArgs:
    """struct stacking_args restricted to what the hot path consumes (stacking.h:65-117).

    `frames` replaces seq + image_indices: the selected frames as ONE array [N][C][H][W] in memory
    (FITS) row order — a torch CUDA tensor (device-resident: the timed `value` path) or a numpy
    array / list of numpy frames on the host (the end-to-end path: H2D + stack + D2H)."""
    method: Optional[Callable[["StackingArgs"], int]] = None
    frames: Any = None
    ref_image: int = 0
    lite_norm: bool = False
    normalize: int = NO_NORM
    coeff: NormCoeff = field(default_factory=NormCoeff)
    force_norm: bool = False
    output_norm: bool = False
    use_32bit_output: bool = True
    reglayer: int = -1
    equalizeRGB: bool = False
    shiftx: Optional[np.ndarray] = None   # round_to_int(dx) per frame (median_and_mean.c:1621-1623)
    shifty: Optional[np.ndarray] = None   # round_to_int(dy) per frame (:416-418)
    type_of_rejection: int = NO_REJEC
    sig: tuple = (3.0, 3.0)
    critical_value: Optional[np.ndarray] = None
    create_rejmaps: bool = False
    merge_lowhigh_rejmaps: bool = False
    weighting_type: int = NO_WEIGHT
    weights: Optional[np.ndarray] = None
    weighted_fwhm: Optional[np.ndarray] = None    # regparam[reglayer][idx].weighted_fwhm of the selected frames (WFWHM_WEIGHT)
    number_of_stars: Optional[np.ndarray] = None  # regparam[reglayer][idx].number_of_stars (NBSTARS_WEIGHT)
    stack_count: Optional[np.ndarray] = None      # STACKCNT keyword of each frame (NBSTACK_WEIGHT, median_and_mean.c:150-159)
    bgnoise: Optional[np.ndarray] = None  # seq->stats[layer][idx]->bgnoise, filled by do_normalization
    overlap_norm: bool = False      # args->overlap_norm (with maximize_framing): normalization by the frames' overlaps (normalization.c:64-67)
    reg_dx: Optional[np.ndarray] = None   # translation_from_H of every frame (registration.c:301-304), unrounded: the overlaps
    reg_dy: Optional[np.ndarray] = None
    overlap_stats: Any = None       # filled by compute_normalization_overlaps: {(layer, i, j): OverlapStats} (seq->ostats)
    feather_dist: int = 0           # > 0: masking (stacking.h:112); `masks` then holds the up-scaled distance masks
    masks: Any = None               # [N][H][W] float32 (torch CUDA tensor or numpy), data->mask[frame] before the ramp (:483-521)
    drizzle: bool = False           # args->drizzle; `drizzle_weights` [N][C][H][W] float32 = data->drizz[frame] (:523-541)
    drizzle_weights: Any = None
    source_bitpix: int = 0          # seq->bitpix of the frames on disk; 8: normalize_to16bit with a 16-bit output_norm result
    engine: int = 0
    retval: int = ST_OK
    result: StackResult = field(default_factory=StackResult)
    ctx: Optional[_lib.Context] = None

    @property
    def nb_images_to_stack(self):
        return len(self.frames)


def _is_torch_cuda(x):
    return type(x).__module__.startswith("torch") and getattr(x, "is_cuda", False)


def _context(args):
    if args.ctx is None:
        dev, stream = 0, None
        if _is_torch_cuda(args.frames):
            import torch
            dev = args.frames.device.index or 0
            stream = torch.cuda.current_stream(args.frames.device).cuda_stream
        args.ctx = _lib.Context(dev, stream)
    return args.ctx


def _shape_dtype(frames):
    if _is_torch_cuda(frames):
        import torch
        n, c, h, w = frames.shape
        if frames.dtype == torch.float32:
            return n, c, h, w, DATA_FLOAT
        if frames.dtype in (torch.uint16, torch.int16):
            return n, c, h, w, DATA_USHORT
        raise TypeError(f"unsupported frame dtype {frames.dtype}")
    f0 = frames[0]
    c, h, w = f0.shape
    dt = DATA_FLOAT if f0.dtype == np.float32 else DATA_USHORT
    if f0.dtype not in (np.float32, np.uint16):
        raise TypeError(f"unsupported frame dtype {f0.dtype}")
    return len(frames), c, h, w, dt


def _fill_cargs(args, n, c, is_mean, keep):
    a = _lib.StackArgs()
    a.is_mean = int(is_mean)
    a.type_of_rejection = args.type_of_rejection
    a.sig[0], a.sig[1] = float(args.sig[0]), float(args.sig[1])

    def arr(x, dt, count):
        if x is None:
            return None
        x = np.ascontiguousarray(x, dtype=dt).reshape(-1)
        if x.size != count:
            raise ValueError(f"array has {x.size} entries, expected {count}")
        keep.append(x)
        return x.ctypes.data_as(C.c_void_p)
    if is_mean and args.type_of_rejection == GESDT:
        if args.critical_value is None:   # stack_mean_or_median computes it itself (:1479-1486)
            args.critical_value = gesdt_critical_values(n, args.sig[0], args.sig[1])
        a.critical_value = arr(args.critical_value, np.float32, len(args.critical_value))
        a.nb_critical = len(args.critical_value)
    a.normalize = args.normalize
    if args.normalize != NO_NORM:
        a.scale = arr(args.coeff.scale, np.float64, n * c)
        a.offset = arr(args.coeff.offset, np.float64, n * c)
        a.mul = arr(args.coeff.mul, np.float64, n * c)
    # the median path ignores weights (src/core/command.c:11514-11515, 11660)
    if is_mean and args.weighting_type != NO_WEIGHT:
        if args.weights is None:
            _compute_weights(args, n, c)
        a.weights = arr(args.weights, np.float64, n * c)
    a.shiftx = arr(args.shiftx, np.int32, n)
    a.output_norm = int(args.output_norm)
    a.use_32bit_output = int(args.use_32bit_output)
    a.create_rejmaps = int(args.create_rejmaps)
    a.merge_lowhigh_rejmaps = int(args.merge_lowhigh_rejmaps)
    a.engine = args.engine
    a.source_bitpix = args.source_bitpix
    return a


def _compute_weights(args, n, c):
    """median_and_mean.c:1519-1539: NOISE_WEIGHT from the normalization statistics, the others from the
    registration data / STACKCNT keywords the caller holds for the selected frames"""
    L = _lib.lib()
    w = np.zeros((c, n), dtype=np.float64)
    for kind, src, dt, fn, what in ((WFWHM_WEIGHT, args.weighted_fwhm, np.float32, L.sgpu_wfwhm_weights, "wFWHM"),
                                    (NBSTARS_WEIGHT, args.number_of_stars, np.int32, L.sgpu_nbstars_weights, "number of stars"),
                                    (NBSTACK_WEIGHT, args.stack_count, np.int32, L.sgpu_nbstack_weights, "stack count")):
        if args.weighting_type == kind:
            if src is None:    # :1144-1147 / :1192-1195
                raise _lib.SgpuError(ST_GENERIC_ERROR, f"Sequence does not have registration info, cannot use weighing by {what}, aborting")
            a = np.ascontiguousarray(src, dtype=dt)
            if a.size != n:
                raise ValueError(f"{what}: {a.size} entries for {n} frames")
            rv = fn(c, n, a.ctypes.data_as(C.c_void_p), w.ctypes.data_as(C.c_void_p))
            if rv:
                raise _lib.SgpuError(rv, f"weights by {what}")
            args.weights = w
            return
    if args.weighting_type != NOISE_WEIGHT:
        raise _lib.SgpuError(ST_GENERIC_ERROR, f"unknown weighting type {args.weighting_type}")
    if args.normalize == NO_NORM or args.bgnoise is None:
        # src/core/command.c:11700-11703: noise weighting needs normalization statistics
        raise _lib.SgpuError(ST_GENERIC_ERROR, "Weighting by noise requires normalization")
    scale = np.ascontiguousarray(args.coeff.scale, dtype=np.float64).reshape(c, n)
    bg = np.ascontiguousarray(args.bgnoise, dtype=np.float64).reshape(c, n)
    w = np.zeros((c, n), dtype=np.float64)
    rv = _lib.lib().sgpu_noise_weights(c, n, scale.ctypes.data_as(C.c_void_p), bg.ctypes.data_as(C.c_void_p),
                                       w.ctypes.data_as(C.c_void_p))
    if rv:
        raise _lib.SgpuError(rv, "sgpu_noise_weights")
    args.weights = w


def stack_mean_or_median(args: StackingArgs, is_mean: bool) -> int:
    """stack_mean_or_median (median_and_mean.c:1261-1826) on the GPU."""
    L = _lib.lib()
    ctx = _context(args)
    n, c, h, w, dt = _shape_dtype(args.frames)
    keep = []
    try:
        ca = _fill_cargs(args, n, c, is_mean, keep)
    except _lib.SgpuError as e:
        args.retval = e.code
        return e.code
    desc = _lib.BlockDesc(n, c, w, h, dt)
    out32 = bool(args.use_32bit_output)
    nmaps = (1 if args.merge_lowhigh_rejmaps else 2) if (is_mean and args.create_rejmaps) else 0
    res = StackResult()
    if _is_torch_cuda(args.frames):
        import torch
        dev = args.frames.device
        frames = args.frames.contiguous()
        blk = C.c_void_p()
        shifted = args.shifty is not None and np.any(np.asarray(args.shifty) != 0)
        masking = is_mean and args.feather_dist > 0
        drizzle = is_mean and bool(args.drizzle)
        if (masking and args.masks is None) or (drizzle and args.drizzle_weights is None):
            args.retval = ST_SEQUENCE_ERROR     # the reference fails reading the mask / weight areas (:497-501, :529-532)
            return ST_SEQUENCE_ERROR
        try:
            if shifted or masking or drizzle:     # the reference applies the y shift when it loads the block (:407-448)
                ctx.check(L.sgpu_block_create(ctx.h, C.byref(desc), C.byref(blk)))
                es = frames.element_size()
                for f in range(n):
                    ctx.check(L.sgpu_block_upload_frame(blk, f, C.c_void_p(frames.data_ptr() + f * c * h * w * es), h, 0,
                                                        int(args.shifty[f]) if shifted else 0))
                for on, planes, kind, per in ((masking, args.masks, _lib.PLANE_MASK, 1),
                                              (drizzle, args.drizzle_weights, _lib.PLANE_DRIZZLE, c)):
                    if not on:
                        continue    # the planes travel like the frames: same rows, same y shift (:483-541)
                    pl = planes if _is_torch_cuda(planes) else torch.from_numpy(np.ascontiguousarray(planes, dtype=np.float32)).to(dev)
                    pl = pl.contiguous().float()
                    if pl.numel() != n * per * h * w:
                        raise ValueError(f"weight planes have {pl.numel()} samples, expected {n * per * h * w}")
                    keep.append(pl)
                    torch.cuda.current_stream(dev).synchronize()
                    for f in range(n):
                        ctx.check(L.sgpu_block_upload_weight_plane(blk, kind, f, C.c_void_p(pl.data_ptr() + f * per * h * w * 4), h, 0,
                                                                   int(args.shifty[f]) if shifted else 0))
                if masking:
                    ctx.check(L.sgpu_block_mask_ramp(blk, int(args.feather_dist)))
            else:
                ctx.check(L.sgpu_block_wrap(ctx.h, C.byref(desc), C.c_void_p(frames.data_ptr()), C.byref(blk)))
            out = torch.empty((c, h, w), dtype=torch.float32 if out32 else torch.uint16, device=dev)
            rl = torch.zeros((c, h, w), dtype=torch.uint16, device=dev) if nmaps >= 1 else None
            rh = torch.zeros((c, h, w), dtype=torch.uint16, device=dev) if nmaps == 2 else None
            irej = torch.zeros(6, dtype=torch.int64, device=dev)
            rv = L.sgpu_stack_block(ctx.h, blk, C.byref(ca), C.c_void_p(out.data_ptr()),
                                    C.c_void_p(rl.data_ptr()) if rl is not None else None,
                                    C.c_void_p(rh.data_ptr()) if rh is not None else None,
                                    C.c_void_p(irej.data_ptr()))
            if rv == 0 and out32 and args.output_norm:   # :1774-1775
                mm = (C.c_float * 2)()
                rv = L.sgpu_minmax_nonzero(ctx.h, C.c_void_p(out.data_ptr()), out.numel(), 1, mm)
                if rv == 0:
                    rv = L.sgpu_rescale_0_1(ctx.h, C.c_void_p(out.data_ptr()), out.numel(), mm[0], mm[1])
            if rv:
                args.retval = rv
                args.error = (L.sgpu_last_error(ctx.h) or b"").decode()
                return rv
            res.data, res.rejmap_low, res.rejmap_high = out, rl, rh
            res.irej = irej.cpu().numpy().astype(np.uint64).reshape(3, 2)
        finally:
            ctx.synchronize()
            if blk:
                L.sgpu_block_destroy(blk)
    else:
        if is_mean and (args.feather_dist > 0 or args.drizzle):
            raise _lib.SgpuError(ST_GENERIC_ERROR, "mask / drizzle weight planes go with device-resident frames (sgpu_stack_block)")
        frames = [np.ascontiguousarray(f) for f in args.frames]
        ptrs = (C.c_void_p * n)(*[f.ctypes.data_as(C.c_void_p) for f in frames])
        out = np.empty((c, h, w), dtype=np.float32 if out32 else np.uint16)
        rl = np.zeros((c, h, w), dtype=np.uint16) if nmaps >= 1 else None
        rh = np.zeros((c, h, w), dtype=np.uint16) if nmaps == 2 else None
        irej = np.zeros(6, dtype=np.uint64)
        sy = None
        if args.shifty is not None:
            sy = np.ascontiguousarray(args.shifty, dtype=np.int32)
            keep.append(sy)
        rv = L.sgpu_stack_host(ctx.h, C.byref(desc), ptrs, sy.ctypes.data_as(C.c_void_p) if sy is not None else None,
                               C.byref(ca), out.ctypes.data_as(C.c_void_p),
                               rl.ctypes.data_as(C.c_void_p) if rl is not None else None,
                               rh.ctypes.data_as(C.c_void_p) if rh is not None else None,
                               irej.ctypes.data_as(C.c_void_p))
        if rv:
            args.retval = rv
            args.error = (L.sgpu_last_error(ctx.h) or b"").decode()
            return rv
        res.data, res.rejmap_low, res.rejmap_high, res.irej = out, rl, rh, irej.reshape(3, 2)
    args.result = res
    # the callee frees coeff and weights (median_and_mean.c:1799-1805); mirrored as "consumed"
    args.retval = ST_OK
    return ST_OK


def stack_mean_with_rejection(args: StackingArgs) -> int:
    return stack_mean_or_median(args, True)


def stack_median(args: StackingArgs) -> int:
    return stack_mean_or_median(args, False)


def _stack_simple(args: StackingArgs, op: int) -> int:
    """the sum / max / min stackers (sum.c:375-441, stackminmax.c:186-306) on the GPU: frames as for
    stack_mean_or_median; shiftx / shifty are the rounded registration shifts (sum.c:160-173)"""
    import torch
    L = _lib.lib()
    host = not _is_torch_cuda(args.frames)
    frames = args.frames
    if host:        # host frames go up through torch (plumbing); the result comes back as numpy
        arr = np.ascontiguousarray(np.stack([np.asarray(f) for f in frames]))
        if arr.dtype not in (np.float32, np.uint16):
            raise TypeError(f"unsupported frame dtype {arr.dtype}")
        t = torch.from_numpy(arr.view(np.int16) if arr.dtype == np.uint16 else arr).cuda()
        frames = t.view(torch.uint16) if arr.dtype == np.uint16 else t
        torch.cuda.current_stream().synchronize()
    ctx = _context(args)
    n, c, h, w, dt = _shape_dtype(frames)
    is_float = dt == DATA_FLOAT
    out32 = bool(args.use_32bit_output)
    if op == _lib.SIMPLE_SUM and is_float and not out32:
        raise ValueError("a 32-bit sequence is summed to a 32-bit image (sum.c:399-400)")
    frames = frames.contiguous()
    dev = frames.device
    desc = _lib.BlockDesc(n, c, w, h, dt)
    sx = None if args.shiftx is None else np.ascontiguousarray(args.shiftx, dtype=np.int32)
    sy = None if args.shifty is None else np.ascontiguousarray(args.shifty, dtype=np.int32)
    blk = C.c_void_p()
    ptr = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
    try:
        if sy is not None and np.any(sy != 0):
            ctx.check(L.sgpu_block_create(ctx.h, C.byref(desc), C.byref(blk)))
            es = frames.element_size()
            for f in range(n):
                ctx.check(L.sgpu_block_upload_frame(blk, f, C.c_void_p(frames.data_ptr() + f * c * h * w * es), h, 0, int(sy[f])))
        else:
            ctx.check(L.sgpu_block_wrap(ctx.h, C.byref(desc), C.c_void_p(frames.data_ptr()), C.byref(blk)))
        if op == _lib.SIMPLE_SUM:
            acc = torch.empty((c, h, w), dtype=torch.float64 if is_float else torch.int64, device=dev)
            mx = C.c_double(0.0)
            rv = L.sgpu_simple_block(ctx.h, blk, op, ptr(sx), ptr(sy), 0, h, C.c_void_p(acc.data_ptr()), C.byref(mx))
            out = torch.empty((c, h, w), dtype=torch.float32 if out32 else torch.uint16, device=dev)
            if rv == 0:
                rv = L.sgpu_sum_convert(ctx.h, C.c_void_p(acc.data_ptr()), int(is_float), acc.numel(), mx.value,
                                        int(out32), C.c_void_p(out.data_ptr()))
        else:
            out = torch.empty((c, h, w), dtype=frames.dtype, device=dev)
            rv = L.sgpu_simple_block(ctx.h, blk, op, ptr(sx), ptr(sy), 0, h, C.c_void_p(out.data_ptr()), None)
        ctx.synchronize()
        if rv:
            args.retval = rv
            args.error = (L.sgpu_last_error(ctx.h) or b"").decode()
            return rv
    finally:
        ctx.synchronize()
        if blk:
            L.sgpu_block_destroy(blk)
    res = StackResult()
    if host:
        o = out.cpu()
        res.data = (o.view(torch.int16).numpy().view(np.uint16) if o.dtype == torch.uint16 else o.numpy())
    else:
        res.data = out
    args.result = res
    args.retval = ST_OK
    return ST_OK


def stack_summing_generic(args: StackingArgs) -> int:
    """stack_summing_generic (sum.c:375-441): sum of the frames, scaled by its maximum"""
    return _stack_simple(args, _lib.SIMPLE_SUM)


def stack_addmax(args: StackingArgs) -> int:
    """stack_addmax (stackminmax.c:301-303): per-pixel maximum (starts from 0)"""
    return _stack_simple(args, _lib.SIMPLE_MAX)


def stack_addmin(args: StackingArgs) -> int:
    """stack_addmin (stackminmax.c:305-306): per-pixel minimum (starts from 1.0 / 65535)"""
    return _stack_simple(args, _lib.SIMPLE_MIN)


def image_stats(ctx, plane, lite=False):
    """statistics(..., STATS_NORM | STATS_LITENORM) of one plane (torch CUDA tensor or numpy)."""
    L = _lib.lib()
    st = _lib.ImStats()
    if _is_torch_cuda(plane):
        import torch
        plane = plane.contiguous()
        ny, nx = plane.shape
        dt = DATA_FLOAT if plane.dtype == torch.float32 else DATA_USHORT
        ctx.check(L.sgpu_image_stats(ctx.h, C.c_void_p(plane.data_ptr()), 1, dt, nx, ny, int(lite), C.byref(st)))
    else:
        plane = np.ascontiguousarray(plane)
        ny, nx = plane.shape
        dt = DATA_FLOAT if plane.dtype == np.float32 else DATA_USHORT
        ctx.check(L.sgpu_image_stats(ctx.h, plane.ctypes.data_as(C.c_void_p), 0, dt, nx, ny, int(lite), C.byref(st)))
    return st


STATS_WORKERS = int(os.environ.get("SGPU_STATS_WORKERS", "4"))     # concurrent per-plane statistics (contexts / streams) in compute_estimators
_stats_pool = {}       # device index -> worker contexts


def result_bgnoise(args: StackingArgs):
    """background noise of every layer of the stacked result (FnNoise1 through K3's kernels): what
    the reference reports after stacking (bgnoise_image_hook, command.c:6901-6910)"""
    ctx = _context(args)
    data = args.result.data
    return [image_stats(ctx, data[layer], lite=2).bgnoise for layer in range(data.shape[0])]


def image_stats_sharded(ctx, shard, row0, full_height, reducer, lite=False):
    """statistics of a frame plane of which this rank holds rows [row0, row0 + shard rows): every
    rank calls it for the same plane and gets the WHOLE frame's statistics (sgpu_image_stats_sharded)"""
    L = _lib.lib()
    st = _lib.ImStats()
    if _is_torch_cuda(shard):
        import torch
        shard = shard.contiguous()
        ny, nx = shard.shape
        dt = DATA_FLOAT if shard.dtype == torch.float32 else DATA_USHORT
        ptr, on_dev = C.c_void_p(shard.data_ptr()), 1
    else:
        shard = np.ascontiguousarray(shard)
        ny, nx = shard.shape
        dt = DATA_FLOAT if shard.dtype == np.float32 else DATA_USHORT
        ptr, on_dev = shard.ctypes.data_as(C.c_void_p), 0
    ctx.check(L.sgpu_image_stats_sharded(ctx.h, ptr, on_dev, dt, nx, ny, row0, full_height, int(lite),
                                         C.cast(reducer, C.c_void_p), None, C.byref(st)))
    return st


def compute_estimators(args: StackingArgs, frame_indices=None, shard=None):
    """_compute_estimators_for_image over the given frames (all by default): returns
    (offset, mul, scale, bgnoise) arrays [C][N] holding ESTIMATORS (normalization.c:98-141).

    shard = (reducer, row0, full_height): args.frames holds only rows [row0, row0 + h) of every frame
    (one process per GPU); every rank then walks every plane, one after the other and in the same
    order, and the reductions are all-reduced (sharding.make_reducer)."""
    L = _lib.lib()
    ctx = _context(args)
    n, c, h, w, dt = _shape_dtype(args.frames)
    off = np.zeros((c, n)); mul = np.ones((c, n)); sc = np.ones((c, n)); bg = np.zeros((c, n))
    todo = [(f, layer) for f in (range(n) if frame_indices is None else frame_indices) for layer in range(c)]
    if shard is not None:
        reducer, row0, full_h = shard
        for f, layer in todo:
            st = image_stats_sharded(ctx, args.frames[f][layer], row0, full_h, reducer, args.lite_norm)
            o, m, s = C.c_double(), C.c_double(), C.c_double()
            L.sgpu_estimators_from_stats(args.normalize, int(args.lite_norm), C.byref(st), C.byref(o), C.byref(m), C.byref(s))
            off[layer, f], mul[layer, f], sc[layer, f], bg[layer, f] = o.value, m.value, s.value, st.bgnoise
        return off, mul, sc, bg

    def one(wctx, f, layer):
        st = image_stats(wctx, args.frames[f][layer], args.lite_norm)
        o, m, s = C.c_double(), C.c_double(), C.c_double()
        L.sgpu_estimators_from_stats(args.normalize, int(args.lite_norm), C.byref(st), C.byref(o), C.byref(m), C.byref(s))
        off[layer, f], mul[layer, f], sc[layer, f], bg[layer, f] = o.value, m.value, s.value, st.bgnoise

    workers = min(STATS_WORKERS, len(todo)) if _is_torch_cuda(args.frames) else 1
    if workers <= 1:
        for f, layer in todo:
            one(ctx, f, layer)
        return off, mul, sc, bg
    # frame-parallel like the reference (compute_normalization runs its frame loop under OpenMP,
    # normalization.c:239-265): one private context (stream + scratch) per worker thread, so the
    # host round trips of one plane's passes overlap the kernels of the others.  Every plane's
    # statistics are independent and deterministic, so the result does not depend on the schedule.
    import queue
    import threading
    ctx.synchronize()
    dev = args.frames.device.index or 0
    pool = _stats_pool.setdefault(dev, [])
    while len(pool) < workers:
        pool.append(_lib.Context(dev, None))
    q = queue.Queue()
    for item in todo:
        q.put(item)
    errors = []

    def run(wctx):
        try:
            while True:
                try:
                    f, layer = q.get_nowait()
                except queue.Empty:
                    return
                one(wctx, f, layer)
        except Exception as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=run, args=(pool[i],)) for i in range(workers)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if errors:
        raise errors[0]
    return off, mul, sc, bg


def finalize_factors(args: StackingArgs, off, mul, sc):
    L = _lib.lib()
    c, n = off.shape
    off, mul, sc = (np.ascontiguousarray(a, dtype=np.float64) for a in (off, mul, sc))
    rv = L.sgpu_compute_factors(args.normalize, c, n, args.ref_image, args.reglayer, int(args.equalizeRGB),
                                off.ctypes.data_as(C.c_void_p), mul.ctypes.data_as(C.c_void_p), sc.ctypes.data_as(C.c_void_p))
    if rv:
        raise _lib.SgpuError(rv, "sgpu_compute_factors")
    args.coeff = NormCoeff(offset=off, mul=mul, scale=sc)


def compute_normalization_overlaps(args: StackingArgs) -> int:
    """compute_normalization_overlaps (normalization.c:666-938) on device-resident frames [N][C][H][W]: every pair of
    frames is levelled on the pixels both have (sgpu_compute_overlap + sgpu_overlap_stats_pair per layer, the
    reference's `_compute_estimators_for_images`), then one linear system per layer (sgpu_overlap_solve)."""
    L = _lib.lib()
    ctx = _context(args)
    if not _is_torch_cuda(args.frames):
        raise _lib.SgpuError(ST_GENERIC_ERROR, "overlap normalization takes device-resident frames")
    frames = args.frames.contiguous()
    n, c, h, w, dt = _shape_dtype(frames)
    if not 0 <= args.ref_image < n:     # :690-695
        raise _lib.SgpuError(ST_GENERIC_ERROR, "The reference image is not in the selected set of images.")
    dx = np.zeros(n) if args.reg_dx is None else np.asarray(args.reg_dx, dtype=np.float64)
    dy = np.zeros(n) if args.reg_dy is None else np.asarray(args.reg_dy, dtype=np.float64)
    es = frames.element_size()
    Nij = np.zeros((c, n, n), dtype=np.uint64)
    Mij, Sij = np.zeros((c, n, n)), np.zeros((c, n, n))
    args.overlap_stats = {}
    for i in range(n):
        for j in range(i + 1, n):
            ai, aj = _lib.Rect(), _lib.Rect()
            if not L.sgpu_compute_overlap(w, h, w, h, dx[i], dy[i], dx[j], dy[j], C.byref(ai), C.byref(aj)):
                continue    # "No overlap between image %d and %d"
            for layer in range(c):
                st = _lib.OverlapStats()
                ctx.check(L.sgpu_overlap_stats_pair(ctx.h, frames.data_ptr() + ((i * c + layer) * h * w) * es, w, h,
                                                    frames.data_ptr() + ((j * c + layer) * h * w) * es, w, h, 1, dt,
                                                    C.byref(ai), C.byref(aj), int(args.lite_norm), C.byref(st)))
                args.overlap_stats[(layer, i, j)] = st
                if st.Nij == 0:
                    continue
                if args.lite_norm:  # :796-806
                    Mij[layer, i, j], Mij[layer, j, i], Sij[layer, i, j], Sij[layer, j, i] = st.medij, st.medji, st.madij, st.madji
                else:
                    Mij[layer, i, j], Mij[layer, j, i], Sij[layer, i, j], Sij[layer, j, i] = st.locij, st.locji, st.scaij, st.scaji
                Nij[layer, i, j] = Nij[layer, j, i] = st.Nij
    off, mul, sc = np.zeros((c, n)), np.ones((c, n)), np.ones((c, n))    # init_coeffs
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    for layer in range(c):
        m = np.ascontiguousarray(Mij[layer])
        rv = L.sgpu_overlap_solve(args.normalize, n, args.ref_image, p(np.ascontiguousarray(Nij[layer])), p(m),
                                  p(np.ascontiguousarray(Sij[layer])), p(off[layer]), p(mul[layer]), p(sc[layer]))
        if rv:
            raise _lib.SgpuError(rv, "overlap normalization: the system is singular (a frame overlaps no other)")
    args.coeff = NormCoeff(offset=off, mul=mul, scale=sc)
    return ST_OK


def do_normalization(args: StackingArgs) -> int:
    """do_normalization (normalization.c:46-78) -> compute_normalization (:217-294), or, with args->overlap_norm,
    compute_normalization_overlaps (:64-67, :666-938)."""
    if args.normalize == NO_NORM:
        return ST_OK
    try:
        if args.overlap_norm:
            return compute_normalization_overlaps(args)
        off, mul, sc, bg = compute_estimators(args)
        args.bgnoise = bg
        finalize_factors(args, off, mul, sc)
    except _lib.SgpuError as e:
        args.retval = e.code
        return e.code
    return ST_OK


def main_stack(args: StackingArgs) -> int:
    """main_stack (stacking.c:76-108): 1. normalization 2. (upscale: not on this path) 3. stack"""
    if args.coeff.scale is None or args.force_norm:
        args.retval = do_normalization(args)
        if args.retval:
            return args.retval
    args.retval = args.method(args)
    return args.retval
