"""Host mirror of the block reader for FITS sequences: stack_read_block_data
(median_and_mean.c:382-545) over uncompressed primary-HDU files.  The host only moves file bytes
(pread -> pinned staging -> H2D inside libsgpu.so); big-endian decode, BZERO, the DATAMAX rescale,
y shift and zero fill run on the device (csrc/fitsread.cu).  No CPU decode path exists here."""
import ctypes as C
import os

import numpy as np

from . import _lib
from .enums import DATA_FLOAT, DATA_USHORT, ST_SEQUENCE_ERROR


def fits_probe(path):
    """sgpu_fits_probe: BITPIX / NAXISn / BZERO / BSCALE / DATAMAX and the data-unit offset."""
    info = _lib.FitsInfo()
    rv = _lib.lib().sgpu_fits_probe(os.fsencode(path), C.byref(info))
    if rv:
        raise _lib.SgpuError(rv, f"{path}: not a readable FITS image")
    return info


def read_block_data(ctx, paths, row0=0, height=None, shifty=None, source="fd"):
    """data->pix[] of one row block: a CUDA tensor [frame][layer][height][width] holding rows
    [row0, row0 + height) of every file in `paths` (frame f shifted by shifty[f], zeros outside).
    source: "fd" (pread into the library's pinned staging), "mmap" (pageable mapping of the file) or
    "pinned" (the data unit read into sgpu_host_alloc memory first)."""
    import torch
    L = _lib.lib()
    infos = [fits_probe(p) for p in paths]
    i0 = infos[0]
    w, fh, c = i0.naxes[0], i0.naxes[1], i0.naxes[2]
    for p, i in zip(paths, infos):
        if (i.naxes[0], i.naxes[1], i.naxes[2], i.bitpix) != (w, fh, c, i0.bitpix):
            raise _lib.SgpuError(ST_SEQUENCE_ERROR, f"{p}: geometry differs from the first image")
        if not i.device_readable:
            raise _lib.SgpuError(ST_SEQUENCE_ERROR, f"{p}: BITPIX {i.bitpix} BZERO {i.bzero} BSCALE {i.bscale} needs cfitsio")
    height = fh - row0 if height is None else height
    dt = DATA_FLOAT if i0.bitpix == -32 else DATA_USHORT
    n = len(paths)
    dev = torch.device("cuda", ctx.device)
    out = torch.empty((n, c, height, w), dtype=torch.float32 if dt == DATA_FLOAT else torch.uint16, device=dev)
    desc = _lib.BlockDesc(n, c, w, height, dt)
    blk = C.c_void_p()
    ctx.check(L.sgpu_block_wrap(ctx.h, C.byref(desc), C.c_void_p(out.data_ptr()), C.byref(blk)))
    try:
        for f, (p, i) in enumerate(zip(paths, infos)):
            sy = int(shifty[f]) if shifty is not None else 0
            resc = -1 if not i.has_datamax else int(i.datamax > 10.0)   # image_format_fits.c:995-1011
            nbytes = w * fh * c * abs(i.bitpix) // 8
            if source == "fd":
                fd = os.open(p, os.O_RDONLY)
                try:
                    ctx.check(L.sgpu_block_upload_fits_fd(blk, f, fd, i.data_offset, i.bitpix, fh, row0, sy, resc))
                finally:
                    os.close(fd)
            elif source == "mmap":
                m = np.memmap(p, dtype=np.uint8, mode="r", offset=i.data_offset, shape=(nbytes,))
                ctx.check(L.sgpu_block_upload_fits(blk, f, C.c_void_p(m.ctypes.data), i.bitpix, fh, row0, sy, resc))
                ctx.synchronize()
                del m
            else:
                hp = C.c_void_p()
                ctx.check(L.sgpu_host_alloc(ctx.h, nbytes, C.byref(hp)))
                try:
                    buf = (C.c_ubyte * nbytes).from_address(hp.value)
                    with open(p, "rb") as fh_:
                        fh_.seek(i.data_offset)
                        fh_.readinto(buf)
                    ctx.check(L.sgpu_block_upload_fits(blk, f, hp, i.bitpix, fh, row0, sy, resc))
                    ctx.synchronize()
                finally:
                    L.sgpu_host_free(ctx.h, hp)
        ctx.synchronize()
    finally:
        L.sgpu_block_destroy(blk)
    return out


def stack_ser(args, path, frame_indices=None, is_mean=True):
    """stack_mean_or_median for frames of a SER file (sgpu_stack_ser); results as stack_fits"""
    from . import stacking
    L = _lib.lib()
    ctx = stacking._context(args)
    si = _lib.SerInfo()
    if L.sgpu_ser_probe(os.fsencode(path), C.byref(si)):
        args.retval = -2
        return -2
    idx = list(range(si.frame_count)) if frame_indices is None else [int(i) for i in frame_indices]
    n, c, h, w = len(idx), 1, si.height, si.width
    keep = []
    try:
        ca = stacking._fill_cargs(args, n, c, is_mean, keep)
    except _lib.SgpuError as e:
        args.retval = e.code
        return e.code
    out32 = bool(args.use_32bit_output)
    nmaps = (1 if args.merge_lowhigh_rejmaps else 2) if (is_mean and args.create_rejmaps) else 0
    out = np.empty((c, h, w), dtype=np.float32 if out32 else np.uint16)
    rl = np.zeros((c, h, w), dtype=np.uint16) if nmaps >= 1 else None
    rh = np.zeros((c, h, w), dtype=np.uint16) if nmaps == 2 else None
    irej = np.zeros(6, dtype=np.uint64)
    sy = np.ascontiguousarray(args.shifty, dtype=np.int32) if args.shifty is not None else None
    ia = np.ascontiguousarray(idx, dtype=np.int32)
    rv = L.sgpu_stack_ser(ctx.h, os.fsencode(path), n, ia.ctypes.data_as(C.c_void_p),
                          sy.ctypes.data_as(C.c_void_p) if sy is not None else None, C.byref(ca), out.ctypes.data_as(C.c_void_p),
                          rl.ctypes.data_as(C.c_void_p) if rl is not None else None,
                          rh.ctypes.data_as(C.c_void_p) if rh is not None else None, irej.ctypes.data_as(C.c_void_p))
    args.retval = rv
    if rv:
        args.error = (L.sgpu_last_error(ctx.h) or b"").decode()
        return rv
    res = stacking.StackResult()
    res.data, res.rejmap_low, res.rejmap_high, res.irej = out, rl, rh, irej.reshape(3, 2)
    args.result = res
    return rv


def fitseq_probe(path):
    """the frames of a FITS sequence file (sgpu_fitseq_probe; _find_hdus, io/fits_sequence.c:39-150) as FitsInfo records"""
    L = _lib.lib()
    n = C.c_int(0)
    if L.sgpu_fitseq_probe(os.fsencode(path), None, 0, C.byref(n)):
        raise _lib.SgpuError(-2, f"{path}: not a FITS sequence file")
    infos = (_lib.FitsInfo * max(n.value, 1))()
    if L.sgpu_fitseq_probe(os.fsencode(path), infos, n.value, C.byref(n)):
        raise _lib.SgpuError(-2, f"{path}: not a FITS sequence file")
    return [infos[i] for i in range(n.value)]


def stack_fits(args, paths, is_mean=True, frame_indices=None):
    """stack_mean_or_median (median_and_mean.c:1261-1826) for a sequence of FITS files on disk:
    sgpu_stack_fits.  `args` is a StackingArgs without frames; the result lands in args.result as
    numpy arrays, like the host-frame path of stack_mean_or_median."""
    from . import stacking
    L = _lib.lib()
    ctx = stacking._context(args)
    fitseq = isinstance(paths, (str, bytes, os.PathLike))     # one FITS sequence file instead of a list of files
    if fitseq:
        frames = fitseq_probe(paths)
        idx = list(range(len(frames))) if frame_indices is None else [int(i) for i in frame_indices]
        if not frames or any(i < 0 or i >= len(frames) for i in idx):
            args.retval = -2
            return -2
        i0, n = frames[idx[0]], len(idx)
    else:
        i0, n = fits_probe(paths[0]), len(paths)
    w, h, c = i0.naxes[0], i0.naxes[1], i0.naxes[2]
    keep = []
    try:
        ca = stacking._fill_cargs(args, n, c, is_mean, keep)
    except _lib.SgpuError as e:
        args.retval = e.code
        return e.code
    out32 = bool(args.use_32bit_output)
    nmaps = (1 if args.merge_lowhigh_rejmaps else 2) if (is_mean and args.create_rejmaps) else 0
    out = np.empty((c, h, w), dtype=np.float32 if out32 else np.uint16)
    rl = np.zeros((c, h, w), dtype=np.uint16) if nmaps >= 1 else None
    rh = np.zeros((c, h, w), dtype=np.uint16) if nmaps == 2 else None
    irej = np.zeros(6, dtype=np.uint64)
    sy = np.ascontiguousarray(args.shifty, dtype=np.int32) if args.shifty is not None else None
    outs = (out.ctypes.data_as(C.c_void_p), rl.ctypes.data_as(C.c_void_p) if rl is not None else None,
            rh.ctypes.data_as(C.c_void_p) if rh is not None else None, irej.ctypes.data_as(C.c_void_p))
    syp = sy.ctypes.data_as(C.c_void_p) if sy is not None else None
    if fitseq:
        ia = np.ascontiguousarray(idx, dtype=np.int32)
        rv = L.sgpu_stack_fitseq(ctx.h, os.fsencode(paths), n, ia.ctypes.data_as(C.c_void_p), syp, C.byref(ca), *outs)
    else:
        cp = (C.c_char_p * n)(*[os.fsencode(p) for p in paths])
        rv = L.sgpu_stack_fits(ctx.h, n, cp, syp, C.byref(ca), *outs)
    args.retval = rv
    if rv:
        args.error = (L.sgpu_last_error(ctx.h) or b"").decode()
        return rv
    res = stacking.StackResult()
    res.data, res.rejmap_low, res.rejmap_high, res.irej = out, rl, rh, irej.reshape(3, 2)
    args.result = res
    return rv
