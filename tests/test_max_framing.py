"""Maximized framing / variable-size frames (median_and_mean.c:160-190 canvas, :363-380 + :394-480 the
read of a frame of its own size into the canvas-wide block)."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle as O


def test_canvas_matches_the_reference(ref):
    from siril_b200 import _lib
    L, R = _lib.lib(), ref.lib
    R.ref_maximized_framing.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    rng = np.random.default_rng(2)
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    for k in range(300):
        n = int(rng.integers(1, 12))
        rx, ry = rng.integers(50, 4000, n).astype(np.int32), rng.integers(50, 3000, n).astype(np.int32)
        h02, h12 = rng.uniform(-900, 900, n), rng.uniform(-900, 900, n)
        if k % 5 == 0:
            h02, h12 = np.round(h02), np.round(h12) + 0.5
        up = int(k % 3 == 0)
        a0, a1, off = C.c_long(), C.c_long(), (C.c_int * 2)()
        assert L.sgpu_maximized_framing(n, p(rx), p(ry), p(h02), p(h12), up, C.byref(a0), C.byref(a1), off) == 0
        en, eo = np.zeros(2, dtype=np.int64), np.zeros(2, dtype=np.int32)
        assert R.ref_maximized_framing(n, p(rx), p(ry), p(h02), p(h12), up, p(en), p(eo)) == 0
        assert [a0.value, a1.value, off[0], off[1]] == [en[0], en[1], eo[0], eo[1]]


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
def test_variable_size_frames_on_a_maximized_canvas(ref, dtype):
    """frames of their own sizes, registered by translations, stacked on the canvas that holds them all: the GPU block is
    filled by sgpu_block_upload_frame_sized from the bare frames; the oracle gets the same frames padded by hand to the
    canvas (top-left aligned, which is where the reference's read puts them) and the same integer shifts"""
    torch = pytest.importorskip("torch")
    import siril_b200 as sb
    from siril_b200 import _lib
    from test_gpu_parity import compare, to_np
    L = _lib.lib()
    ctx = _lib.Context(0)
    rng = np.random.default_rng(21)
    N, Cn = 7, 2
    rx = rng.integers(60, 100, N).astype(np.int32); ry = rng.integers(40, 70, N).astype(np.int32)
    h02 = rng.uniform(-20, 20, N); h12 = rng.uniform(-15, 15, N)
    a0, a1, off = C.c_long(), C.c_long(), (C.c_int * 2)()
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    assert L.sgpu_maximized_framing(N, p(rx), p(ry), p(h02), p(h12), 0, C.byref(a0), C.byref(a1), off) == 0
    Wc, Hc = a0.value, a1.value
    # integer shifts as the block loop computes them (median_and_mean.c:416-418, :1619-1623): dx = h02 - offset[0], dy = -h12 - offset[1]
    rnd = lambda v: int(np.floor(v + 0.5)) if v >= 0 else -int(np.floor(-v + 0.5))
    sx = np.array([rnd(h02[f] - off[0]) for f in range(N)], dtype=np.int32)
    sy = np.array([rnd(-h12[f] - off[1]) for f in range(N)], dtype=np.int32)
    frames, padded = [], np.zeros((N, Cn, Hc, Wc), dtype=dtype)
    for f in range(N):
        fr = (0.2 + 0.02 * rng.standard_normal((Cn, ry[f], rx[f])))
        fr[rng.random(fr.shape) < 0.02] += 0.3
        fr = np.clip(fr, 1e-3, 1)
        fr = np.round(fr * 65535).astype(np.uint16) if dtype == np.uint16 else fr.astype(np.float32)
        frames.append(np.ascontiguousarray(fr))
        padded[f, :, Hc - ry[f]:, : rx[f]] = fr          # memory rows are bottom-up: the frame's top row on the canvas' top row
    desc = _lib.BlockDesc(N, Cn, Wc, Hc, 1 if dtype == np.float32 else 0)
    blk = C.c_void_p()
    ctx.check(L.sgpu_block_create(ctx.h, C.byref(desc), C.byref(blk)))
    try:
        for f in range(N):
            ctx.check(L.sgpu_block_upload_frame_sized(blk, f, p(frames[f]), int(rx[f]), int(ry[f]), Hc, 0, int(sy[f])))
        args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=None, type_of_rejection=O.SIGMA, sig=(3, 3),
                               shiftx=sx, create_rejmaps=True, ctx=ctx, use_32bit_output=True)
        keep = []
        ca = sb.stacking._fill_cargs(args, N, Cn, True, keep)
        out = torch.empty((Cn, Hc, Wc), dtype=torch.float32, device="cuda")
        lo = torch.zeros((Cn, Hc, Wc), dtype=torch.int16, device="cuda"); hi = torch.zeros_like(lo)
        irej = torch.zeros(6, dtype=torch.int64, device="cuda")
        ctx.check(L.sgpu_stack_block(ctx.h, blk, C.byref(ca), out.data_ptr(), lo.data_ptr(), hi.data_ptr(), irej.data_ptr()))
        ctx.synchronize()
    finally:
        L.sgpu_block_destroy(blk)
    g = dict(out=out.cpu().numpy(), rej_low=lo.cpu().numpy().view(np.uint16), rej_high=hi.cpu().numpy().view(np.uint16),
             irej=irej.cpu().numpy().astype(np.uint64).reshape(3, 2))
    o = ref.stack(padded, rejection=O.SIGMA, sig=(3, 3), shiftx=sx, shifty=sy, create_rejmaps=True)
    compare(g, o, "maximized canvas")
    assert (o["out"] == 0).any() and (o["out"] > 0).any()       # the canvas has corners no frame reaches
