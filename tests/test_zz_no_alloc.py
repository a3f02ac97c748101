"""sgpu_minmax_nonzero_static — the non-allocating variant of sgpu_minmax_nonzero (scratch in a
file-scope static device array, include/sgpu.h) — against the existing entry point, through the C
ABI, on calls in a row.  Kept in a file pytest collects LAST."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def both(ctx, t, skip_first):
    a, b = (C.c_float * 2)(), (C.c_float * 2)()
    ctx.check(ctx.L.sgpu_minmax_nonzero(ctx.h, C.c_void_p(t.data_ptr()), t.numel(), skip_first, a))
    ctx.check(ctx.L.sgpu_minmax_nonzero_static(ctx.h, C.c_void_p(t.data_ptr()), t.numel(), skip_first, b))
    return (a[0], a[1]), (b[0], b[1])


def test_static_scratch_variant_matches_the_allocating_entry_point():
    import siril_b200 as sb
    ctx = sb.Context(0)
    rng = np.random.default_rng(99)
    # calls in a row on ONE context, sizes going down as well as up: a call must never read a partial
    # an earlier, larger call left in the shared array
    for n in (3_000_001, 70_000, 65_537, 65_536, 1, 2, 1_000_003, 255):
        img = rng.standard_normal(n).astype(np.float32)
        img[rng.random(n) < 0.01] = 0.0
        img[0] = 1e9                                   # only skip_first keeps it out
        if n > 2:
            img[rng.integers(1, n)] = 0.0
        t = torch.from_numpy(img).cuda()
        for skip in (0, 1):
            exp, got = both(ctx, t, skip)
            assert got == exp, (n, skip, got, exp)   # bit-exact, as the parity tests of this path
            nz = img[skip:][img[skip:] != 0]
            if len(nz):
                assert got == (float(nz.min()), float(nz.max())), (n, skip)
    # all zeros after a populated call: the empty result, not the last call's
    z = torch.zeros(100_000, device="cuda")
    exp, got = both(ctx, z, 0)
    assert got == exp == (float(np.finfo(np.float32).max), -float(np.finfo(np.float32).max))
    ctx.close()


def test_static_scratch_limit_is_enforced():
    import siril_b200 as sb
    ctx = sb.Context(0)
    t = torch.ones(16, device="cuda")
    out = (C.c_float * 2)()
    limit = (65536 // 8 - 1) * 65536                    # SGPU_STATIC_MAX_PIXELS
    rv = ctx.L.sgpu_minmax_nonzero_static(ctx.h, C.c_void_p(t.data_ptr()), limit + 1, 0, out)   # refused before any access
    assert rv == -10                                    # SGPU_ALLOC_ERROR
    msg = (ctx.L.sgpu_last_error(ctx.h) or b"").decode()
    assert "beyond the static scratch's limit" in msg and str(limit) in msg, msg
    # and the context still works
    rv = ctx.L.sgpu_minmax_nonzero_static(ctx.h, C.c_void_p(t.data_ptr()), 16, 0, out)
    assert rv == 0 and (out[0], out[1]) == (1.0, 1.0)
    ctx.close()
