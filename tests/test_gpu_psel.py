"""GPU parity tests of the selection engine (engine 6, siril_b200/csrc/stack_psel.cuh): every
compiled variant (lanes per pixel x register slots), both sample types, the TMA and the row-load
tile paths, nulls, ties, weights, normalization — against the CPU oracle, and strips of the
benchmark workloads (sgpu_synth_fill data, full width, real depth) against the reference's own
code.  Bar: rejection counts / maps and integer outputs bit-exact, float pixels within 1e-6."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle as O
import test_gpu_parity as tp
from test_gpu_parity import compare, synth

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

PSEL = 6
# every (lanes, slots) variant of stack_psel.cu: the shallowest and the deepest stack it serves
DEPTHS = [41, 50, 51, 64, 65, 80, 81, 100, 101, 128, 129, 160, 161, 200, 201, 256, 257, 300, 301, 320,
          321, 400, 401, 512, 513, 600, 601, 800, 801, 1024, 1025, 1152]


@pytest.fixture(autouse=True)
def force_psel():
    tp.ENGINE = PSEL
    yield
    tp.ENGINE = 0


def run(fr, **kw):
    return tp.run_gpu(fr, **kw)


@pytest.mark.parametrize("N", DEPTHS)
def test_every_variant(orc, N):
    """sigma clip (f32, TMA tile), median (u16, row loads), percentile (f32, x shifts: row loads)"""
    rng = np.random.default_rng(7000 + N)
    fr = synth(rng, N, 1, 5, 72, np.float32, zeros=0.001, hot=0.002)
    kw = dict(rejection=O.SIGMA, sig=(3, 3), create_rejmaps=True)
    compare(run(fr, **dict(kw)), orc.stack(fr, **kw), f"sigma N={N}")
    fu = synth(rng, N, 1, 3, 40, np.uint16, zeros=0.001, hot=0.002)
    kw = dict(is_mean=False)
    compare(run(fu, **dict(kw)), orc.stack(fu, **kw), f"median u16 N={N}")
    kw = dict(rejection=O.PERCENTILE, sig=(0.2, 0.3), create_rejmaps=True, merge_rejmaps=True, shiftx=rng.integers(-3, 4, N))
    compare(run(fr, **dict(kw)), orc.stack(fr, **kw), f"percentile shifted N={N}")


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("algo", [O.NO_REJEC, O.PERCENTILE, O.SIGMA], ids=["none", "percentile", "sigma"])
@pytest.mark.parametrize("N", [64, 100, 200, 300])
def test_norm_weights_nulls(orc, algo, dtype, N):
    rng = np.random.default_rng(300 + N + algo)
    Cn, H, W = 3, 4, 68
    fr = synth(rng, N, Cn, H, W, dtype, zeros=0.05, hot=0.004)
    fr[: N // 2, :, 0, :8] = 0           # a border where half of the frames are null
    fr[:, :, 1, :4] = fr[0, :, 1, :4]    # constant stacks
    k = 65535 if dtype == np.uint16 else 1
    sig = tp.SIGS[algo]
    for cs in (dict(), dict(norm=O.ADDITIVE_SCALING), dict(norm=O.MULTIPLICATIVE, weighted=True), dict(weighted=True, norm=O.ADDITIVE)):
        kw = dict(rejection=algo, sig=sig, create_rejmaps=True)
        if "norm" in cs:
            kw.update(normalize=cs["norm"], scale=1 + 0.03 * rng.standard_normal((Cn, N)),
                      offset=0.005 * rng.standard_normal((Cn, N)) * k, mul=1 + 0.03 * rng.standard_normal((Cn, N)))
        if cs.get("weighted"):
            kw["weights"] = rng.uniform(0.5, 1.5, (Cn, N))
        compare(run(fr, **dict(kw)), orc.stack(fr, **kw), f"{cs} N={N}")


@pytest.mark.parametrize("N", [100, 300])
def test_ties_and_quantized(orc, N):
    """few distinct values: long runs of equal samples around the median"""
    rng = np.random.default_rng(N)
    fr = synth(rng, N, 1, 4, 64, np.float32, zeros=0.0, hot=0.002)
    fr = (np.round(fr * 400) / 400).astype(np.float32)
    fu = (np.round(fr * 2000)).astype(np.uint16)
    for f in (fr, fu):
        for kw in (dict(rejection=O.SIGMA, sig=(3, 3), create_rejmaps=True), dict(is_mean=False)):
            compare(run(f, **dict(kw)), orc.stack(f, **kw), f"ties N={N} {f.dtype}")


def synth_fill(n, c, h, w, dtype, seed, row0=0, full_h=None):
    import siril_b200 as sb
    from siril_b200 import _lib
    ctx = sb.Context(0, torch.cuda.current_stream().cuda_stream)
    L = _lib.lib()
    tdt = torch.float32 if dtype == np.float32 else torch.uint16
    fr = torch.empty((n, c, h, w), dtype=tdt, device="cuda")
    desc = _lib.BlockDesc(n, c, w, h, 1 if dtype == np.float32 else 0)
    blk = C.c_void_p()
    ctx.check(L.sgpu_block_wrap(ctx.h, C.byref(desc), C.c_void_p(fr.data_ptr()), C.byref(blk)))
    ctx.check(L.sgpu_synth_fill(ctx.h, blk, seed, row0, full_h or h))
    ctx.synchronize()
    L.sgpu_block_destroy(blk)
    return fr


@pytest.mark.parametrize("name,n,c,w,full_h,dtype,seed,kw", [
    ("target", 300, 1, 9576, 6388, np.float32, 1004, dict(rejection=O.SIGMA, sig=(3, 3), create_rejmaps=True)),
    ("C3", 200, 3, 6000, 4000, np.float32, 1003, dict(is_mean=False)),
    ("C2-sigma", 100, 1, 6248, 4176, np.float32, 1002, dict(rejection=O.SIGMA, sig=(3, 3), create_rejmaps=True)),
    ("C5-sigma", 500, 1, 9576, 6388, np.uint16, 1005, dict(rejection=O.SIGMA, sig=(3, 3), create_rejmaps=True)),
    ("C2-norm-sigma", 100, 1, 6248, 4176, np.float32, 1002, dict(rejection=O.SIGMA, sig=(3, 3), create_rejmaps=True, normalize=O.ADDITIVE_SCALING)),
], ids=["target", "C3", "C2-sigma", "C5-sigma", "C2-norm-sigma"])
def test_config_strips(ref, name, n, c, w, full_h, dtype, seed, kw):
    """the frames the benchmark stacks (sgpu_synth_fill), full width, real depth, 16 rows from the
    middle of the image, against the reference's own code"""
    import siril_b200 as sb
    rows = 16
    fr = synth_fill(n, c, rows, w, dtype, seed, row0=full_h // 2, full_h=full_h)
    host = tp.to_np(fr)
    kw = dict(kw)
    args = sb.StackingArgs(method=sb.stack_median if kw.get("is_mean") is False else sb.stack_mean_with_rejection, frames=fr,
                           type_of_rejection=kw.get("rejection", O.NO_REJEC), sig=kw.get("sig", (3, 3)),
                           create_rejmaps=kw.get("create_rejmaps", False), engine=PSEL)
    if kw.get("normalize"):     # the configuration's real coefficients: do_normalization on the strip's own statistics
        args.normalize = kw["normalize"]
        assert sb.do_normalization(args) == 0
        kw.update(scale=args.coeff.scale.copy(), offset=args.coeff.offset.copy(), mul=args.coeff.mul.copy())
        assert args.method(args) == 0, getattr(args, "error", "")
    else:
        assert sb.main_stack(args) == 0, getattr(args, "error", "")
    assert args.ctx.last_engine == "stack_psel_kernel"
    handed = args.ctx.handovers
    r = args.result
    g = dict(out=tp.to_np(r.data), rej_low=tp.to_np(r.rejmap_low), rej_high=tp.to_np(r.rejmap_high), irej=r.irej)
    compare(g, ref.stack(host, **kw), name)
    assert handed <= 1e-3 * c * rows * w + 2, handed     # the selection engine decides (nearly) every pixel itself
