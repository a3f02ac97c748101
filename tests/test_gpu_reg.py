"""GPU parity of the REGISTER engine (sgpu_stack_args.engine = 3: stack_reg.cuh / reg.cuh) against
the CPU oracle, through the C ABI: stack depths 32..512 (both lane groupings, every register-slot
variant), normalization, x/y shifts, ragged widths, rejection maps, null samples, ties and heavy
outliers (pixels the engine cannot decide identically are handed to the exact engine — the result
must be identical either way), and the share of pixels it decides itself on benchmark-like data."""
import numpy as np
import pytest

from oracle import oracle as O
from test_gpu_parity import NAMES, SIGS, compare, synth, to_dev, to_np

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def run_reg(fr, **kw):
    import siril_b200 as sb
    is_mean = kw.pop("is_mean", True)
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection if is_mean else sb.stack_median, frames=to_dev(fr),
                           type_of_rejection=kw.pop("rejection", O.NO_REJEC), sig=kw.pop("sig", (3, 3)),
                           normalize=kw.pop("normalize", O.NO_NORM),
                           coeff=sb.NormCoeff(offset=kw.pop("offset", None), mul=kw.pop("mul", None), scale=kw.pop("scale", None)),
                           shiftx=kw.pop("shiftx", None), shifty=kw.pop("shifty", None),
                           output_norm=kw.pop("output_norm", False),
                           create_rejmaps=kw.pop("create_rejmaps", False), merge_lowhigh_rejmaps=kw.pop("merge_rejmaps", False),
                           engine=3)
    assert not kw, kw
    def go():
        return args.method(args) if args.normalize != O.NO_NORM else sb.main_stack(args)
    rv = go()
    if rv != 0 and "does not cover" in str(getattr(args, "error", "")):
        args.engine = 0          # x shifts / unaligned planes: the automatic choice takes over
        rv = go()
    assert rv == 0, getattr(args, "error", rv)
    r = args.result
    return dict(out=to_np(r.data), rej_low=to_np(r.rejmap_low), rej_high=to_np(r.rejmap_high), irej=r.irej), args.ctx.handovers


REG_ALGOS = [O.NO_REJEC, O.PERCENTILE, O.SIGMA, O.WINSORIZED, None]


@pytest.mark.parametrize("algo", REG_ALGOS, ids=lambda a: "median" if a is None else NAMES[a])
@pytest.mark.parametrize("N", [32, 33, 64, 100, 101, 152, 153, 200, 256, 257, 300, 304, 305, 400, 401, 512])
def test_reg_depths(orc, algo, N):
    rng = np.random.default_rng(7 * N + (algo or 11))
    fr = synth(rng, N, 1, 5, 76, np.float32, zeros=0.0005, hot=0.002)   # 380 px: a plane TMA can address
    kw = dict(is_mean=False) if algo is None else dict(rejection=algo, sig=SIGS[algo], create_rejmaps=True)
    g, _ = run_reg(fr, **dict(kw))
    compare(g, orc.stack(fr, **kw), f"reg N={N}")


@pytest.mark.parametrize("algo", [O.SIGMA, O.WINSORIZED, O.PERCENTILE, None], ids=lambda a: "median" if a is None else NAMES[a])
def test_reg_norm_shifts_layers(orc, algo):
    rng = np.random.default_rng(90 + (algo or 0))
    N, Cn, H, W = 100, 3, 9, 68
    fr = synth(rng, N, Cn, H, W, np.float32, zeros=0.0)
    base = dict(is_mean=False) if algo is None else dict(rejection=algo, sig=SIGS[algo])
    coeffs = dict(scale=1 + 0.05 * rng.standard_normal((Cn, N)), offset=0.01 * rng.standard_normal((Cn, N)),
                  mul=1 + 0.05 * rng.standard_normal((Cn, N)))
    cases = [dict(normalize=O.ADDITIVE_SCALING, **coeffs), dict(normalize=O.MULTIPLICATIVE_SCALING, **coeffs),
             dict(normalize=O.ADDITIVE, shiftx=rng.integers(-5, 6, N), shifty=rng.integers(-3, 4, N), **coeffs),
             dict(normalize=O.MULTIPLICATIVE, output_norm=True, **coeffs),
             dict(shiftx=rng.integers(-5, 6, N))]
    for cs in cases:
        kw = dict(base, **cs)
        if algo is not None:
            kw.update(create_rejmaps=True, merge_rejmaps=bool(cs.get("output_norm")))
        g, _ = run_reg(fr, **dict(kw))
        compare(g, orc.stack(fr, **kw), str({k: v for k, v in cs.items() if k in ("normalize", "output_norm")}))


@pytest.mark.parametrize("algo", [O.SIGMA, O.WINSORIZED], ids=["sigma", "winsorized"])
def test_reg_adversarial(orc, algo):
    """null samples, ties, heavy outliers, a constant region: whatever the engine hands over must
    come back identical from the exact engine"""
    rng = np.random.default_rng(3 + algo)
    for N in (40, 100, 300):
        fr = synth(rng, N, 1, 12, 64, np.float32, zeros=0.01, hot=0.05)
        fr[:, :, :3, :] = np.round(fr[:, :, :3, :] * 200) / 200          # ties
        fr[:, :, 3, :] = 0.25                                             # all samples equal
        fr[:, :, 4, :32] = 0.0                                            # all null
        fr[: N // 3, :, 5, :] += 0.3                                      # a third of the stack displaced
        kw = dict(rejection=algo, sig=(2.0, 2.5), create_rejmaps=True)
        g, _ = run_reg(fr, **dict(kw))
        compare(g, orc.stack(fr, **kw), f"adversarial N={N}")


@pytest.mark.parametrize("case", [(100, O.WINSORIZED, O.ADDITIVE_SCALING), (300, O.SIGMA, O.NO_NORM), (200, None, O.ADDITIVE_SCALING)],
                         ids=["C2-like", "target-like", "C3-like"])
def test_reg_benchmark_like(orc, case):
    """benchmark-like frames (the generator bench.py uses): parity on a 64-row strip and nearly all
    pixels decided in registers"""
    import ctypes as C
    import siril_b200 as sb
    from siril_b200 import _lib
    N, algo, nrm = case
    H, W = 64, 1024
    ctx = sb.Context(0, torch.cuda.current_stream().cuda_stream)
    fr = torch.empty((N, 1, H, W), dtype=torch.float32, device="cuda")
    L = _lib.lib()
    desc = _lib.BlockDesc(N, 1, W, H, 1)
    blk = C.c_void_p()
    ctx.check(L.sgpu_block_wrap(ctx.h, C.byref(desc), C.c_void_p(fr.data_ptr()), C.byref(blk)))
    ctx.check(L.sgpu_synth_fill(ctx.h, blk, 1002, 0, H))
    ctx.synchronize()
    L.sgpu_block_destroy(blk)
    host = fr.cpu().numpy()
    rng = np.random.default_rng(5)
    kw = dict(is_mean=False) if algo is None else dict(rejection=algo, sig=(3, 3), create_rejmaps=True)
    if nrm != O.NO_NORM:
        kw.update(normalize=nrm, scale=1 + 0.02 * rng.standard_normal((1, N)), offset=0.002 * rng.standard_normal((1, N)),
                  mul=np.ones((1, N)))
    g, handed = run_reg(host, **dict(kw))
    compare(g, orc.stack(host, **kw), str(case))
    assert handed <= 0.005 * H * W, (handed, H * W)
