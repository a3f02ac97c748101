"""GPU parity tests proper: the CUDA path (through the C ABI) against the CPU oracle on the same
seeded inputs.  Bar (BASELINE.json north_star): rejection counts / maps and integer outputs
bit-exact; float stacked pixels within 1e-6 relative."""
import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

ALGOS = [O.NO_REJEC, O.PERCENTILE, O.SIGMA, O.MAD, O.SIGMEDIAN, O.WINSORIZED, O.LINEARFIT, O.GESDT]
NAMES = ["none", "percentile", "sigma", "mad", "sigmedian", "winsorized", "linearfit", "gesdt"]
SIGS = {O.NO_REJEC: (3, 3), O.PERCENTILE: (0.2, 0.3), O.SIGMA: (3, 3), O.MAD: (3, 3), O.SIGMEDIAN: (3, 3),
        O.WINSORIZED: (3, 3), O.LINEARFIT: (5, 5), O.GESDT: (0.3, 0.05)}
REL_TOL = 1e-6   # float stacked pixels (north_star)
ENGINE = 0       # sgpu_stack_args.engine: 0 automatic, 1 exact, 2 sorted, 3 register, 4 thread-sort, 5 lane-sort, 6 selection engine


@pytest.fixture(autouse=True, params=[0, 1, 2, 5, 6], ids=["auto", "exact", "sorted", "lsort", "psel"])
def engine(request):
    """every parity test runs through the automatic choice (register engine, else sorted engine,
    each with its exact-engine hand-over list), the exact thread-per-pixel engine alone, and the
    sorted and thread-sort engines wherever they apply (test_gpu_reg.py forces the register engine)"""
    global ENGINE
    ENGINE = request.param
    yield request.param
    ENGINE = 0


def synth(rng, N, Cn, H, W, dtype, zeros=0.002, hot=0.01):
    sky = 0.1 * (1 + 0.05 * rng.standard_normal((N, 1, 1, 1)))
    gain = 1 + 0.1 * rng.standard_normal((N, 1, 1, 1))
    yy, xx = np.mgrid[0:H, 0:W]
    sig = 0.2 * np.exp(-((yy - H / 2) ** 2 + (xx - W / 2) ** 2) / (0.1 * H * W))
    fr = gain * (sky + sig[None, None]) + 0.01 * rng.standard_normal((N, Cn, H, W))
    m = rng.random((N, Cn, H, W)) < hot
    fr[m] += rng.uniform(0.2, 0.6, m.sum())
    fr = np.clip(fr, 1e-4, 1.0)
    if zeros:
        fr[rng.random((N, Cn, H, W)) < zeros] = 0.0
    if dtype == np.uint16:
        return np.round(fr * 65535).astype(np.uint16)
    return fr.astype(np.float32)


def to_dev(fr):
    if fr.dtype == np.uint16:
        return torch.from_numpy(fr.view(np.int16)).cuda().view(torch.uint16)
    return torch.from_numpy(fr).cuda()


def to_np(t):
    if t is None:
        return None
    if t.dtype == torch.uint16:
        return t.view(torch.int16).cpu().numpy().view(np.uint16)
    return t.cpu().numpy()


def run_gpu(fr, device_resident=True, **kw):
    import siril_b200 as sb
    is_mean = kw.pop("is_mean", True)
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection if is_mean else sb.stack_median,
                           frames=to_dev(fr) if device_resident else fr,
                           type_of_rejection=kw.pop("rejection", O.NO_REJEC), sig=kw.pop("sig", (3, 3)),
                           critical_value=kw.pop("crit", None),
                           normalize=kw.pop("normalize", O.NO_NORM),
                           coeff=sb.NormCoeff(offset=kw.pop("offset", None), mul=kw.pop("mul", None), scale=kw.pop("scale", None)),
                           weights=kw.pop("weights", None), shiftx=kw.pop("shiftx", None), shifty=kw.pop("shifty", None),
                           output_norm=kw.pop("output_norm", False), use_32bit_output=kw.pop("use_32bit_output", True),
                           create_rejmaps=kw.pop("create_rejmaps", False), merge_lowhigh_rejmaps=kw.pop("merge_rejmaps", False),
                           engine=ENGINE)
    if args.weights is not None:
        args.weighting_type = sb.NOISE_WEIGHT
    assert not kw, kw
    def go():
        if args.normalize != O.NO_NORM:
            return args.method(args)       # coefficients supplied: step 3 of main_stack only
        return sb.main_stack(args)
    rv = go()
    if rv != 0 and ENGINE in (2, 5, 6) and "does not cover" in str(getattr(args, "error", "")):
        args.engine = 0                    # configuration outside the forced engine: automatic choice
        rv = go()
    assert rv == 0, getattr(args, "error", rv)
    r = args.result
    return dict(out=to_np(r.data) if device_resident else r.data,
                rej_low=to_np(r.rejmap_low) if device_resident else r.rejmap_low,
                rej_high=to_np(r.rejmap_high) if device_resident else r.rejmap_high, irej=r.irej)


def compare(g, o, what=""):
    assert np.array_equal(g["irej"], o["irej"]), (what, g["irej"], o["irej"])
    for k in ("rej_low", "rej_high"):
        assert (g[k] is None) == (o[k] is None), (what, k)
        if g[k] is not None:
            bad = np.argwhere(g[k] != o[k])
            assert len(bad) == 0, (what, k, bad[:5], g[k][tuple(bad[0])], o[k][tuple(bad[0])])
    a, b = g["out"], o["out"]
    assert a.dtype == b.dtype and a.shape == b.shape
    if a.dtype == np.uint16:
        assert np.array_equal(a, b), what
    else:
        both_nan = np.isnan(a) & np.isnan(b)
        rel = np.abs(a - b) / np.maximum(np.abs(b), 1e-30)
        rel[both_nan] = 0
        rel[(a == b)] = 0
        assert np.nanmax(rel) <= REL_TOL, (what, float(np.nanmax(rel)), np.unravel_index(np.nanargmax(rel), rel.shape))


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("algo", ALGOS, ids=NAMES)
@pytest.mark.parametrize("N", [3, 8, 9, 20, 50])
def test_rejection_parity(orc, algo, dtype, N):
    rng = np.random.default_rng(1000 * N + algo)
    fr = synth(rng, N, 1, 24, 40, dtype)
    kw = dict(rejection=algo, sig=SIGS[algo], create_rejmaps=True)
    compare(run_gpu(fr, **kw), orc.stack(fr, **kw), f"{NAMES[algo]} N={N}")


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("algo", ALGOS, ids=NAMES)
def test_norm_weights_shifts(orc, algo, dtype):
    rng = np.random.default_rng(50 + algo)
    N, Cn, H, W = 18, 3, 12, 36
    fr = synth(rng, N, Cn, H, W, dtype)
    k = 65535 if dtype == np.uint16 else 1
    cases = [dict(normalize=O.ADDITIVE_SCALING), dict(normalize=O.MULTIPLICATIVE_SCALING, weighted=True),
             dict(normalize=O.ADDITIVE, shifts=True, create_rejmaps=True, weighted=True),
             dict(normalize=O.MULTIPLICATIVE, create_rejmaps=True, merge_rejmaps=True, output_norm=True),
             dict(shifts=True)]
    for cs in cases:
        kw = dict(rejection=algo, sig=SIGS[algo])
        if "normalize" in cs:
            kw.update(normalize=cs["normalize"], scale=1 + 0.05 * rng.standard_normal((Cn, N)),
                      offset=0.01 * rng.standard_normal((Cn, N)) * k, mul=1 + 0.05 * rng.standard_normal((Cn, N)))
        if cs.get("weighted"):
            kw["weights"] = rng.uniform(0.5, 1.5, (Cn, N))
        if cs.get("shifts"):
            kw.update(shiftx=rng.integers(-5, 6, N), shifty=rng.integers(-3, 4, N))
        for key in ("create_rejmaps", "merge_rejmaps", "output_norm"):
            if cs.get(key):
                kw[key] = True
        outs = [True, False] if (dtype == np.uint16 and not cs.get("output_norm")) else [True]
        for out32 in outs:
            compare(run_gpu(fr, use_32bit_output=out32, **dict(kw)), orc.stack(fr, use_32bit_output=out32, **kw), str(cs))


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("N", [2, 7, 8, 9, 20, 33, 64])
def test_median_parity(orc, dtype, N):
    rng = np.random.default_rng(N)
    fr = synth(rng, N, 1, 16, 50, dtype)
    k = 65535 if dtype == np.uint16 else 1
    kw = dict(is_mean=False, normalize=O.ADDITIVE_SCALING, scale=1 + 0.05 * rng.standard_normal((1, N)),
              offset=0.01 * rng.standard_normal((1, N)) * k, mul=np.ones((1, N)))
    g, o = run_gpu(fr, **dict(kw)), orc.stack(fr, **kw)
    compare(g, o, f"median N={N}")
    if dtype == np.uint16:   # 16-bit median must be bit-exact
        g, o = run_gpu(fr, use_32bit_output=False, **dict(kw)), orc.stack(fr, use_32bit_output=False, **kw)
        assert np.array_equal(g["out"], o["out"])


@pytest.mark.parametrize("algo", [O.SIGMA, O.WINSORIZED, O.LINEARFIT, O.GESDT, O.MAD], ids=["sigma", "winsorized", "linearfit", "gesdt", "mad"])
def test_adversarial_stacks(orc, algo):
    """many outliers, tiny N, zeros, ties: the `N - r <= 4` guard regime where results depend on
    the order quickselect leaves the stack in"""
    rng = np.random.default_rng(31 + algo)
    for N in (5, 6, 7, 10, 12, 30):
        if algo == O.GESDT and N < 4:
            continue
        for dtype in (np.float32, np.uint16):
            fr = synth(rng, N, 1, 16, 32, dtype, zeros=0.1, hot=0.3)
            fr[:, :, :4, :] = np.round(fr[:, :, :4, :].astype(np.float64) * (1 if dtype == np.uint16 else 50)) / (1 if dtype == np.uint16 else 50)
            fr = fr.astype(dtype)
            kw = dict(rejection=algo, sig=(1.0, 1.0) if algo != O.GESDT else (0.3, 0.05), create_rejmaps=True)
            compare(run_gpu(fr, **kw), orc.stack(fr, **kw), f"adversarial N={N} {dtype.__name__}")


def test_host_path_and_ragged(orc):
    """end-to-end host entry point (H2D + stack + D2H), widths that are not multiples of the tile"""
    rng = np.random.default_rng(8)
    for (N, H, W) in [(20, 5, 1), (12, 3, 33), (9, 17, 257)]:
        fr = synth(rng, N, 1, H, W, np.float32)
        kw = dict(rejection=O.SIGMA, sig=(3, 3), create_rejmaps=True, shifty=rng.integers(-2, 3, N))
        compare(run_gpu(fr, device_resident=False, **dict(kw)), orc.stack(fr, **kw), f"host {N}x{H}x{W}")


@pytest.mark.parametrize("nblk", [1, 2, 3, 7])
@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
def test_host_path_row_blocks(orc, monkeypatch, nblk, dtype):
    """sgpu_stack_host cut into row blocks (upload of block k + 1 overlapping the stack of block k):
    the same result for every block count, three layers, y shifts, both rejection maps"""
    monkeypatch.setenv("SGPU_HOST_BLOCKS", str(nblk))
    rng = np.random.default_rng(80 + nblk)
    N, Cn, H, W = 14, 3, 23, 70
    fr = synth(rng, N, Cn, H, W, dtype)
    kw = dict(rejection=O.WINSORIZED, sig=(3, 3), create_rejmaps=True, shifty=rng.integers(-3, 4, N),
              shiftx=rng.integers(-2, 3, N))
    compare(run_gpu(fr, device_resident=False, **dict(kw)), orc.stack(fr, **kw), f"host blocks {nblk}")
    kw = dict(is_mean=False)
    compare(run_gpu(fr, device_resident=False, **dict(kw)), orc.stack(fr, **kw), f"host blocks {nblk} median")


def test_all_zero_and_single_valid(orc):
    fr = np.zeros((10, 1, 4, 8), dtype=np.float32)
    fr[3, 0, 1, :] = 0.25
    fr[:, 0, 2, :] = 0.5
    for algo in ALGOS:
        kw = dict(rejection=algo, sig=SIGS[algo])
        compare(run_gpu(fr, **kw), orc.stack(fr, **kw), f"zeros {NAMES[algo]}")


def test_error_codes():
    import siril_b200 as sb
    fr = torch.zeros((1, 1, 4, 4), device="cuda")
    a = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=fr)
    assert sb.main_stack(a) == sb.ST_GENERIC_ERROR          # "Select at least two frames"
    fr = torch.zeros((2, 1, 4, 4), device="cuda")
    a = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=fr, type_of_rejection=sb.GESDT, sig=(0.3, 0.05))
    assert sb.main_stack(a) == sb.ST_GENERIC_ERROR          # GESDT needs three frames


def test_config_c1_shape(orc):
    """BASELINE config C1: 20 x 1024x1024 mono f32, sigma clip 3/3, no normalization"""
    import siril_b200 as sb
    ctx = sb.Context(0, torch.cuda.current_stream().cuda_stream)
    fr = torch.empty((20, 1, 1024, 1024), dtype=torch.float32, device="cuda")
    import ctypes as C
    from siril_b200 import _lib
    L = _lib.lib()
    desc = _lib.BlockDesc(20, 1, 1024, 1024, 1)
    blk = C.c_void_p()
    ctx.check(L.sgpu_block_wrap(ctx.h, C.byref(desc), C.c_void_p(fr.data_ptr()), C.byref(blk)))
    ctx.check(L.sgpu_synth_fill(ctx.h, blk, 1001, 0, 1024))
    ctx.synchronize()
    L.sgpu_block_destroy(blk)
    host = fr.cpu().numpy()
    kw = dict(rejection=O.SIGMA, sig=(3, 3), create_rejmaps=True)
    g = run_gpu(host, **kw)
    o = orc.stack(host, **kw)
    compare(g, o, "C1")
    assert g["irej"][0].sum() > 0


@pytest.mark.parametrize("case", [
    (100, np.float32, O.WINSORIZED), (100, np.float32, O.SIGMA), (200, np.float32, None), (300, np.float32, O.GESDT),
    (300, np.float32, O.SIGMA), (500, np.uint16, O.LINEARFIT), (130, np.uint16, O.GESDT), (257, np.uint16, O.SIGMA),
    (65, np.float32, O.PERCENTILE), (600, np.float32, O.SIGMA), (1100, np.uint16, None),
    (90, np.float32, O.SIGMA), (180, np.float32, O.WINSORIZED), (350, np.float32, O.SIGMA), (700, np.uint16, O.SIGMA), (330, np.float32, None)],
    ids=lambda c: f"{c[0]}-{np.dtype(c[1]).name}-{'median' if c[2] is None else NAMES[c[2]]}")
def test_deep_stacks(orc, case):
    """stack depths of the benchmark configs (BASELINE.json C2..C5) on a small image: every
    register-tile variant of the sorted engine (32 .. 2048 slots)"""
    N, dtype, algo = case
    rng = np.random.default_rng(N)
    fr = synth(rng, N, 1, 6, 70, dtype, zeros=0.001, hot=0.002)
    k = 65535 if dtype == np.uint16 else 1
    kw = dict(normalize=O.ADDITIVE_SCALING, scale=1 + 0.02 * rng.standard_normal((1, N)),
              offset=0.002 * rng.standard_normal((1, N)) * k, mul=np.ones((1, N)))
    if algo is None:
        kw.update(is_mean=False)
    else:
        kw.update(rejection=algo, sig=SIGS[algo], create_rejmaps=True)
    compare(run_gpu(fr, **dict(kw)), orc.stack(fr, **kw), str(case))


@pytest.mark.parametrize("N,s0", [(20, 0.35), (10, 0.7), (40, 0.35), (10, 0.9)])
def test_gesdt_outlier_budget_is_a_float_product(orc, N, s0):
    """median_and_mean.c:1479: the critical-value table has (int) floor(nb_frames * sig[0]) entries with the
    product taken in FLOAT — one more than the double product for these depths; a shorter table is refused"""
    import ctypes as C
    import siril_b200 as sb
    rng = np.random.default_rng(N)
    fr = synth(rng, N, 1, 8, 40, np.float32, zeros=0.0, hot=0.05)
    kw = dict(rejection=O.GESDT, sig=(s0, 0.05), create_rejmaps=True)
    crit = O.gesdt_critical_values(N, s0, 0.05)
    assert len(crit) == int(np.floor(np.float32(N) * np.float32(s0)))
    compare(run_gpu(fr, **dict(kw)), orc.stack(fr, crit=crit, **kw), f"gesdt N={N} sig0={s0}")
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=to_dev(fr), type_of_rejection=O.GESDT, sig=(s0, 0.05),
                           critical_value=crit[:-1], engine=ENGINE)
    assert sb.main_stack(args) == sb.ST_GENERIC_ERROR and "critical values" in args.error


@pytest.mark.parametrize("case", [
    # N, dtype, sig0, zeros, what
    (450, np.float32, 0.3, 0.001, "budget past the 128 removals the lane groups take: one-thread code"),
    (300, np.float32, 0.3, 0.06, "many nulls: every pixel its own budget"),
    (64, np.uint16, 0.3, 0.03, "16-bit samples, nulls"),
    (12, np.float32, 0.3, 0.3, "budget spent on the nulls, stacks of one and two samples"),
    (33, np.float32, 0.47, 0.0, "flagged positions could collide: exact engine"),
    (200, np.float32, 0.3, 0.002, "frame weights"),
    (100, np.float32, 0.3, 0.002, "samples of both signs: the sums are not exact"),
    (128, np.float32, 0.3, 0.0, "a full 32-bit word of decisions and some"),
    (428, np.uint16, 0.3, 0.0, "128 removals exactly"),
], ids=lambda c: f"{c[0]}-{np.dtype(c[1]).name}-{c[2]}-{c[3]}")
def test_gesdt_lane_groups(orc, case):
    """GESDT in the lane-sort engine (stack_lsort.cuh: gesdt_decide / coop_gesdt): the decisions of one thread per pixel and the
    Grubbs tests of four lanes per pixel against rejection_float.c:301-348 + confirm_outliers, at the edges of what the
    groups take.  The image is not a multiple of the tile, so the last CTA has lanes without a pixel."""
    N, dtype, s0, zeros, what = case
    rng = np.random.default_rng(7 * N + 1)
    fr = synth(rng, N, 1, 5, 75, dtype, zeros=zeros, hot=0.03)
    kw = dict(rejection=O.GESDT, sig=(s0, 0.05), create_rejmaps=True)
    if "weights" in what:
        kw["weights"] = rng.uniform(0.5, 1.5, (1, N))
    if "both signs" in what:
        kw.update(normalize=O.ADDITIVE, scale=np.ones((1, N)), mul=np.ones((1, N)),
                  offset=0.12 + 0.002 * rng.standard_normal((1, N)))
    g = run_gpu(fr, **dict(kw))
    o = orc.stack(fr, **kw)
    compare(g, o, what)
    if s0 < 0.4 and zeros < 0.1:
        assert g["irej"][0].sum() > 0, what
