"""GPU parity of the normalization inputs (K3/K4): per-image statistics, estimators, factors and
noise weights against the oracle.  Histogram-derived quantities (ngoodpix, min, max, median, MAD,
IKSS location) are order-independent and must be bit-exact; double reductions (mean, sigma,
bgnoise, IKSS scale) are summed in a different (tree) order than the sequential CPU loop and are
held to 1e-12 relative."""
import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

EXACT = ("total", "ngoodpix", "min", "max", "median", "mad", "location", "normValue")
CLOSE = ("mean", "sigma", "bgnoise", "scale")


def make_image(rng, H, W, dtype, zeros=0.01):
    yy, xx = np.mgrid[0:H, 0:W]
    img = 0.1 + 0.15 * np.exp(-((yy - H / 2) ** 2 + (xx - W / 3) ** 2) / (0.05 * H * W)) + 0.01 * rng.standard_normal((H, W))
    hot = rng.random((H, W)) < 0.002
    img[hot] += rng.uniform(0.2, 0.7, hot.sum())
    img = np.clip(img, 1e-4, 1.0)
    if zeros:
        img[rng.random((H, W)) < zeros] = 0
        if W > 20:
            img[:3, :] = 0          # black border rows, as registered frames have
            img[:, -5:] = 0
    if dtype == np.uint16:
        return np.round(img * 65535).astype(np.uint16)
    return img.astype(np.float32)


def check(got, exp):
    for k in EXACT:
        assert got[k] == exp[k], (k, got[k], exp[k])
    for k in CLOSE:
        # sigma = sqrt(sum2/n - mean^2) cancels ~1 digit and the CPU's own sequential sum over
        # ~2e6 samples carries ~1e-13 relative rounding noise: 1e-9 (it is logged, never consumed)
        tol = 1e-9 if k == "sigma" else 1e-12
        assert got[k] == pytest.approx(exp[k], rel=tol, abs=1e-300), (k, got[k], exp[k])


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("lite", [False, True], ids=["ikss", "lite"])
@pytest.mark.parametrize("shape", [(64, 96), (257, 1031), (40, 5), (1200, 1600), (24, 9576)])   # 9576: rows that need the shared-memory opt-in
def test_image_stats(orc, dtype, lite, shape):
    import siril_b200 as sb
    from siril_b200.stacking import image_stats
    rng = np.random.default_rng(shape[0] * 7 + shape[1])
    img = make_image(rng, *shape, dtype)
    ctx = sb.Context(0, torch.cuda.current_stream().cuda_stream)
    exp = orc.image_stats(img, lite)
    got_host = image_stats(ctx, img, lite).asdict()
    check(got_host, exp)
    t = torch.from_numpy(img.view(np.int16) if dtype == np.uint16 else img).cuda()
    if dtype == np.uint16:
        t = t.view(torch.uint16)
    got_dev = image_stats(ctx, t, lite).asdict()
    assert got_dev == got_host     # deterministic reductions: host-staged and resident agree bit for bit


def odd_image(rng, kind, H, W, dtype):
    """distributions whose median is far from where the shared-memory histogram window is placed"""
    if kind == "bimodal":            # 45% near 0.02, 55% near 0.9: the mean sits in the empty middle
        img = np.where(rng.random((H, W)) < 0.45, 0.02, 0.9) + 0.004 * rng.standard_normal((H, W))
    elif kind == "uniform":
        img = rng.uniform(0.001, 1.0, (H, W))
    elif kind == "spike":            # > half of the samples share one value, a few huge outliers
        img = np.full((H, W), 0.25) + (rng.random((H, W)) < 0.3) * rng.uniform(0, 0.01, (H, W))
        img[rng.random((H, W)) < 0.001] = 1.0
    elif kind == "tail":             # exponential: median well below the mean
        img = 0.001 + 0.05 * rng.exponential(1.0, (H, W))
    img = np.clip(img, 1e-4, 1.0)
    img[rng.random((H, W)) < 0.01] = 0
    if dtype == np.uint16:
        return np.round(img * 65535).astype(np.uint16)
    return img.astype(np.float32)


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("lite", [False, True], ids=["ikss", "lite"])
@pytest.mark.parametrize("kind", ["bimodal", "uniform", "spike", "tail"])
def test_image_stats_window(orc, dtype, lite, kind, monkeypatch):
    """windowed and full histograms give the same medians, whether or not the window holds the median"""
    import siril_b200 as sb
    from siril_b200.stacking import image_stats
    rng = np.random.default_rng(len(kind) * 31 + int(lite))
    img = odd_image(rng, kind, 300, 410, dtype)
    ctx = sb.Context(0, torch.cuda.current_stream().cuda_stream)
    try:
        exp = orc.image_stats(img, lite)
    except Exception:
        exp = None                   # the reference gives up on this image (null MAD): so must we
    if exp is None:
        with pytest.raises(sb.SgpuError):
            image_stats(ctx, img, lite)
    else:
        check(image_stats(ctx, img, lite).asdict(), exp)
    if kind == "bimodal":
        assert ctx.stats_window_misses > 0       # the fallback path really ran
    monkeypatch.setenv("SGPU_STATS_FULL_HIST", "1")
    full = sb.Context(0, torch.cuda.current_stream().cuda_stream)
    if exp is None:
        with pytest.raises(sb.SgpuError):
            image_stats(full, img, lite)
    else:
        check(image_stats(full, img, lite).asdict(), exp)
    assert full.stats_window_misses == 0


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("nrm", [O.ADDITIVE, O.MULTIPLICATIVE, O.ADDITIVE_SCALING, O.MULTIPLICATIVE_SCALING], ids=["add", "mul", "addscale", "mulscale"])
def test_do_normalization_and_weights(orc, dtype, nrm):
    """main_stack steps 1+3 end to end: stats -> estimators -> factors (-> noise weights) -> stack"""
    import siril_b200 as sb
    rng = np.random.default_rng(17)
    N, Cn, H, W = 7, 3, 48, 80
    fr = np.stack([np.stack([make_image(rng, H, W, dtype, zeros=0.005) for _ in range(Cn)]) for _ in range(N)])
    gains = 1 + 0.2 * rng.random(N)
    if dtype == np.uint16:
        fr = np.clip(np.round(fr * gains[:, None, None, None] * 0.8), 0, 65535).astype(np.uint16)
        dev = torch.from_numpy(fr.view(np.int16)).cuda().view(torch.uint16)
    else:
        fr = (fr * gains[:, None, None, None] * 0.8).astype(np.float32)
        dev = torch.from_numpy(fr).cuda()
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=dev, normalize=nrm, ref_image=2,
                           type_of_rejection=sb.WINSORIZED, sig=(3, 3), weighting_type=sb.NOISE_WEIGHT,
                           equalizeRGB=True, reglayer=1, create_rejmaps=True)
    assert sb.do_normalization(args) == 0
    # oracle: the same pipeline on the CPU
    off = np.zeros((Cn, N)); mul = np.ones((Cn, N)); sc = np.ones((Cn, N)); bg = np.zeros((Cn, N))
    for f in range(N):
        for l in range(Cn):
            st = orc.image_stats(fr[f, l], False)
            bg[l, f] = st["bgnoise"]
            if nrm in (O.ADDITIVE_SCALING, O.MULTIPLICATIVE_SCALING):
                sc[l, f] = st["scale"]
            if nrm in (O.ADDITIVE, O.ADDITIVE_SCALING):
                off[l, f] = st["location"]
            else:
                mul[l, f] = st["location"]
    eo, em, es = orc.compute_factors(nrm, off, mul, sc, 2, reglayer=1, equalize_rgb=True)
    for got, exp in ((args.coeff.offset, eo), (args.coeff.mul, em), (args.coeff.scale, es)):
        assert np.allclose(got, exp, rtol=1e-12, atol=0)
    np.testing.assert_allclose(args.bgnoise, bg, rtol=1e-12)
    ew = orc.noise_weights(es, bg)
    # stack with the GPU's own coefficients on both sides: stacking parity is independent of the
    # 1e-15 differences in the coefficients
    coeff = sb.NormCoeff(offset=args.coeff.offset.copy(), mul=args.coeff.mul.copy(), scale=args.coeff.scale.copy())
    assert args.method(args) == 0
    np.testing.assert_allclose(args.weights, ew, rtol=1e-11)
    o = orc.stack(fr, rejection=O.WINSORIZED, sig=(3, 3), normalize=nrm, scale=coeff.scale, offset=coeff.offset, mul=coeff.mul,
                  weights=args.weights, create_rejmaps=True)
    out = args.result.data.cpu().numpy()
    assert np.array_equal(args.result.irej, o["irej"])
    rel = np.abs(out - o["out"]) / np.maximum(np.abs(o["out"]), 1e-30)
    assert rel.max() <= 1e-6


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
def test_result_bgnoise(orc, dtype):
    """noise estimate of the stacked result (STATS_BASIC | STATS_BGNOISE only)"""
    import siril_b200 as sb
    rng = np.random.default_rng(23)
    fr = np.stack([np.stack([make_image(rng, 60, 90, dtype, zeros=0.002) for _ in range(3)]) for _ in range(9)])
    dev = torch.from_numpy(fr.view(np.int16) if dtype == np.uint16 else fr).cuda()
    if dtype == np.uint16:
        dev = dev.view(torch.uint16)
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=dev, type_of_rejection=sb.SIGMA, sig=(3, 3))
    assert args.method(args) == 0
    got = sb.result_bgnoise(args)
    res = args.result.data.cpu().numpy()
    for layer in range(3):
        exp = orc.image_stats(res[layer], True)
        assert got[layer] == pytest.approx(exp["bgnoise"], rel=1e-12)
    st = sb.stacking.image_stats(args.ctx, args.result.data[0], lite=2).asdict()
    exp = orc.image_stats(res[0], True)
    for k in ("total", "ngoodpix", "min", "max"):
        assert st[k] == exp[k]
    assert st["median"] == 0 and st["mad"] == 0
