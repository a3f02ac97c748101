"""SURVEY.md §8f row n3 on the GPU: feather masks and drizzle weights as per-sample weights of the mean stack
(median_and_mean.c:483-542 read + ramp, :1687-1692 gather, :1043-1070 weighted combine, rejection_float.c:116-126
drizzle-null removal) through the C ABI (sgpu_block_upload_weight_plane, sgpu_block_mask_ramp, sgpu_stack_block)
against the reference's own mean_and_reject (oracle/_ref) and the port.  Counts / maps / u16 bit-exact, f32 <= 1e-6."""
import numpy as np
import pytest

from oracle import oracle as O
from test_gpu_parity import ALGOS, NAMES, SIGS, compare, synth, to_dev, to_np

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def planes(rng, N, Cn, H, W):
    masks = (rng.random((N, H, W)) * 60).astype(np.float32)      # distances to the frame border, up-scaled (before the ramp)
    masks[rng.random(masks.shape) < 0.2] = 0
    masks[:, : H // 4] = 0                                         # a band every frame masks out: norm == 0 fallback (:1071-1082)
    dz = rng.random((N, Cn, H, W)).astype(np.float32)
    dz[rng.random(dz.shape) < 0.15] = 0
    return masks, dz


def run(fr, engine=0, **kw):
    import siril_b200 as sb
    masks, dz = kw.pop("masks", None), kw.pop("drizzle_weights", None)
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=to_dev(fr),
                           type_of_rejection=kw.pop("rejection", O.NO_REJEC), sig=kw.pop("sig", (3, 3)),
                           critical_value=kw.pop("crit", None), normalize=kw.pop("normalize", O.NO_NORM),
                           coeff=sb.NormCoeff(offset=kw.pop("offset", None), mul=kw.pop("mul", None), scale=kw.pop("scale", None)),
                           weights=kw.pop("weights", None), shiftx=kw.pop("shiftx", None), shifty=kw.pop("shifty", None),
                           use_32bit_output=kw.pop("use_32bit_output", True), create_rejmaps=kw.pop("create_rejmaps", False),
                           feather_dist=kw.pop("feather_dist", 0), masks=masks if masks is None else torch.from_numpy(masks).cuda(),
                           drizzle=dz is not None, drizzle_weights=dz, engine=engine)
    if args.weights is not None:
        args.weighting_type = sb.NOISE_WEIGHT
    assert not kw, kw
    rv = args.method(args)
    assert rv == 0, getattr(args, "error", rv)
    if dz is not None or engine == 1:       # drizzle weights change what a null sample is: the exact engine's job
        assert args.ctx.last_engine == "stack_exact_kernel"
    elif engine:
        assert args.ctx.last_engine == {2: "stack_sorted_kernel", 5: "stack_lsort_kernel"}[engine]
    else:
        assert args.ctx.last_engine != "stack_psel_kernel"
    r = args.result
    return dict(out=to_np(r.data), rej_low=to_np(r.rejmap_low), rej_high=to_np(r.rejmap_high), irej=r.irej)


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("algo", ALGOS, ids=NAMES)
def test_masks_and_drizzle_parity(any_oracle, algo, dtype):
    rng = np.random.default_rng(31 + algo)
    N, Cn, H, W = 14, 1, 20, 52
    fr = synth(rng, N, Cn, H, W, dtype, zeros=0.02)
    masks, dz = planes(rng, N, Cn, H, W)
    w = rng.uniform(0.5, 1.5, (Cn, N))
    for kw in (dict(masks=masks, feather_dist=25), dict(drizzle_weights=dz),
               dict(masks=masks, feather_dist=25, drizzle_weights=dz, weights=w)):
        kw.update(rejection=algo, sig=SIGS[algo], create_rejmaps=True)
        compare(run(fr, **kw), any_oracle.stack(fr, **kw), f"{NAMES[algo]} {sorted(kw)}")


@pytest.mark.parametrize("engine", [0, 1, 2, 5], ids=["auto", "exact", "sorted", "lsort"])
@pytest.mark.parametrize("algo", [O.NO_REJEC, O.SIGMA, O.WINSORIZED, O.LINEARFIT, O.GESDT], ids=["none", "sigma", "winsorized", "linearfit", "gesdt"])
@pytest.mark.parametrize("N", [12, 70])
def test_masks_through_the_fast_engines(ref, N, algo, engine):
    """feather masks only: the sorted / lane-sort engines carry them into the weighted combine (a 70-frame stack would go
    to the selection engine, which does not: the automatic choice skips it)"""
    import siril_b200 as sb
    rng = np.random.default_rng(100 + algo + N)
    Cn, H, W = 1, 16, 72
    fr = synth(rng, N, Cn, H, W, np.float32, zeros=0.01)
    masks, _ = planes(rng, N, Cn, H, W)
    sx = rng.integers(-3, 4, N) if algo == O.SIGMA else None
    kw = dict(masks=masks, feather_dist=25, rejection=algo, sig=SIGS[algo], create_rejmaps=True, weights=rng.uniform(0.5, 1.5, (Cn, N)))
    if sx is not None:
        kw["shiftx"] = sx
    try:
        g = run(fr, engine=engine, **kw)
    except AssertionError as e:
        if engine in (2, 5) and "does not cover" in str(e):
            pytest.skip("configuration outside the forced engine")
        raise
    compare(g, ref.stack(fr, **kw), f"{N} {algo} {engine}")


def test_selection_engine_refuses_masks():
    import siril_b200 as sb
    fr = torch.rand((70, 1, 8, 64), device="cuda") * 0.1 + 0.2
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=fr, type_of_rejection=O.SIGMA, feather_dist=5,
                           masks=torch.ones((70, 8, 64), device="cuda") * 3, engine=6)
    assert args.method(args) == sb.ST_GENERIC_ERROR and "mask" in args.error


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
def test_masks_shifts_rgb_norm(ref, dtype):
    """the planes travel with the frames: y shift at upload, x shift at the gather, three layers (one mask per frame
    for all of them, drizzle weights per layer), normalization, 16-bit output"""
    rng = np.random.default_rng(5)
    N, Cn, H, W = 9, 3, 18, 40
    fr = synth(rng, N, Cn, H, W, dtype)
    masks, dz = planes(rng, N, Cn, H, W)
    sx, sy = rng.integers(-4, 5, N), rng.integers(-3, 4, N)
    sx[0] = sy[0] = 0
    scale = 1 + 0.05 * rng.standard_normal((Cn, N))
    offset = 0.01 * rng.standard_normal((Cn, N)) * (65535 if dtype == np.uint16 else 1)
    kw = dict(rejection=O.WINSORIZED, sig=(3, 3), create_rejmaps=True, masks=masks, feather_dist=10, drizzle_weights=dz,
              shiftx=sx, shifty=sy, normalize=O.ADDITIVE_SCALING, scale=scale, offset=offset, mul=np.ones((Cn, N)),
              use_32bit_output=dtype == np.float32)
    compare(run(fr, **kw), ref.stack(fr, **kw), "shifted rgb")


def test_mask_ramp_values(ref):
    """sgpu_block_mask_ramp against the reference's own ramp lines (via a stack whose result exposes the weights)"""
    rng = np.random.default_rng(9)
    N, H, W = 2, 8, 1001
    fr = np.empty((N, 1, H, W), dtype=np.float32)
    fr[0], fr[1] = 0.25, 0.75
    masks = np.empty((N, H, W), dtype=np.float32)
    masks[0] = 30.0                                              # weight 1
    masks[1] = (np.arange(H * W, dtype=np.float32).reshape(H, W) % 1001) * np.float32(0.02) * rng.uniform(0.9, 1.1)
    kw = dict(rejection=O.NO_REJEC, masks=masks, feather_dist=20)
    g, o = run(fr, **kw), ref.stack(fr, **kw)
    assert np.array_equal(g["out"], o["out"])                    # same weights to the last bit => same weighted mean
    assert len(np.unique(o["out"])) > 500


def test_missing_planes_is_a_sequence_error():
    import siril_b200 as sb
    fr = torch.zeros((3, 1, 4, 8), device="cuda") + 0.5
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=fr, feather_dist=10)
    assert args.method(args) == sb.ST_SEQUENCE_ERROR
