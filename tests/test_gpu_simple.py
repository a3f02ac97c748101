"""GPU parity of the sum / max / min stackers (sum.c, stackminmax.c) through the C ABI, against
the reference's own loops (oracle/_ref, ref_simple.c) and the restatement.  Integer work (WORD
sums, min, max) is bit exact; float sums are added in frame order on both sides, so they are bit
exact as well."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def frames_of(rng, N, C, H, W, dtype):
    fr = 0.05 + 0.9 * rng.random((N, C, H, W))
    fr[rng.random(fr.shape) < 0.01] = 0
    if dtype == np.uint16:
        return np.round(fr * 65535).astype(np.uint16)
    fr = fr.astype(np.float32)
    fr[0, 0, 0, :3] = [-0.25, 1.5, 0.0]        # out-of-range samples go through unclamped
    return fr


def run(frames, op, shiftx, shifty, out32, on_device):
    import siril_b200 as sb
    fn = {"sum": sb.stack_summing_generic, "max": sb.stack_addmax, "min": sb.stack_addmin}[op]
    src = frames
    if on_device:
        t = torch.from_numpy(frames.view(np.int16) if frames.dtype == np.uint16 else frames).cuda()
        src = t.view(torch.uint16) if frames.dtype == np.uint16 else t
    args = sb.StackingArgs(method=fn, frames=src, shiftx=shiftx, shifty=shifty, use_32bit_output=out32)
    rv = args.method(args)
    if rv:
        return None
    d = args.result.data
    if on_device:
        d = d.cpu()
        d = d.view(torch.int16).numpy().view(np.uint16) if d.dtype == torch.uint16 else d.numpy()
    return d


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("op", ["sum", "max", "min"])
@pytest.mark.parametrize("shape", [(9, 1, 37, 53), (20, 3, 24, 130), (3, 1, 5, 1031), (70, 1, 16, 64)])
@pytest.mark.parametrize("shifted", [False, True], ids=["aligned", "shifted"])
def test_simple_stackers(any_oracle, dtype, op, shape, shifted):
    rng = np.random.default_rng(sum(shape) + len(op))
    N, C, H, W = shape
    fr = frames_of(rng, N, C, H, W, dtype)
    sx = rng.integers(-7, 8, N).astype(np.int32) if shifted else None
    sy = rng.integers(-4, 5, N).astype(np.int32) if shifted else None
    if shifted:
        sx[0] = W + 3           # a frame shifted entirely out of the image contributes nothing
    for out32 in ((True, False) if (op == "sum" and dtype == np.uint16) else (True,)):
        exp = any_oracle.stack_simple(fr, op, sx, sy, out32)
        for on_device in (True, False):
            got = run(fr, op, sx, sy, out32, on_device)
            assert got is not None and got.dtype == exp.dtype
            assert np.array_equal(got, exp), (op, out32, on_device, np.abs(got.astype(np.float64) - exp).max())


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("op", ["sum", "max", "min"])
@pytest.mark.parametrize("shape", [(9, 1, 37, 56), (20, 3, 24, 136), (70, 1, 16, 64), (5, 1, 3, 8)])
@pytest.mark.parametrize("yshift", [False, True], ids=["aligned", "yshifted"])
def test_simple_stackers_wide_loads(any_oracle, dtype, op, shape, yshift):
    """no x shift and a width that is a multiple of 16 bytes: the 16-byte-per-thread kernel"""
    rng = np.random.default_rng(sum(shape) + len(op) + 1)
    N, C, H, W = shape
    fr = frames_of(rng, N, C, H, W, dtype)
    sy = rng.integers(-4, 5, N).astype(np.int32) if yshift else None
    sx = np.zeros(N, np.int32) if yshift else None
    if yshift:
        sy[0] = H + 2
    for out32 in ((True, False) if (op == "sum" and dtype == np.uint16) else (True,)):
        exp = any_oracle.stack_simple(fr, op, sx, sy, out32)
        for on_device in (True, False):
            got = run(fr, op, sx, sy, out32, on_device)
            assert got is not None and got.dtype == exp.dtype
            assert np.array_equal(got, exp), (op, out32, on_device, np.abs(got.astype(np.float64) - exp).max())


def test_sum_saturates_to_word(ref):
    """WORD sums above 65535 are rescaled by 65535 / max (sum.c:329-341)"""
    rng = np.random.default_rng(5)
    fr = rng.integers(20000, 65535, (40, 1, 16, 48)).astype(np.uint16)
    exp = ref.stack_simple(fr, "sum", None, None, False)
    got = run(fr, "sum", None, None, False, True)
    assert exp.max() == 65535 and np.array_equal(got, exp)


def test_empty_sum_is_an_error(ref):
    fr = np.zeros((4, 1, 8, 16), np.float32)
    assert ref.stack_simple(fr, "sum") is None
    assert run(fr, "sum", None, None, True, True) is None


def test_sum_at_bench_size_is_linear():
    """61 MP x 32 frames of WORD data: sum of constant frames is the constant image scaled to the maximum"""
    import siril_b200 as sb
    N, H, W = 32, 2048, 4096
    base = torch.arange(H * W, device="cuda", dtype=torch.int32).remainder(60000).to(torch.uint16).view(1, 1, H, W)
    fr = base.expand(N, 1, H, W).contiguous()
    args = sb.StackingArgs(method=sb.stack_summing_generic, frames=fr, use_32bit_output=True)
    assert args.method(args) == 0
    exp = (base[0].to(torch.float64) * N * (1.0 / (59999.0 * N))).to(torch.float32)
    assert torch.equal(args.result.data, exp)
