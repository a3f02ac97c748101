"""GPU parity of the raw FITS block reader: pread -> pinned staging -> H2D ->
device decode, against the numpy restatement of internal_read_partial_fits
(oracle/fits_ref.py).  Byte / integer work and a single float multiply: bit exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def to_np(t):
    t = t.cpu()
    return t.view(torch.int16).numpy().view(np.uint16) if t.dtype == torch.uint16 else t.numpy()


def same_samples(got, want):
    """bit-for-bit, except that a NaN is a NaN: the device multiply returns the canonical quiet NaN
    where x86 keeps the payload (only the DATAMAX rescale touches the bits at all)"""
    if got.dtype != want.dtype or got.shape != want.shape:
        return False
    if got.dtype != np.float32:
        return np.array_equal(got, want)
    gn, wn = np.isnan(got), np.isnan(want)
    return np.array_equal(gn, wn) and np.array_equal(got.view(np.uint32)[~gn], want.view(np.uint32)[~wn])


def make(rng, kind, c, h, w):
    if kind == "u16":
        return rng.integers(0, 65536, (c, h, w)).astype(np.uint16), {}
    if kind == "i16":
        return rng.integers(-32768, 32768, (c, h, w)).astype(np.int16), {}
    if kind == "u8":
        return rng.integers(0, 256, (c, h, w)).astype(np.uint8), {}
    a = rng.random((c, h, w)).astype(np.float32)
    a[0, 0, :2] = [np.nan, -0.5] if w >= 2 else a[0, 0, :2]
    if kind == "f32":
        return a, {"datamax": 1.0}
    if kind == "f32_nokey":
        return a, {}
    if kind == "f32_datamax":
        return a * np.float32(65535), {"datamax": 65535.0}
    return a * np.float32(60000) + np.float32(11), {}        # f32_big_nokey: the three-sample guess


@pytest.mark.parametrize("kind", ["u16", "i16", "u8", "f32", "f32_nokey", "f32_datamax", "f32_big_nokey"])
@pytest.mark.parametrize("shape", [(1, 37, 53), (3, 24, 130), (1, 5, 1031), (1, 2, 1)])
@pytest.mark.parametrize("source", ["fd", "mmap", "pinned"])
def test_fits_block_reader(tmp_path, kind, shape, source):
    import siril_b200 as sb
    from siril_b200 import fitsread
    from oracle import fits_ref
    rng = np.random.default_rng(sum(shape) + len(kind))
    c, h, w = shape
    n = 4
    paths = []
    for f in range(n):
        a, kw = make(rng, kind, c, h, w)
        p = str(tmp_path / f"img_{f:05d}.fit")
        fits_ref.write_fits(p, a, **kw)
        paths.append(p)
    ctx = sb.Context(0)
    for row0, hh, sy in [(0, None, None), (0, None, [0, 3, -2, h + 4]), (h // 3, max(1, h // 2), [1, 0, -1, 2])]:
        got = to_np(fitsread.read_block_data(ctx, paths, row0, hh, sy, source))
        want = fits_ref.read_block(paths, row0, hh, sy)
        assert same_samples(got, want), (row0, hh, sy)
    ctx.close()


@pytest.mark.parametrize("kind", ["u16", "f32"])
def test_fits_reader_multi_chunk(tmp_path, kind):
    """frames larger than two 32 MiB staging chunks: the double buffer is recycled"""
    import siril_b200 as sb
    from siril_b200 import fitsread
    from oracle import fits_ref
    rng = np.random.default_rng(5)
    c, h, w = 1, (4100 if kind == "f32" else 5000), 5003
    paths = []
    for f in range(2):
        a = rng.integers(0, 65536, (c, h, w)).astype(np.uint16)
        if kind == "f32":
            a = (a / np.float32(65535)).astype(np.float32)
        p = str(tmp_path / f"big_{f}.fit")
        fits_ref.write_fits(p, a, datamax=1.0 if kind == "f32" else None)
        paths.append(p)
    ctx = sb.Context(0)
    got = to_np(fitsread.read_block_data(ctx, paths, 7, h - 11, [0, 5]))
    want = fits_ref.read_block(paths, 7, h - 11, [0, 5])
    assert same_samples(got, want)
    ctx.close()


def test_stack_from_fits_equals_stack_from_arrays(tmp_path, orc):
    """the block read from FITS files feeds the stacker like an uploaded block"""
    import siril_b200 as sb
    from siril_b200 import fitsread
    from oracle import fits_ref
    rng = np.random.default_rng(11)
    N, C, H, W = 12, 1, 40, 75
    fr = (0.1 + 0.01 * rng.standard_normal((N, C, H, W))).astype(np.float32)
    fr[rng.random(fr.shape) < 0.01] += 0.4
    paths = []
    for f in range(N):
        p = str(tmp_path / f"s_{f:05d}.fit")
        fits_ref.write_fits(p, fr[f], datamax=1.0)
        paths.append(p)
    ctx = sb.Context(0)
    blk = fitsread.read_block_data(ctx, paths)
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=blk, type_of_rejection=sb.SIGMA,
                           sig=(3.0, 3.0), create_rejmaps=True)
    assert sb.main_stack(args) == 0
    ref = orc.stack(fr, rejection=sb.SIGMA, sig=(3.0, 3.0), create_rejmaps=True)
    out = args.result.data.cpu().numpy()
    assert np.array_equal(args.result.irej, ref["irej"])
    rel = np.abs(out - ref["out"]) / np.maximum(np.abs(ref["out"]), 1e-30)
    assert rel.max() <= 1e-6


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
@pytest.mark.parametrize("nblk", [1, 3])
def test_stack_fits_files(tmp_path, orc, monkeypatch, dtype, nblk):
    """sgpu_stack_fits: files on disk -> stacked image, row blocks filled by the raw reader while the
    block before is stacked; against the oracle on the decoded frames"""
    import siril_b200 as sb
    from siril_b200 import fitsread
    from oracle import fits_ref
    monkeypatch.setenv("SGPU_HOST_BLOCKS", str(nblk))
    rng = np.random.default_rng(21 + nblk)
    N, C, H, W = 15, 3, 26, 90
    fr = 0.1 + 0.01 * rng.standard_normal((N, C, H, W))
    fr[rng.random(fr.shape) < 0.01] += 0.4
    fr = fr.astype(np.float32) if dtype == np.float32 else np.round(fr * 65535).astype(np.uint16)
    paths = []
    for f in range(N):
        p = str(tmp_path / f"q_{f:05d}.fit")
        fits_ref.write_fits(p, fr[f], datamax=1.0 if dtype == np.float32 else None)
        paths.append(p)
    sy = rng.integers(-3, 4, N).astype(np.int32)
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=None, type_of_rejection=sb.SIGMA,
                           sig=(3.0, 3.0), create_rejmaps=True, shifty=sy)
    assert fitsread.stack_fits(args, paths) == 0, getattr(args, "error", "")
    ref = orc.stack(fr, rejection=sb.SIGMA, sig=(3.0, 3.0), create_rejmaps=True, shifty=sy)
    assert np.array_equal(args.result.irej, ref["irej"])
    assert np.array_equal(args.result.rejmap_low, ref["rej_low"]) and np.array_equal(args.result.rejmap_high, ref["rej_high"])
    rel = np.abs(args.result.data - ref["out"]) / np.maximum(np.abs(ref["out"]), 1e-30)
    assert rel.max() <= 1e-6


@pytest.mark.parametrize("nblk", [1, 3])
def test_stack_fits_byte_images_normalize_to16bit(tmp_path, orc, monkeypatch, nblk):
    """8-bit sequences with a 16-bit output and output_norm: the reference scales the result by 65535/255
    before rounding (normalize_to16bit, median_and_mean.c:547-555, 1729-1733); without output_norm it does not"""
    import siril_b200 as sb
    from siril_b200 import fitsread
    from oracle import fits_ref
    monkeypatch.setenv("SGPU_HOST_BLOCKS", str(nblk))
    rng = np.random.default_rng(77)
    N, C, H, W = 12, 1, 21, 64
    fr = np.clip(np.round(40 + 6 * rng.standard_normal((N, C, H, W))), 1, 255).astype(np.uint8)
    fr[rng.random(fr.shape) < 0.01] = 250
    paths = []
    for f in range(N):
        p = str(tmp_path / f"b_{f:05d}.fit")
        fits_ref.write_fits(p, fr[f])
        paths.append(p)
    wide = fr.astype(np.uint16)      # BYTE_IMG is read into WORD buffers as it is (image_format_fits.c: convert_data_ushort)
    for output_norm in (True, False):
        args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=None, type_of_rejection=sb.SIGMA, sig=(3.0, 3.0),
                               use_32bit_output=False, output_norm=output_norm)
        assert fitsread.stack_fits(args, paths) == 0, getattr(args, "error", "")
        ref = orc.stack(wide, rejection=sb.SIGMA, sig=(3.0, 3.0), use_32bit_output=False, output_norm=output_norm, source_bitpix=8)
        assert np.array_equal(args.result.irej, ref["irej"])
        assert np.array_equal(args.result.data, ref["out"]), output_norm
        assert (args.result.data.max() > 255) == output_norm


def test_stack_fits_float_range_guess_is_per_file(tmp_path, orc, monkeypatch):
    """float files without DATAMAX whose samples exceed 10 only in some rows: the 1/65535 rescale is decided
    once per file (the samples the reference looks at when it reads the first layer whole), not per row block"""
    import siril_b200 as sb
    from siril_b200 import fitsread
    from oracle import fits_ref
    rng = np.random.default_rng(78)
    N, H, W = 6, 24, 48
    fr = (3000.0 + 300 * rng.standard_normal((N, 1, H, W))).astype(np.float32)     # ADU-like floats
    fr[:, :, :8, :] = 5.0          # a dark band: a block made of these rows alone would guess "already in [0, 1]"
    paths = []
    for f in range(N):
        p = str(tmp_path / f"g_{f:05d}.fit")
        fits_ref.write_fits(p, fr[f])      # no DATAMAX
        paths.append(p)
    outs = []
    for nblk in (1, 3):
        monkeypatch.setenv("SGPU_HOST_BLOCKS", str(nblk))
        args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=None, type_of_rejection=sb.NO_REJEC, output_norm=True)
        assert fitsread.stack_fits(args, paths) == 0, getattr(args, "error", "")
        outs.append(args.result.data.copy())
    assert np.array_equal(outs[0], outs[1])
    # whole-layer sampling: indices 0, n/3, 2n/3 (, n - n%3): sample 0 is 5.0, the others ~3000 -> rescaled everywhere
    scaled = fr * np.float32(1.0 / 65535.0)
    ref = orc.stack(scaled.astype(np.float32), rejection=sb.NO_REJEC, output_norm=True)
    rel = np.abs(outs[0] - ref["out"]) / np.maximum(np.abs(ref["out"]), 1e-30)
    assert rel.max() <= 1e-6


def test_stack_fits_rejects_mixed_sequences(tmp_path):
    import siril_b200 as sb
    from siril_b200 import fitsread
    from oracle import fits_ref
    a, b = str(tmp_path / "a.fit"), str(tmp_path / "b.fit")
    fits_ref.write_fits(a, np.full((1, 8, 8), 0.5, np.float32), datamax=1.0)
    fits_ref.write_fits(b, np.full((1, 8, 9), 0.5, np.float32), datamax=1.0)
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=None)
    assert fitsread.stack_fits(args, [a, b, a]) == sb.ST_SEQUENCE_ERROR
    assert fitsread.stack_fits(args, [a, str(tmp_path / "missing.fit")]) == sb.ST_SEQUENCE_ERROR


@pytest.mark.parametrize("dtype", [np.float32, np.uint16], ids=["f32", "u16"])
def test_stack_fitseq_file(tmp_path, orc, dtype):
    """sgpu_stack_fitseq: the frames of ONE FITS sequence file (one image per HDU), a subset of them by index, through the
    same raw reader; against the oracle on the decoded frames"""
    import siril_b200 as sb
    from siril_b200 import fitsread
    from oracle import fits_ref
    rng = np.random.default_rng(33)
    N, C, H, W = 11, 1, 30, 70
    fr = 0.1 + 0.01 * rng.standard_normal((N, C, H, W))
    fr[rng.random(fr.shape) < 0.01] += 0.4
    fr = fr.astype(np.float32) if dtype == np.float32 else np.round(fr * 65535).astype(np.uint16)
    p = str(tmp_path / "seq.fit")
    hdus = []
    for f in range(N):
        hdus.append((fr[f], None, 1.0 if dtype == np.float32 else None))
        if f == 3:
            hdus.append(("table", 5000))
            hdus.append(((rng.random((1, 4, 4)) * 10).astype(fr.dtype), "Thumbnail", None))
    fits_ref.write_fitseq(p, hdus)
    idx = [0, 2, 3, 4, 5, 7, 8, 10]
    sy = rng.integers(-2, 3, len(idx)).astype(np.int32)
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=None, type_of_rejection=sb.SIGMA,
                           sig=(3.0, 3.0), create_rejmaps=True, shifty=sy)
    assert fitsread.stack_fits(args, p, frame_indices=idx) == 0, getattr(args, "error", "")
    ref = orc.stack(fr[idx], rejection=sb.SIGMA, sig=(3.0, 3.0), create_rejmaps=True, shifty=sy)
    assert np.array_equal(args.result.irej, ref["irej"])
    assert np.array_equal(args.result.rejmap_low, ref["rej_low"]) and np.array_equal(args.result.rejmap_high, ref["rej_high"])
    rel = np.abs(args.result.data - ref["out"]) / np.maximum(np.abs(ref["out"]), 1e-30)
    assert rel.max() <= 1e-6


@pytest.mark.parametrize("kind", ["u8", "u16le", "u16be", "u12le"])
def test_stack_ser_file(tmp_path, ref, kind):
    """sgpu_stack_ser: frames of a SER file (rows top-down, 8 / 16-bit, either byte order) through the raw reader; the
    samples the stack must see are the reference's own conversion (ser_manage_endianess_and_depth, compiled into
    oracle/_ref) of the bytes read_area_from_image reads, turned bottom-up as Siril holds images"""
    import ctypes as C
    import siril_b200 as sb
    from siril_b200 import fitsread
    from oracle import fits_ref
    rng = np.random.default_rng(44)
    N, H, W = 12, 37, 90
    depth = {"u8": 8, "u16le": 16, "u16be": 16, "u12le": 12}[kind]
    level = 0.3 + 0.02 * rng.standard_normal((N, H, W))
    level[rng.random(level.shape) < 0.01] += 0.4
    disp = np.clip(np.round(level * (2 ** depth - 1)), 1, 2 ** depth - 1).astype(np.uint8 if depth == 8 else np.uint16)
    p = str(tmp_path / "movie.ser")
    fits_ref.write_ser(p, disp, depth, big_endian=kind == "u16be")
    si = sb._lib.SerInfo()
    assert sb._lib.lib().sgpu_ser_probe(p.encode(), C.byref(si)) == 0
    assert (si.width, si.height, si.frame_count, si.byte_depth, si.planes, si.device_readable) == (W, H, N, 1 if depth == 8 else 2, 1, 1)
    # the reference's view of the file: raw bytes of every frame -> WORDs by its own conversion, then bottom-up
    raw = open(p, "rb").read()[178:]
    bpp = si.byte_depth
    frames = np.zeros((N, 1, H, W), dtype=np.uint16)
    R = ref.lib
    R.ref_ser_convert.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_long]
    for f in range(N):
        buf = np.zeros(H * W, dtype=np.uint16)
        chunk = np.frombuffer(raw, dtype=np.uint8, count=H * W * bpp, offset=f * H * W * bpp)
        buf.view(np.uint8)[: chunk.size] = chunk
        R.ref_ser_convert(bpp, si.little_endian, buf.ctypes.data, H * W)
        frames[f, 0] = buf.reshape(H, W)[::-1]
    assert np.array_equal(frames[:, 0, ::-1], disp.astype(np.uint16))
    idx = [0, 1, 2, 4, 5, 6, 8, 9, 10, 11]
    sy = rng.integers(-3, 4, len(idx)).astype(np.int32)
    for out32 in (True, False):
        args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=None, type_of_rejection=sb.WINSORIZED,
                               sig=(3.0, 3.0), create_rejmaps=True, shifty=sy, use_32bit_output=out32)
        assert fitsread.stack_ser(args, p, frame_indices=idx) == 0, getattr(args, "error", "")
        o = ref.stack(frames[idx], rejection=sb.WINSORIZED, sig=(3.0, 3.0), create_rejmaps=True, shifty=sy, use_32bit_output=out32,
                      source_bitpix=8 if depth == 8 else 16)
        assert np.array_equal(args.result.irej, o["irej"])
        assert np.array_equal(args.result.rejmap_low, o["rej_low"]) and np.array_equal(args.result.rejmap_high, o["rej_high"])
        if out32:
            rel = np.abs(args.result.data - o["out"]) / np.maximum(np.abs(o["out"]), 1e-30)
            assert rel.max() <= 1e-6
        else:
            assert np.array_equal(args.result.data, o["out"])


def test_ser_probe_rejects_colour_and_truncated(tmp_path):
    import ctypes as C
    import siril_b200 as sb
    from oracle import fits_ref
    L = sb._lib.lib()
    si = sb._lib.SerInfo()
    p = str(tmp_path / "rgb.ser")
    fits_ref.write_ser(p, np.zeros((2, 4, 12), dtype=np.uint8), 8, color_id=100)      # RGB: 3 planes, the file is too short for that
    assert L.sgpu_ser_probe(p.encode(), C.byref(si)) == -2
    q = str(tmp_path / "bayer.ser")
    fits_ref.write_ser(q, np.ones((2, 4, 12), dtype=np.uint8), 8, color_id=8)         # a colour filter array: read as mono
    assert L.sgpu_ser_probe(q.encode(), C.byref(si)) == 0 and si.device_readable == 1
    assert L.sgpu_ser_probe(str(tmp_path / "none.ser").encode(), C.byref(si)) == -2
