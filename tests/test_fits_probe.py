"""CPU checks of the FITS leg: the header probe of libsgpu.so (host code, no GPU) against the
oracle's parser, and the oracle's decode against hand-made big-endian vectors."""
import ctypes as C
import os

import numpy as np
import pytest


@pytest.mark.parametrize("dtype,c", [(np.uint16, 1), (np.int16, 3), (np.uint8, 1), (np.float32, 3)])
def test_probe_matches_oracle_parser(tmp_path, dtype, c):
    from siril_b200 import _lib
    from oracle import fits_ref
    rng = np.random.default_rng(3)
    a = (rng.random((c, 9, 14)) * 200).astype(dtype)
    p = str(tmp_path / "x.fit")
    dm = 65535.0 if dtype == np.float32 else None
    fits_ref.write_fits(p, a, datamax=dm, extra_cards=[f"{'COMMENT':<8} = not a value", f"{'EXPTIME':<8}= {30.0:>20.1f}"] * 20)
    info = _lib.FitsInfo()
    assert _lib.lib().sgpu_fits_probe(os.fsencode(p), C.byref(info)) == 0
    bitpix, (w, h, cc), bzero, datamax, off = fits_ref.parse_header(p)
    assert (info.bitpix, info.naxes[0], info.naxes[1], info.naxes[2]) == (bitpix, w, h, cc)
    assert info.bzero == bzero and info.bscale == 1.0 and info.data_offset == off and off == 2880 * 2
    assert bool(info.has_datamax) == (datamax is not None) and (datamax is None or info.datamax == datamax)
    assert info.device_readable == 1


def test_probe_rejects_non_fits_and_truncated(tmp_path):
    from siril_b200 import _lib
    from oracle import fits_ref
    L = _lib.lib()
    info = _lib.FitsInfo()
    p = str(tmp_path / "junk.fit")
    open(p, "wb").write(b"not a fits file" * 400)
    assert L.sgpu_fits_probe(os.fsencode(p), C.byref(info)) == -2
    assert L.sgpu_fits_probe(os.fsencode(str(tmp_path / "missing.fit")), C.byref(info)) == -2
    q = str(tmp_path / "short.fit")
    off = fits_ref.write_fits(q, np.zeros((1, 64, 64), np.float32))
    os.truncate(q, off + 100)
    assert L.sgpu_fits_probe(os.fsencode(q), C.byref(info)) == -2
    r = str(tmp_path / "scaled.fit")
    fits_ref.write_fits(r, np.zeros((1, 8, 8), np.int16), bzero=100)
    assert L.sgpu_fits_probe(os.fsencode(r), C.byref(info)) == 0 and info.device_readable == 0


def test_oracle_decode_known_bytes(tmp_path):
    """FITS standard: BITPIX 16 big-endian two's complement, physical = BZERO + raw"""
    from oracle import fits_ref
    p = str(tmp_path / "k.fit")
    fits_ref.write_fits(p, np.array([[0, 1, 32768, 65535]], np.uint16))
    raw = open(p, "rb").read()[2880:2888]
    assert raw == bytes([0x80, 0x00, 0x80, 0x01, 0x00, 0x00, 0x7f, 0xff])
    assert fits_ref.read_block([p])[0, 0, 0].tolist() == [0, 1, 32768, 65535]
    fits_ref.write_fits(p, np.array([[-32768, -1, 0, 32767]], np.int16))
    assert fits_ref.read_block([p])[0, 0, 0].tolist() == [0, 32767, 32768, 65535]    # convert_data_ushort SHORT_IMG
    fits_ref.write_fits(p, np.array([[1.0, 65535.0, 0.5, 20.0]], np.float32), datamax=65535.0)
    got = fits_ref.read_block([p])[0, 0, 0]
    assert got[1] == np.float32(65535.0) * np.float32(.000015259022)


def test_fitseq_probe_follows_find_hdus(tmp_path):
    """FITS sequence files (io/fits_sequence.c:39-150): which HDUs are frames, and where their data units start"""
    from siril_b200 import fitsread
    from oracle import fits_ref
    rng = np.random.default_rng(8)
    fr = [(rng.random((1, 11, 17)) * 60000).astype(np.uint16) for _ in range(4)]
    p = str(tmp_path / "seq.fit")
    hdus = [(fr[0], None, None), ("table", 3000), (fr[1], "SCI", None), ((rng.random((3, 11, 17)) * 100).astype(np.uint16), None, None),
            ((rng.random((1, 5, 5)) * 100).astype(np.uint16), "Thumbnail", None), (fr[2], None, None),
            (rng.random((1, 11, 17)).astype(np.float32), None, 1.0), (fr[3], "ICCProfile", None)]
    fits_ref.write_fitseq(p, hdus)
    infos = fitsread.fitseq_probe(p)
    assert len(infos) == 3          # the RGB HDU, the thumbnail, the float HDU and the ICC profile are not frames; the table is stepped over
    raw = open(p, "rb").read()
    for info, want in zip(infos, fr[:3]):
        assert (info.bitpix, info.bzero, info.naxes[0], info.naxes[1], info.naxes[2], info.device_readable) == (16, 32768.0, 17, 11, 1, 1)
        got = np.frombuffer(raw, dtype=">i2", count=11 * 17, offset=info.data_offset).astype(np.int32) + 32768
        assert np.array_equal(got.reshape(11, 17), want[0])
    # an image of another size fails the file ("different parameters, which is not allowed")
    q = str(tmp_path / "bad.fit")
    fits_ref.write_fitseq(q, [(fr[0], None, None), ((rng.random((1, 12, 17)) * 100).astype(np.uint16), None, None)])
    with pytest.raises(Exception):
        fitsread.fitseq_probe(q)
    # image in the primary HDU, extensions after it
    r = str(tmp_path / "prim.fit")
    fits_ref.write_fitseq(r, [(fr[0], None, None), (fr[1], None, None)], empty_primary=False)
    assert len(fitsread.fitseq_probe(r)) == 2
