"""TEST INFRASTRUCTURE (oracle): CPU restatement of the FITS read leg of the block reader, numpy.

Follows internal_read_partial_fits (src/io/image_format_fits.c:927-1015) for uncompressed
primary-HDU images: cfitsio's fits_read_subset is the FITS standard's big-endian decode with BZERO
applied (cfitsio is an un-vendored, unpinned system dependency of the reference and absent here:
parity of this leg is pinned on the FITS standard's definition of BITPIX / BZERO, not on cfitsio
output — "parity unpinned" against cfitsio itself), followed by convert_data_ushort (:390-416) and,
for float data with DATAMAX > 10 — or, without the key, with three or four samples of the region
above 10 (:997-1003) — convert_floats (:648-672).  Also writes the minimal FITS files the tests use.
Only tests/ may import this module."""
import numpy as np

INV_USHRT_MAX_SINGLE = np.float32(.000015259022)   # core/siril.h:94


def write_fits(path, data, bzero=None, datamax=None, extra_cards=()):
    """data: [layer][row][col] (or [row][col]) in file order; uint16 is stored as BITPIX 16 with
    BZERO 32768 (the FITS convention cfitsio's USHORT_IMG stands for), int16 as BITPIX 16, uint8 as
    BITPIX 8, float32 as BITPIX -32."""
    a = np.asarray(data)
    if a.ndim == 2:
        a = a[None]
    c, h, w = a.shape
    if a.dtype == np.uint16:
        bitpix, raw = 16, (a.astype(np.int32) - 32768).astype(">i2")
        bzero = 32768 if bzero is None else bzero
    elif a.dtype == np.int16:
        bitpix, raw = 16, a.astype(">i2")
    elif a.dtype == np.uint8:
        bitpix, raw = 8, a
    elif a.dtype == np.float32:
        bitpix, raw = -32, a.astype(">f4")
    else:
        raise TypeError(a.dtype)
    cards = [f"{'SIMPLE':<8}= {'T':>20}", f"{'BITPIX':<8}= {bitpix:>20d}", f"{'NAXIS':<8}= {3 if c > 1 else 2:>20d}",
             f"{'NAXIS1':<8}= {w:>20d}", f"{'NAXIS2':<8}= {h:>20d}"]
    if c > 1:
        cards.append(f"{'NAXIS3':<8}= {c:>20d}")
    if bzero is not None:
        cards += [f"{'BZERO':<8}= {float(bzero):>20.1f} / offset data range to that of unsigned short", f"{'BSCALE':<8}= {1.0:>20.1f}"]
    if datamax is not None:
        cards.append(f"{'DATAMAX':<8}= {float(datamax):>20.6E}")
    cards += list(extra_cards) + ["END"]
    hdr = "".join(f"{x:<80}" for x in cards)
    hdr += " " * (-len(hdr) % 2880)
    body = raw.tobytes()
    with open(path, "wb") as f:
        f.write(hdr.encode("ascii"))
        f.write(body)
        f.write(b"\0" * (-len(body) % 2880))
    return len(hdr)


def _hdu_bytes(data, primary, extname=None, bzero=None, datamax=None):
    a = np.asarray(data)
    if a.ndim == 2:
        a = a[None]
    c, h, w = a.shape
    if a.dtype == np.uint16:
        bitpix, raw = 16, (a.astype(np.int32) - 32768).astype(">i2")
        bzero = 32768 if bzero is None else bzero
    elif a.dtype == np.int16:
        bitpix, raw = 16, a.astype(">i2")
    elif a.dtype == np.uint8:
        bitpix, raw = 8, a
    elif a.dtype == np.float32:
        bitpix, raw = -32, a.astype(">f4")
    else:
        raise TypeError(a.dtype)
    first = f"{'SIMPLE':<8}= {'T':>20}" if primary else f"{'XTENSION':<8}= {repr('IMAGE   '):<20}"
    cards = [first, f"{'BITPIX':<8}= {bitpix:>20d}", f"{'NAXIS':<8}= {3 if c > 1 else 2:>20d}",
             f"{'NAXIS1':<8}= {w:>20d}", f"{'NAXIS2':<8}= {h:>20d}"]
    if c > 1:
        cards.append(f"{'NAXIS3':<8}= {c:>20d}")
    if not primary:
        cards += [f"{'PCOUNT':<8}= {0:>20d}", f"{'GCOUNT':<8}= {1:>20d}"]
    if bzero is not None:
        cards += [f"{'BZERO':<8}= {float(bzero):>20.1f}", f"{'BSCALE':<8}= {1.0:>20.1f}"]
    if datamax is not None:
        cards.append(f"{'DATAMAX':<8}= {float(datamax):>20.6E}")
    if extname:
        cards.append(f"{'EXTNAME':<8}= {repr(extname):<20}")
    cards.append("END")
    hdr = "".join(f"{x:<80}" for x in cards)
    hdr += " " * (-len(hdr) % 2880)
    body = raw.tobytes()
    return hdr.encode("ascii") + body + b"\0" * (-len(body) % 2880)


def write_fitseq(path, hdus, empty_primary=True):
    """a FITS sequence file (SEQ_FITSEQ, io/fits_sequence.c): `hdus` is a list of (data [layer][row][col], extname or None,
    datamax or None) image HDUs, or ("table", nbytes) for a binary table the reader must step over; with empty_primary the
    primary HDU holds no data (NAXIS = 0), as astronomy pipelines write it"""
    with open(path, "wb") as f:
        first = True
        if empty_primary:
            cards = [f"{'SIMPLE':<8}= {'T':>20}", f"{'BITPIX':<8}= {8:>20d}", f"{'NAXIS':<8}= {0:>20d}", f"{'EXTEND':<8}= {'T':>20}", "END"]
            hdr = "".join(f"{x:<80}" for x in cards)
            f.write((hdr + " " * (-len(hdr) % 2880)).encode("ascii"))
            first = False
        for h in hdus:
            if isinstance(h[0], str) and h[0] == "table":
                nbytes = int(h[1])
                cards = [f"{'XTENSION':<8}= {repr('BINTABLE'):<20}", f"{'BITPIX':<8}= {8:>20d}", f"{'NAXIS':<8}= {2:>20d}",
                         f"{'NAXIS1':<8}= {nbytes:>20d}", f"{'NAXIS2':<8}= {1:>20d}", f"{'PCOUNT':<8}= {0:>20d}", f"{'GCOUNT':<8}= {1:>20d}",
                         f"{'TFIELDS':<8}= {0:>20d}", "END"]
                hdr = "".join(f"{x:<80}" for x in cards)
                f.write((hdr + " " * (-len(hdr) % 2880)).encode("ascii"))
                f.write(b"\x7f" * nbytes + b"\0" * (-nbytes % 2880))
                continue
            data, extname, datamax = h
            f.write(_hdu_bytes(data, first, extname=extname, datamax=datamax))
            first = False


def write_ser(path, frames, bit_depth, big_endian=False, color_id=0):
    """a SER file (io/ser.c): frames [n][rows][cols] of uint8 / uint16 samples in DISPLAY order (row 0 = top row, as the
    format stores them); the header's LittleEndian field is written the way Siril and the first SER programs use it
    (0 = little-endian data, 1 = big-endian, io/ser.h:71-80)"""
    import struct
    a = np.asarray(frames)
    n, h, w = a.shape
    hdr = b"LUCAM-RECORDER" + struct.pack("<7i", 0, color_id, 1 if big_endian else 0, w, h, bit_depth, n)
    hdr += b"obs".ljust(40) + b"instr".ljust(40) + b"scope".ljust(40) + struct.pack("<2q", 0, 0)
    assert len(hdr) == 178
    raw = a.astype(np.uint8) if bit_depth <= 8 else a.astype(">u2" if big_endian else "<u2")
    with open(path, "wb") as f:
        f.write(hdr)
        f.write(raw.tobytes())


def parse_header(path):
    """BITPIX, (w, h, c), BZERO, DATAMAX or None, data offset"""
    kv, off = {}, 0
    with open(path, "rb") as f:
        while True:
            blk = f.read(2880).decode("ascii")
            off += 2880
            cards = [blk[i:i + 80] for i in range(0, 2880, 80)]
            for cd in cards:
                if cd[8:10] == "= ":
                    kv[cd[:8].strip()] = cd[10:].split("/")[0].strip()
            if any(cd.startswith("END     ") for cd in cards):
                break
    g = lambda k, d=None: float(kv[k]) if k in kv else d
    return int(kv["BITPIX"]), (int(kv["NAXIS1"]), int(kv["NAXIS2"]), int(kv.get("NAXIS3", 1))), g("BZERO", 0.0), g("DATAMAX"), off


def read_block(paths, row0=0, height=None, shifty=None):
    """stack_read_block_data (median_and_mean.c:382-545) for one row block of every file:
    [frame][layer][height][width], zeros outside a shifted frame."""
    out = None
    for f, p in enumerate(paths):
        bitpix, (w, fh, c), bzero, datamax, off = parse_header(p)
        hh = fh - row0 if height is None else height
        if bitpix == 16:
            raw = np.fromfile(p, dtype=">i2", count=w * fh * c, offset=off).reshape(c, fh, w)
            img = (raw.astype(np.int32) + 32768).astype(np.uint16)   # TUSHORT with BZERO 32768 / convert_data_ushort SHORT_IMG
        elif bitpix == 8:
            img = np.fromfile(p, dtype=np.uint8, count=w * fh * c, offset=off).reshape(c, fh, w).astype(np.uint16)
        else:
            img = np.fromfile(p, dtype=">f4", count=w * fh * c, offset=off).reshape(c, fh, w).astype(np.float32)
        if out is None:
            out = np.zeros((len(paths), c, hh, w), dtype=img.dtype)
        sy = int(shifty[f]) if shifty is not None else 0
        lo, hi = max(sy - row0, 0), min(fh + sy - row0, hh)
        if hi <= lo:
            continue
        for l in range(c):
            reg = img[l, row0 + lo - sy: row0 + hi - sy, :].copy()
            if bitpix == -32:
                dm = datamax
                if dm is None and reg.size > 3:     # image_format_fits.c:997-1003
                    dm = 0.0
                    flat = reg.reshape(-1)
                    for i in range(0, reg.size, reg.size // 3):
                        x = float(flat[i])
                        dm = dm if dm > x else x
                if dm is not None and dm > 10.0:
                    reg = reg * INV_USHRT_MAX_SINGLE
            out[f, l, lo:hi, :] = reg
    return out
