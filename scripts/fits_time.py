"""Times the raw FITS block reader on the GPU box: N frames of the C2 geometry (6248x4176 f32, and the
same as 16-bit) written to /dev/shm (page-cache resident: the file system is not what is measured),
read into a resident block through pread -> pinned staging -> H2D -> device decode."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import siril_b200 as sb  # noqa: E402
from siril_b200 import fitsread  # noqa: E402

N, H, W = 8, 4176, 6248
d = "/dev/shm/sgpu_fits_time"
os.makedirs(d, exist_ok=True)
ctx = sb.Context(0)
for kind, bitpix, dt in (("f32", -32, ">f4"), ("u16", 16, ">i2")):
    rng = np.random.default_rng(1)
    paths = []
    for f in range(N):
        p = f"{d}/{kind}_{f}.fit"
        cards = [f"{'SIMPLE':<8}= {'T':>20}", f"{'BITPIX':<8}= {bitpix:>20d}", f"{'NAXIS':<8}= {2:>20d}",
                 f"{'NAXIS1':<8}= {W:>20d}", f"{'NAXIS2':<8}= {H:>20d}"]
        cards += [f"{'BZERO':<8}= {32768.0:>20.1f}", f"{'BSCALE':<8}= {1.0:>20.1f}"] if bitpix == 16 else [f"{'DATAMAX':<8}= {1.0:>20.6E}"]
        hdr = "".join(f"{x:<80}" for x in cards + ["END"])
        hdr += " " * (-len(hdr) % 2880)
        a = rng.random((H, W), dtype=np.float32)
        body = (a.astype(dt) if bitpix == -32 else (a * 65535 - 32768).astype(dt)).tobytes()
        with open(p, "wb") as fo:
            fo.write(hdr.encode()); fo.write(body); fo.write(b"\0" * (-len(body) % 2880))
        paths.append(p)
    for source in ("fd", "mmap", "pinned"):
        fitsread.read_block_data(ctx, paths[:2], source=source)      # warm: staging buffers, page cache
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        blk = fitsread.read_block_data(ctx, paths, source=source)
        torch.cuda.synchronize()
        dt_s = time.perf_counter() - t0
        gb = N * H * W * abs(bitpix) / 8 / 1e9
        print(f"{kind} {source}: {N} frames {gb:.2f} GB in {dt_s * 1e3:.1f} ms = {gb / dt_s:.1f} GB/s "
              f"({N * H * W / dt_s / 1e9:.2f} G samples/s; 'pinned' includes the host read into pinned memory)")
        del blk
    # files on disk -> stacked image in one call (sgpu_stack_fits): reader, stack and read-back pipelined by row blocks
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=None, type_of_rejection=sb.SIGMA, sig=(3.0, 3.0), ctx=ctx)
    assert fitsread.stack_fits(args, paths) == 0, getattr(args, "error", "")
    t0 = time.perf_counter()
    assert fitsread.stack_fits(args, paths) == 0
    dt_s = time.perf_counter() - t0
    print(f"{kind} stack_fits (sigma 3/3): {N} files -> result in {dt_s * 1e3:.1f} ms = {N * H * W / dt_s / 1e9:.2f} G px-frames/s, "
          f"{N * H * W * abs(bitpix) / 8 / dt_s / 1e9:.1f} GB/s of file data")
    for p in paths:
        os.remove(p)
ctx.close()
