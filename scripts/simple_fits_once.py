"""one pass of the simple stackers (C2 geometry, 32 frames) and of the FITS reader, for the ncu launch list"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import siril_b200 as sb
from siril_b200 import fitsread

N, H, W = 32, 4176, 6248
for dt in ("f32", "u16"):
    fr = torch.rand((N, 1, H, W), device="cuda")
    if dt == "u16":
        fr = (fr * 65535).to(torch.int32).to(torch.uint16)
    for fn in (sb.stack_summing_generic, sb.stack_addmax, sb.stack_addmin):
        args = sb.StackingArgs(method=fn, frames=fr, use_32bit_output=True)
        for _ in range(2):
            assert args.method(args) == 0
    del fr
d = "/dev/shm/sgpu_once"
os.makedirs(d, exist_ok=True)
ctx = sb.Context(0)
for bitpix, dt in ((-32, ">f4"), (16, ">i2")):
    cards = [f"{'SIMPLE':<8}= {'T':>20}", f"{'BITPIX':<8}= {bitpix:>20d}", f"{'NAXIS':<8}= {2:>20d}",
             f"{'NAXIS1':<8}= {W:>20d}", f"{'NAXIS2':<8}= {H:>20d}", f"{'DATAMAX':<8}= {1.0:>20.6E}", "END"]
    hdr = "".join(f"{x:<80}" for x in cards); hdr += " " * (-len(hdr) % 2880)
    body = np.random.default_rng(0).random((H, W), dtype=np.float32)
    body = (body.astype(dt) if bitpix == -32 else (body * 60000 - 32768).astype(dt)).tobytes()
    p = f"{d}/x{bitpix}.fit"
    with open(p, "wb") as fo:
        fo.write(hdr.encode()); fo.write(body); fo.write(b"\0" * (-len(body) % 2880))
    fitsread.read_block_data(ctx, [p, p])
    os.remove(p)
ctx.close()
