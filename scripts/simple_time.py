"""time the sum / max / min stackers on a resident block (C2 geometry by default)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import siril_b200 as sb

N, H, W = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (100, 4176, 6248)
dt = sys.argv[4] if len(sys.argv) > 4 else "f32"
fr = torch.rand((N, 1, H, W), device="cuda")
if dt == "u16":
    fr = (fr * 65535).to(torch.int32).to(torch.uint16)
shift = np.random.default_rng(1).integers(-9, 10, N).astype(np.int32)
for name, fn in (("sum", sb.stack_summing_generic), ("max", sb.stack_addmax), ("min", sb.stack_addmin)):
    for sx in (None, shift):
        args = sb.StackingArgs(method=fn, frames=fr, shiftx=sx, use_32bit_output=True)
        for _ in range(2):
            assert args.method(args) == 0
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            args.method(args)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        gb = fr.numel() * fr.element_size() / 1e9
        print(f"{name} {dt} N={N} {'shifted' if sx is not None else 'aligned'}: {ms:.3f} ms, {gb / ms * 1e3:.0f} GB/s of input")
