"""time sgpu_image_stats on one resident plane (C2 geometry) — used alone and under ncu"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import siril_b200 as sb
from siril_b200.stacking import image_stats

H, W = 4176, 6248
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
dt = sys.argv[2] if len(sys.argv) > 2 else "f32"
g = torch.Generator(device="cuda").manual_seed(1)
img = 0.1 + 0.01 * torch.randn((H, W), device="cuda", generator=g)
img += (torch.rand((H, W), device="cuda", generator=g) < 0.002) * 0.5
img = img.clamp_(1e-4, 1.0)
if dt == "u16":
    img = (img * 65535).round().to(torch.int32).to(torch.uint16)
ctx = sb.Context(0, torch.cuda.current_stream().cuda_stream)
for lite in (False, True):
    image_stats(ctx, img, lite)
    torch.cuda.synchronize()
    l0 = ctx.launches
    t0 = time.perf_counter()
    for _ in range(reps):
        st = image_stats(ctx, img, lite)
    torch.cuda.synchronize()
    dtm = (time.perf_counter() - t0) / reps * 1e3
    print(f"{dt} lite={lite}: {dtm:.3f} ms per plane, {(ctx.launches - l0) // reps} launches, misses {ctx.stats_window_misses}, median {st.median:.6f}")
