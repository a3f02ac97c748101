#!/usr/bin/env python3
"""bench.py — rejection-stacking hot path: input px-frames/s and fraction of the HBM roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload C2] [--impl reference]

One "step" = one pass of the stacking call (normalization apply + rejection + combine + output
write; for N > 1 also the NCCL gather of the output rows to rank 0) over one synthetic frame
block resident in HBM.  Workloads are BASELINE.json's configs; the default (N = 1) is configs[1]:
C2 = 100 x 6248x4176 mono float32, winsorized sigma clip 3/3, additive+scaling normalization.
For N > 1 the image is row-sharded, one shard of the workload's height per rank (weak scaling).

Prints ONE JSON line (rank 0).  `--impl reference` times the reference's own CPU implementation
(oracle/_ref built from /root/reference, else the oracle port) on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

NO_REJEC, PERCENTILE, SIGMA, MAD, SIGMEDIAN, WINSORIZED, LINEARFIT, GESDT = range(8)
NO_NORM, ADDITIVE, MULTIPLICATIVE, ADDITIVE_SCALING, MULTIPLICATIVE_SCALING = range(5)

# name: (frames, layers, height, width, dtype, is_mean, rejection, sig, normalize, seed, label)
WORKLOADS = {
    "C1": (20, 1, 1024, 1024, "f32", True, SIGMA, (3.0, 3.0), NO_NORM, 1001,
           "20x1024x1024 mono f32, sigma clip 3/3, no normalization"),
    "C2": (100, 1, 4176, 6248, "f32", True, WINSORIZED, (3.0, 3.0), ADDITIVE_SCALING, 1002,
           "100x6248x4176 mono f32 (IMX455), winsorized 3/3, additive+scaling normalization"),
    "C3": (200, 3, 4000, 6000, "f32", False, NO_REJEC, (3.0, 3.0), ADDITIVE_SCALING, 1003,
           "200x6000x4000 RGB f32, median, additive+scaling normalization (noise weights computed, not applied: reference ignores them for median)"),
    "C4": (300, 1, 6388, 9576, "f32", True, GESDT, (0.3, 0.05), NO_NORM, 1004,
           "300x9576x6388 mono f32 (IMX461), GESDT 0.3/0.05"),
    "C5": (500, 1, 6388, 9576, "u16", True, LINEARFIT, (5.0, 5.0), MULTIPLICATIVE, 1005,
           "500x9576x6388 mono u16, linear-fit clip 5/5, multiplicative normalization"),
    "target": (300, 1, 6388, 9576, "f32", True, SIGMA, (3.0, 3.0), NO_NORM, 1004,
               "300x9576x6388 mono f32, sigma clip 3/3 (north-star target)"),
}
REJ_NAMES = ["none", "percentile", "sigma", "mad", "sigmedian", "winsorized", "linearfit", "gesdt"]
METRIC = "input px-frames/s, rejection stacking"


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                pass
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                continue
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


def algorithmic_bytes(n, c, h, w, es, out_es=4, n_rejmaps=0):
    """SURVEY.md §8d / BASELINE.md §3: every input sample read once, every output pixel (and
    rejection-map pixel) written once"""
    return w * h * c * (n * es + out_es + 2 * n_rejmaps)


def cpu_reference_run(frames_np, wl, coeff, nthreads, budget_s=15.0):
    """Time the reference's CPU implementation on a bounded row sample of the same workload."""
    from oracle import oracle as O
    kind = "reference_fast" if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libsiril_ref_fast.so")) else "port"
    orc = O.Oracle(kind)
    n, c, h, w, dt, is_mean, rej, sig, nrm, seed, label = wl
    kw = dict(is_mean=is_mean, rejection=rej, sig=sig, normalize=nrm, nthreads=nthreads)
    if nrm != NO_NORM:
        kw.update(scale=coeff[2], offset=coeff[0], mul=coeff[1])
    rows_avail = frames_np.shape[2]
    # the reference's block planner needs at least a few rows per thread
    nthreads = max(1, min(nthreads, rows_avail // 4))
    probe = min(rows_avail, max(16, 4 * nthreads))
    t0 = time.perf_counter()
    orc.stack(frames_np[:, :, :probe, :], **kw)
    tp = max(time.perf_counter() - t0, 1e-4)
    rows = int(max(probe, min(rows_avail, budget_s / tp * probe)))
    t0 = time.perf_counter()
    orc.stack(frames_np[:, :, :rows, :], **kw)
    dt_s = time.perf_counter() - t0
    units = n * c * rows * w
    return {"value": units / dt_s, "unit": "px-frames/s", "cores": nthreads,
            "kind": "reference" if kind.startswith("reference") else "port",
            "sample": f"{n} frames x {rows} rows x {w} px x {c} layer(s) of the workload ({units:.3g} px-frames, {dt_s:.1f} s), frames in RAM"}, dt_s


def synth_numpy(n, c, h, w, dt, seed):
    """host-side synthetic rows for --impl reference when no GPU generator is available"""
    rng = np.random.default_rng(seed)
    sky = 0.10 * (1 + 0.02 * rng.uniform(-1, 1, (n, 1, 1, 1)))
    gain = 1 + 0.05 * rng.uniform(-1, 1, (n, 1, 1, 1))
    fr = gain * sky + 0.01 * rng.standard_normal((n, c, h, w))
    hot = rng.random((n, c, h, w)) < 1e-3
    fr[hot] += rng.uniform(0.2, 0.6, int(hot.sum()))
    fr = np.clip(fr, 1e-4, 1.0)
    return np.round(fr * 65535).astype(np.uint16) if dt == "u16" else fr.astype(np.float32)


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n, c, h, w, dt, is_mean, rej, sig, nrm, seed, label = wl
    nthreads = os.cpu_count() or 1
    rows = min(h, max(256, 8 * nthreads))
    frames = synth_numpy(n, c, rows, w, dt, seed)
    coeff = None
    if nrm != NO_NORM:
        rng = np.random.default_rng(1)
        coeff = (0.002 * rng.standard_normal((c, n)) * (65535 if dt == "u16" else 1),
                 1 + 0.02 * rng.standard_normal((c, n)), 1 + 0.02 * rng.standard_normal((c, n)))
    per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    vals, times = [], []
    for i in range(args.warmup + args.steps):
        cb, dt_s = cpu_reference_run(frames, wl, coeff, nthreads, budget_s=per_step)
        if i >= args.warmup:
            vals.append(cb["value"]); times.append(dt_s)
    v = float(np.mean(vals))
    cb["value"] = v
    line = {"metric": METRIC, "value": v, "unit": "px-frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": float(np.mean(times) * 1e3), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": dt, "data": "synthetic", "impl": "reference",
            "config": {"workload": label, "name": args.workload, "rejection": REJ_NAMES[rej] if is_mean else "median"},
            "cpu_baseline": cb, "e2e": {"value": v, "unit": "px-frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def run_ours(args, wl):
    import ctypes as C
    import torch
    import torch.distributed as dist
    import siril_b200 as sb
    from siril_b200 import _lib

    n, c, h, w, dt, is_mean, rej, sig, nrm, seed, label = wl
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the stacking path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"      # keep NCCL's version banner off stdout: one JSON line
        dist.init_process_group("nccl", device_id=dev)
    L = _lib.lib()
    stream = torch.cuda.current_stream(dev)
    ctx = sb.Context(local, stream.cuda_stream)
    tdt = torch.float32 if dt == "f32" else torch.uint16
    es = 4 if dt == "f32" else 2

    # ---- synthetic frame block for this rank's row shard (device-side generator) ----
    frames = torch.empty((n, c, h, w), dtype=tdt, device=dev)
    desc = _lib.BlockDesc(n, c, w, h, sb.DATA_FLOAT if dt == "f32" else sb.DATA_USHORT)
    blk = C.c_void_p()
    ctx.check(L.sgpu_block_wrap(ctx.h, C.byref(desc), C.c_void_p(frames.data_ptr()), C.byref(blk)))
    ctx.check(L.sgpu_synth_fill(ctx.h, blk, seed, rank * h, world * h))
    ctx.synchronize()

    # ---- normalization coefficients (K3/K4): frame-parallel estimators, all-gathered ----
    sargs = sb.StackingArgs(method=sb.stack_mean_with_rejection if is_mean else sb.stack_median, frames=frames,
                            type_of_rejection=rej, sig=sig, normalize=nrm, ctx=ctx,
                            weighting_type=sb.NOISE_WEIGHT if args.workload == "C3" else sb.NO_WEIGHT)
    stats_ms = None
    if nrm != NO_NORM:
        from siril_b200.stacking import compute_estimators, finalize_factors
        if world > 1:       # the first collective pays the communicator set-up: keep it out of stats_ms
            dist.all_reduce(torch.zeros(1, device=dev))
        mine = list(range(rank, n, world))
        from siril_b200.stacking import STATS_WORKERS
        compute_estimators(sargs, mine[:STATS_WORKERS])     # warm-up: worker contexts, scratch, kernel attributes
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        off, mul, sc, bg = compute_estimators(sargs, mine)
        if world > 1:
            from siril_b200 import sharding
            off, mul, sc, bg = sharding.merge_estimators(dist, [off, mul, sc, bg], mine, device=dev)
        finalize_factors(sargs, off, mul, sc)
        sargs.bgnoise = bg
        torch.cuda.synchronize()
        stats_ms = (time.perf_counter() - t0) * 1e3
    coeff_np = (sargs.coeff.offset, sargs.coeff.mul, sargs.coeff.scale) if nrm != NO_NORM else None

    # ---- device-resident step ----
    keep = []
    cargs = sb.stacking._fill_cargs(sargs, n, c, is_mean, keep)
    out = torch.empty((c, h, w), dtype=torch.float32, device=dev)
    irej = torch.zeros(6, dtype=torch.int64, device=dev)
    gathered = [torch.empty_like(out) for _ in range(world)] if (world > 1 and rank == 0) else None

    def step():
        ctx.check(L.sgpu_stack_block(ctx.h, blk, C.byref(cargs), C.c_void_p(out.data_ptr()), None, None, C.c_void_p(irej.data_ptr())))
        if world > 1:
            dist.gather(out, gathered, dst=0)

    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = ctx.launches
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    kev = []
    torch.cuda.synchronize()
    ev[0].record(stream)
    for i in range(args.steps):
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        k0.record(stream)
        ctx.check(L.sgpu_stack_block(ctx.h, blk, C.byref(cargs), C.c_void_p(out.data_ptr()), None, None, C.c_void_p(irej.data_ptr())))
        k1.record(stream)
        if world > 1:
            dist.gather(out, gathered, dst=0)
        ev[i + 1].record(stream)
        kev.append((k0, k1))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in kev]))
    launches = ctx.launches - launches0
    handed = ctx.handovers
    kernel_name = ctx.last_engine
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([total_ms, kernel_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, kernel_ms_max = float(t[0]), float(t[1])
    ms_per_step = total_ms / args.steps
    units = n * c * h * w * world
    value = units / (ms_per_step * 1e-3)

    # ---- end-to-end step through the host entry point (pinned host frames, H2D + D2H inside) ----
    e2e = None
    if not args.no_e2e:
        frame_bytes = c * h * w * es
        host = torch.empty((n, c, h, w), dtype=tdt).pin_memory()
        host.copy_(frames)
        torch.cuda.synchronize()
        hout = torch.empty((c, h, w), dtype=torch.float32).pin_memory()
        ptrs = (C.c_void_p * n)(*[host.data_ptr() + f * frame_bytes for f in range(n)])
        hirej = (C.c_uint64 * 6)()
        e2e_steps = max(1, min(args.steps, 3))

        def e2e_step():
            ctx.check(L.sgpu_stack_host(ctx.h, C.byref(desc), ptrs, None, C.byref(cargs), C.c_void_p(hout.data_ptr()), None, None, hirej))
        e2e_step()   # warm-up
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        torch.cuda.synchronize()
        dt_s = (time.perf_counter() - t0) / e2e_steps
        tt = torch.tensor([dt_s], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        e2e = {"value": units / float(tt[0]), "unit": "px-frames/s", "h2d_bytes_per_step": n * frame_bytes * world,
               "d2h_bytes_per_step": (c * h * w * 4 + 48) * world, "ms_per_step": float(tt[0]) * 1e3, "steps": e2e_steps,
               "path": "sgpu_stack_host: pinned host frames -> H2D -> stack -> D2H result"}
        del host

    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        B = algorithmic_bytes(n, c, h, w, es)
        achieved = B / (kernel_ms_max * 1e-3) / 1e9
        traffic = None    # DRAM bytes per launch from the committed ncu --set full capture (per px-frame x units)
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            t = json.load(open(tp)).get(args.workload)
            if t and t.get("kernel") == kernel_name:
                traffic = t["dram_bytes_per_px_frame"] * n * c * h * w
        cb = None
        if world == 1 or True:
            rows = min(h, max(256, 8 * (os.cpu_count() or 1)))
            sample = frames[:, :, :rows, :].contiguous()
            sample_np = (sample.view(torch.int16).cpu().numpy().view(np.uint16) if dt == "u16" else sample.cpu().numpy())
            cb, _ = cpu_reference_run(sample_np, wl, coeff_np, os.cpu_count() or 1, budget_s=args.cpu_budget)
        line = {
            "metric": METRIC, "value": value, "unit": "px-frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": dt, "data": "synthetic",
            "config": {"workload": label, "name": args.workload, "rejection": REJ_NAMES[rej] if is_mean else "median",
                       "frames": n, "width": w, "height_per_gpu": h, "layers": c,
                       "l2": "inputs larger than L2 (%.1f GB per GPU per step)" % (n * c * h * w * es / 1e9),
                       "sharding": "row shards, one per rank; output rows gathered to rank 0 with NCCL" if world > 1 else "single GPU",
                       "stats_ms": stats_ms,
                       "handed_to_exact_engine_px": handed},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src, "kernel": kernel_name, "kernel_ms": kernel_ms_max,
                         "algorithmic_bytes": B, "frac_of_8TBs_nominal": achieved / 8000.0},
            "cpu_baseline": cb, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
        }
        print(json.dumps(line))
    L.sgpu_block_destroy(blk)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C2", choices=sorted(WORKLOADS))
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--height", type=int, default=0, help="profiling only: override the workload's rows per GPU")
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU work for the cpu_baseline sample")
    ap.add_argument("--rej", default="", help="profiling only: override the workload's rejection (none, sigma, winsorized, ...)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    wl = WORKLOADS[args.workload]
    if args.height:
        wl = wl[:2] + (args.height,) + wl[3:-1] + (wl[-1] + f" [PROFILING SUBSET: {args.height} rows]",)
    if args.rej:
        wl = wl[:6] + (REJ_NAMES.index(args.rej),) + wl[7:-1] + (wl[-1] + f" [PROFILING: rejection {args.rej}]",)
    if args.impl == "reference":
        return run_reference(args, wl)
    return run_ours(args, wl)


if __name__ == "__main__":
    sys.exit(main())
