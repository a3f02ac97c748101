#!/usr/bin/env python3
"""bench.py — rejection-stacking hot path: input px-frames/s and fraction of the HBM roofline.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload target] [--impl reference]

One "step" = one pass of the stacking call (normalization apply + rejection + combine + output and
rejection-map writes; for N > 1 every rank's kernels store their rows straight into the image on
rank 0 over NVLink — or, --gather nccl, the rows are sent there after the stack — and the rejection
counters are all-reduced) over one synthetic frame block resident in HBM.  Workloads are BASELINE.json's configs plus the north-star target (sigma clip of 300 x 61 MP
float frames), which is the default.  For N > 1 ONE image is row-sharded over the ranks (strong
scaling): rank g holds rows [H*g/N, H*(g+1)/N) of every frame.

Prints ONE JSON line (rank 0).  `--impl reference` times the reference's own CPU implementation
(oracle/_ref built from /root/reference, else the oracle port) on the host cores, on a bounded row
sample of the same workload, same generator, same coefficients.
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
N_REJMAPS = 2       # both rejection maps are part of the output (median_and_mean.c:1333-1342)


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
        self.index, self.rows, self.proc, self.i0 = index, [], None, 0

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "25"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def mark(self):
        """the timed region starts here: only what is sampled from now on counts"""
        self.i0 = len(self.rows)

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
        for r in (self.rows[self.i0:] or self.rows[-1:]):   # (a region shorter than one sampling period: the sample before it)
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


def config_of(args, wl, world, extra=None):
    """the `config` object: identical keys (and, for the same workload, values) in both arms"""
    n, c, h, w, dt, is_mean, rej, sig, nrm, seed, label = wl
    cfg = {"workload": label, "name": args.workload, "rejection": REJ_NAMES[rej] if is_mean else "median",
           "frames": n, "width": w, "height": h, "layers": c, "rejection_maps": N_REJMAPS if is_mean and rej != NO_REJEC else 0,
           "l2": "inputs larger than L2 (%.1f GB per step)" % (n * c * h * w * (4 if dt == "f32" else 2) / 1e9),
           "sharding": (("rows [H*g/N, H*(g+1)/N) of one image per rank; every rank's kernels store its output rows and both rejection "
                         "maps straight into the image on rank 0 over NVLink (peer mapping), counters all-reduced")
                        if getattr(args, "gather", "peer") == "peer" else
                        ("rows [H*g/N, H*(g+1)/N) of one image per rank; output rows and both rejection maps gathered to rank 0 "
                         "with NCCL, counters all-reduced")) if world > 1 else "single GPU"}
    if extra:
        cfg.update(extra)
    return cfg


def cpu_reference_run(frames_np, wl, coeff, nthreads, budget_s=15.0):
    """Time the reference's CPU implementation on a bounded row sample of the same workload."""
    from oracle import oracle as O
    kind = "reference_fast" if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libsiril_ref_fast.so")) else "port"
    orc = O.Oracle(kind)
    n, c, h, w, dt, is_mean, rej, sig, nrm, seed, label = wl
    kw = dict(is_mean=is_mean, rejection=rej, sig=sig, normalize=nrm, nthreads=nthreads,
              create_rejmaps=bool(is_mean and rej != NO_REJEC))
    if nrm != NO_NORM:
        kw.update(scale=coeff[2], offset=coeff[0], mul=coeff[1])
    rows_avail = frames_np.shape[2]
    # the reference's block planner needs at least a few rows per thread
    nthreads = max(1, min(nthreads, rows_avail // 4))
    kw["nthreads"] = nthreads
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


def synth_numpy(n, c, h, w, dt, seed, full_h):
    """host twin of sgpu_synth_fill's model (misc.cu: per-frame sky / gain, a smooth signal, Gaussian
    noise 0.01, hot and cold outliers) for --impl reference on a box without a GPU; not bit-equal"""
    rng = np.random.default_rng(seed)
    sky = 0.10 * (1 + 0.02 * rng.uniform(-1, 1, (n, c, 1, 1)))
    gain = 1 + 0.05 * rng.uniform(-1, 1, (n, c, 1, 1))
    yy, xx = np.mgrid[0:h, 0:w]
    fx, fy = xx / w - 0.5, (yy + full_h // 2) / full_h - 0.5
    sig = np.abs(0.3 * np.exp(-(fx * fx + fy * fy) * 8) * (0.6 + 0.4 * np.sin(12 * fx) * np.cos(9 * fy)))
    fr = gain * (sky + sig[None, None]) + 0.01 * rng.standard_normal((n, c, h, w))
    u = rng.random((n, c, h, w))
    hot = u < 1e-3
    fr[hot] += 0.2 + 0.4 * rng.random(int(hot.sum()))
    fr[(u >= 1e-3) & (u < 1.2e-3)] -= 0.05
    fr = np.clip(fr, 1e-4, 1.0)
    return np.maximum(np.rint(fr * 65535), 1).astype(np.uint16) if dt == "u16" else fr.astype(np.float32)


def sample_rows(wl, rows):
    """`rows` rows from the middle of the workload's frames, as the GPU arm stacks them: generated by
    sgpu_synth_fill and downloaded when a GPU is there, else by the host twin.  Returns (frames,
    coefficients (offset, mul, scale) or None, how)"""
    n, c, h, w, dt, is_mean, rej, sig, nrm, seed, label = wl
    row0 = max(0, (h - rows) // 2)
    try:
        import ctypes as C
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("no GPU")
        import siril_b200 as sb
        from siril_b200 import _lib
        L = _lib.lib()
        ctx = sb.Context(0, torch.cuda.current_stream().cuda_stream)
        tdt = torch.float32 if dt == "f32" else torch.uint16
        fr = torch.empty((n, c, rows, w), dtype=tdt, device="cuda")
        desc = _lib.BlockDesc(n, c, w, rows, sb.DATA_FLOAT if dt == "f32" else sb.DATA_USHORT)
        blk = C.c_void_p()
        ctx.check(L.sgpu_block_wrap(ctx.h, C.byref(desc), C.c_void_p(fr.data_ptr()), C.byref(blk)))
        ctx.check(L.sgpu_synth_fill(ctx.h, blk, seed, row0, h))
        ctx.synchronize()
        L.sgpu_block_destroy(blk)
        coeff = None
        if nrm != NO_NORM:      # the statistics of the strip stand in for the frame's (same code path, a fraction of the rows)
            sargs = sb.StackingArgs(frames=fr, normalize=nrm, ctx=ctx)
            sb.do_normalization(sargs)
            coeff = (sargs.coeff.offset, sargs.coeff.mul, sargs.coeff.scale)
        host = fr.view(torch.int16).cpu().numpy().view(np.uint16) if dt == "u16" else fr.cpu().numpy()
        del fr
        torch.cuda.empty_cache()
        return host, coeff, "sgpu_synth_fill (the GPU arm's generator), downloaded"
    except Exception as e:  # noqa: BLE001 — no GPU or no library: the host twin
        frames = synth_numpy(n, c, rows, w, dt, seed, h)
        coeff = None
        if nrm != NO_NORM:
            rng = np.random.default_rng(1)
            coeff = (0.002 * rng.standard_normal((c, n)) * (65535 if dt == "u16" else 1),
                     1 + 0.02 * rng.standard_normal((c, n)), 1 + 0.02 * rng.standard_normal((c, n)))
        return frames, coeff, f"host twin of the generator ({type(e).__name__}: no GPU generator here)"


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n, c, h, w, dt, is_mean, rej, sig, nrm, seed, label = wl
    nthreads = os.cpu_count() or 1
    rows = min(h, 640)      # enough rows for a few seconds of CPU work at any depth
    frames, coeff, how = sample_rows(wl, rows)
    per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))      # at least 2 s of CPU work per step
    vals, times = [], []
    for i in range(args.warmup + args.steps):
        cb, dt_s = cpu_reference_run(frames, wl, coeff, nthreads, budget_s=per_step)
        if i >= args.warmup:
            vals.append(cb["value"]); times.append(dt_s)
    v = float(np.mean(vals))
    cb["value"] = v
    cb["sample"] += "; " + how
    world = int(os.environ.get("WORLD_SIZE", "1"))
    line = {"metric": METRIC, "value": v, "unit": "px-frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": float(np.mean(times) * 1e3), "higher_is_better": True, "scaling": "strong" if world > 1 else "weak",
            "vs_baseline": None, "dtype": dt, "data": "synthetic", "impl": "reference",
            "config": config_of(args, wl, world),
            "cpu_baseline": cb, "e2e": {"value": v, "unit": "px-frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


def mem_available_bytes():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) * 1024
    except Exception:
        pass
    return 0


def run_ours(args, wl):
    import ctypes as C
    import torch
    import torch.distributed as dist
    import siril_b200 as sb
    from siril_b200 import _lib, sharding

    n, c, H, w, dt, is_mean, rej, sig, nrm, seed, label = wl
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the stacking path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    json_fd = 1
    if world > 1:
        # NCCL prints its version banner on stdout when the first communicator comes up: keep the descriptor for the ONE
        # JSON line aside and point fd 1 at stderr for everything else
        sys.stdout.flush()
        json_fd = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)
    L = _lib.lib()
    stream = torch.cuda.current_stream(dev)
    ctx = sb.Context(local, stream.cuda_stream)
    tdt = torch.float32 if dt == "f32" else torch.uint16
    es = 4 if dt == "f32" else 2
    r0, r1 = sharding.row_shard(H, rank, world)
    h = r1 - r0                                     # this rank's rows of the ONE image (strong scaling)
    maps = N_REJMAPS if (is_mean and rej != NO_REJEC) else 0

    # ---- synthetic frame block: this rank's rows of every frame (device-side generator) ----
    frames = torch.empty((n, c, h, w), dtype=tdt, device=dev)
    desc = _lib.BlockDesc(n, c, w, h, sb.DATA_FLOAT if dt == "f32" else sb.DATA_USHORT)
    blk = C.c_void_p()
    ctx.check(L.sgpu_block_wrap(ctx.h, C.byref(desc), C.c_void_p(frames.data_ptr()), C.byref(blk)))
    ctx.check(L.sgpu_synth_fill(ctx.h, blk, seed, r0, H))
    ctx.synchronize()

    # ---- normalization coefficients (K3/K4) from whole-frame statistics ----
    sargs = sb.StackingArgs(method=sb.stack_mean_with_rejection if is_mean else sb.stack_median, frames=frames,
                            type_of_rejection=rej, sig=sig, normalize=nrm, ctx=ctx, create_rejmaps=maps > 0,
                            weighting_type=sb.NOISE_WEIGHT if args.workload == "C3" else sb.NO_WEIGHT)
    stats_ms = None
    if nrm != NO_NORM:
        from siril_b200.stacking import compute_estimators, finalize_factors, STATS_WORKERS
        if world > 1 and args.stats == "reduce":
            # every rank holds rows of every frame: the reductions behind the statistics are all-reduced, plane by plane
            dist.all_reduce(torch.zeros(1, device=dev))     # the first collective pays the communicator set-up
            reducer = sharding.make_reducer(dist, dev)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            off, mul, sc, bg = compute_estimators(sargs, shard=(reducer, r0, H))
        elif world > 1:
            # frame-parallel: one all-to-all turns the row shards into whole frames (frames f = rank mod N here), each
            # GPU computes its frames' statistics with the single-GPU kernels, the estimators are all-gathered
            dist.all_reduce(torch.zeros(1, device=dev))
            whole, mine = sharding.frames_from_shards(dist, frames, rank, world)       # warm-up: communicator, allocator
            wargs = sb.StackingArgs(method=sargs.method, frames=whole, normalize=nrm, ctx=ctx)
            compute_estimators(wargs, list(range(min(len(mine), STATS_WORKERS))))
            del whole, wargs
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            whole, mine = sharding.frames_from_shards(dist, frames, rank, world)
            wargs = sb.StackingArgs(method=sargs.method, frames=whole, normalize=nrm, ctx=ctx)
            lo, lm, ls, lb = compute_estimators(wargs)
            part = [np.zeros((c, n)) for _ in range(4)]
            for a, l in zip(part, (lo, lm, ls, lb)):
                a[:, mine] = l
            off, mul, sc, bg = sharding.merge_estimators(dist, part, mine, dev)
            del whole, wargs
            torch.cuda.empty_cache()
        else:
            compute_estimators(sargs, list(range(min(n, STATS_WORKERS))))     # warm-up: worker contexts, scratch, kernel attributes
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            off, mul, sc, bg = compute_estimators(sargs)
        finalize_factors(sargs, off, mul, sc)
        sargs.bgnoise = bg
        torch.cuda.synchronize()
        stats_ms = (time.perf_counter() - t0) * 1e3
    coeff_np = (sargs.coeff.offset, sargs.coeff.mul, sargs.coeff.scale) if nrm != NO_NORM else None

    # ---- device-resident step ----
    keep = []
    cargs = sb.stacking._fill_cargs(sargs, n, c, is_mean, keep)
    out = torch.empty((c, h, w), dtype=torch.float32, device=dev)
    rlo = torch.zeros((c, h, w), dtype=torch.uint16, device=dev) if maps else None
    rhi = torch.zeros((c, h, w), dtype=torch.uint16, device=dev) if maps else None
    irej = torch.zeros(6, dtype=torch.int64, device=dev)
    full = None
    peer = None
    if world > 1 and args.gather == "peer":
        # the whole image lives on rank 0 and every rank's kernels store their rows straight into it over NVLink
        # (CUDA IPC mapping, sharding.PeerImage): no gather step; what remains per step is the counters' all-reduce
        try:
            peer = sharding.PeerImage(dist, ctx, c, H, w, 2 if maps else 0)
            cargs.out_plane_stride = H * w
            p_out, p_lo, p_hi = peer.pointers(r0)
        except RuntimeError as e:       # (raised on every rank alike) no peer mapping on this box: gather with NCCL instead
            print(f"bench.py: {e}; falling back to --gather nccl", file=sys.stderr)
            args.gather = "nccl"
    if world > 1 and peer is None and rank == 0:     # the whole image on rank 0: [C][H][W] result and both maps
        full = [torch.empty((c, H, w), dtype=torch.float32, device=dev)] + \
               ([torch.empty((c, H, w), dtype=torch.uint16, device=dev) for _ in range(2)] if maps else [])
    bounds = [sharding.row_shard(H, g, world) for g in range(world)]

    def stack():
        irej.zero_()
        if peer is not None:
            ctx.check(L.sgpu_stack_block(ctx.h, blk, C.byref(cargs), C.c_void_p(p_out), C.c_void_p(p_lo) if maps else None,
                                         C.c_void_p(p_hi) if maps else None, C.c_void_p(irej.data_ptr())))
            return
        ctx.check(L.sgpu_stack_block(ctx.h, blk, C.byref(cargs), C.c_void_p(out.data_ptr()),
                                     C.c_void_p(rlo.data_ptr()) if maps else None, C.c_void_p(rhi.data_ptr()) if maps else None,
                                     C.c_void_p(irej.data_ptr())))

    def gather():
        """output rows and rejection maps of every rank into the image on rank 0; counters summed"""
        if peer is not None:        # the rows are already in rank 0's image (or on their way: the timed region ends with
            dist.all_reduce(irej)   # a device synchronize on every rank and a barrier)
            return
        # (torch's NCCL binding has no 16-bit integer type: the maps travel as byte views of the same memory)
        i16 = lambda t: t.view(torch.uint8) if t.dtype == torch.uint16 else t    # noqa: E731
        parts = [i16(t) for t in [out] + ([rlo, rhi] if maps else [])]
        ops = []
        if rank == 0:
            for t_full, t_mine in zip([i16(t) for t in full], parts):
                t_full[:, r0:r1, :].copy_(t_mine)
            recv = []
            for g in range(1, world):
                a, b = bounds[g]
                for t_full in [i16(t) for t in full]:
                    view = t_full[:, a:b, :]        # mono: the rows of a rank are one contiguous piece of the image
                    buf = view if view.is_contiguous() else torch.empty(view.shape, dtype=t_full.dtype, device=dev)
                    if buf is not view:
                        recv.append((view, buf))
                    ops.append(dist.P2POp(dist.irecv, buf, g))
            for q in dist.batch_isend_irecv(ops):
                q.wait()
            for view, buf in recv:
                view.copy_(buf)
        else:
            for t_mine in parts:
                ops.append(dist.P2POp(dist.isend, t_mine, 0))
            for q in dist.batch_isend_irecv(ops):
                q.wait()
        dist.all_reduce(irej)

    def step():
        stack()
        if world > 1:
            gather()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()     # (nvidia-smi takes a moment to come up: started before the warm-up, counted from the timed region on)
    for _ in range(args.warmup):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    launches0 = ctx.launches
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    kev = []
    torch.cuda.synchronize()
    sampler.mark()
    ev[0].record(stream)
    for i in range(args.steps):
        k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        k0.record(stream)
        stack()
        k1.record(stream)
        if world > 1:
            gather()
        ev[i + 1].record(stream)
        kev.append((k0, k1))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    total_ms = ev[0].elapsed_time(ev[-1])
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in kev]))
    launches = ctx.launches - launches0
    handed = torch.tensor([ctx.handovers], dtype=torch.int64, device=dev)
    kernel_name = ctx.last_engine
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([total_ms, kernel_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(handed)
    total_ms, kernel_ms_max = float(t[0]), float(t[1])
    ms_per_step = total_ms / args.steps
    units = n * c * H * w                           # the whole job: every rank's rows
    value = units / (ms_per_step * 1e-3)

    # ---- end-to-end step through the host entry point (pinned host frames, H2D + D2H inside) ----
    e2e = None
    if not args.no_e2e:
        frame_bytes = c * h * w * es
        need = n * frame_bytes
        avail = mem_available_bytes()
        e_rows = h
        if need > 0.6 * avail or need > (100 << 30):     # host RAM decides how much of the shard the host arm can hold pinned
            e_rows = max(16, int(h * min(0.6 * avail, 100 << 30) / need) // 4 * 4)
        e_rows = min(e_rows, h)
        eframe = c * e_rows * w * es
        host = torch.empty((n, c, e_rows, w), dtype=tdt).pin_memory()
        host.copy_(frames[:, :, :e_rows, :])
        torch.cuda.synchronize()
        edesc = _lib.BlockDesc(n, c, w, e_rows, sb.DATA_FLOAT if dt == "f32" else sb.DATA_USHORT)
        hout = torch.empty((c, e_rows, w), dtype=torch.float32).pin_memory()
        hlo = torch.empty((c, e_rows, w), dtype=torch.uint16).pin_memory() if maps else None
        hhi = torch.empty((c, e_rows, w), dtype=torch.uint16).pin_memory() if maps else None
        ptrs = (C.c_void_p * n)(*[host.data_ptr() + f * eframe for f in range(n)])
        hirej = (C.c_uint64 * 6)()
        e2e_steps = max(1, min(args.steps, 2))
        del frames      # the resident block makes room for the host arm's row blocks
        L.sgpu_block_destroy(blk)
        blk = None
        torch.cuda.empty_cache()

        def e2e_step():
            ctx.check(L.sgpu_stack_host(ctx.h, C.byref(edesc), ptrs, None, C.byref(cargs), C.c_void_p(hout.data_ptr()),
                                        C.c_void_p(hlo.data_ptr()) if maps else None, C.c_void_p(hhi.data_ptr()) if maps else None, hirej))
        e2e_step()   # warm-up
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        torch.cuda.synchronize()
        dt_s = (time.perf_counter() - t0) / e2e_steps
        e_units = torch.tensor([float(n * c * e_rows * w)], dtype=torch.float64, device=dev)
        tt = torch.tensor([dt_s], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dist.all_reduce(e_units)
        e2e = {"value": float(e_units[0]) / float(tt[0]), "unit": "px-frames/s",
               "h2d_bytes_per_step": int(float(e_units[0]) * es),
               "d2h_bytes_per_step": int(float(e_units[0]) / n * (4 + 2 * maps) + 48 * world),
               "ms_per_step": float(tt[0]) * 1e3, "steps": e2e_steps,
               "rows_per_rank": e_rows, "rows_of_shard": h,
               "path": "sgpu_stack_host: pinned host frames -> H2D in row blocks (three rotating device buffers) -> stack -> D2H result + rejection maps"
                       + ("" if e_rows == h else f"; host RAM holds {e_rows} of the shard's {h} rows")}
        del host

    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        B = algorithmic_bytes(n, c, h, w, es, n_rejmaps=maps)   # rank 0's launch: its rows
        achieved = B / (kernel_ms_max * 1e-3) / 1e9
        traffic = None    # DRAM bytes per launch from the committed ncu --set full capture (per px-frame x units)
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            tj = json.load(open(tp)).get(args.workload)
            if tj and str(tj.get("kernel", "")).split(" + ")[0] == kernel_name:     # ("stack_psel_kernel + stack_psel_finish_kernel": the pair)
                traffic = tj["dram_bytes_per_px_frame"] * n * c * h * w
        cb = None
        if not args.no_cpu:
            rows = min(H, 640)
            sample_np, _, how = sample_rows(wl, rows)
            cb, _ = cpu_reference_run(sample_np, wl, coeff_np, os.cpu_count() or 1, budget_s=args.cpu_budget)
            cb["sample"] += "; " + how
        line = {
            "metric": METRIC, "value": value, "unit": "px-frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if world > 1 else "weak", "vs_baseline": None,
            "dtype": dt, "data": "synthetic",
            "config": config_of(args, wl, world),
            "detail": {"rows_on_rank0": h, "stats_ms": stats_ms, "handed_to_exact_engine_px": int(handed[0]),
                       # the second headline: normalization statistics + coefficients + stack, for workloads that normalize
                       "normalize_plus_stack_ms": (stats_ms + ms_per_step) if stats_ms is not None else None,
                       "px_frames_per_s_with_normalization": (units / ((stats_ms + ms_per_step) * 1e-3)) if stats_ms is not None else None},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src, "kernel": kernel_name, "kernel_ms": kernel_ms_max,
                         "algorithmic_bytes": B, "frac_of_8TBs_nominal": achieved / 8000.0,
                         "note": ("per GPU: the stack call of rank 0's rows (slowest rank's time)" if world > 1 else "the stack call")
                                 + ("; stack_psel_kernel (streaming passes, leaves a record per pixel) + stack_psel_finish_kernel (sort + rejection, "
                                    "one thread per pixel) + the exact engine on the hand-over list: `achieved` is the algorithmic bytes over the time of "
                                    "all three, `traffic` the DRAM bytes of the first two (the records' round trip is on top of the samples)"
                                    if kernel_name == "stack_psel_kernel" else "")},
            "cpu_baseline": cb, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
        }
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if blk is not None:
        L.sgpu_block_destroy(blk)
    if peer is not None:
        torch.cuda.synchronize()
        peer.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="target", choices=sorted(WORKLOADS))
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"],
                    help="N > 1: how the rows reach rank 0 — stored into its image by the stacking kernels over NVLink (peer), "
                         "or sent after the stack with NCCL send / recv (nccl)")
    ap.add_argument("--stats", default="frames", choices=["frames", "reduce"],
                    help="N > 1, workloads with normalization: statistics frame-parallel after an all-to-all of the shards "
                         "(frames) or on the row shards with every reduction all-reduced (reduce)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--height", type=int, default=0, help="profiling only: override the workload's image height")
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU work for the cpu_baseline sample")
    ap.add_argument("--rej", default="", help="profiling only: override the workload's rejection (none, sigma, winsorized, ...)")
    ap.add_argument("--frames", type=int, default=0, help="profiling only: override the workload's stack depth")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    wl = WORKLOADS[args.workload]
    if args.height:
        wl = wl[:2] + (args.height,) + wl[3:-1] + (wl[-1] + f" [PROFILING SUBSET: {args.height} rows]",)
    if args.frames:
        wl = (args.frames,) + wl[1:-1] + (wl[-1] + f" [PROFILING: {args.frames} frames]",)
    if args.rej:
        wl = wl[:6] + (REJ_NAMES.index(args.rej),) + wl[7:-1] + (wl[-1] + f" [PROFILING: rejection {args.rej}]",)
    if args.impl == "reference":
        return run_reference(args, wl)
    return run_ours(args, wl)


if __name__ == "__main__":
    sys.exit(main())
