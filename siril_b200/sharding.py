"""Row-shard plumbing for one process per GPU (torch.distributed; NCCL on GPUs, gloo in the CPU
tests).  The stacking path shards by image rows — the reference's own block unit
(`struct _image_block`, src/stacking/stacking.h:167-169; OpenMP loop median_and_mean.c:1552) — so
no collective is needed while stacking; what is exchanged is
  * the reductions behind the per-frame normalization statistics (counts, sums, minima / maxima,
    integer histograms, per-row noise values): every rank holds only its rows of a frame and the
    estimators are whole-frame quantities, so each reduction is all-reduced before it is used
    (make_reducer + sgpu_image_stats_sharded; stacking.compute_estimators(..., shard=...)),
  * the output rows (and rejection maps), gathered to rank 0,
  * the rejection counters irej[3][2] (median_and_mean.c:1748-1757) and, for output_norm, the
    min/max of the non-zero result pixels (norm_to_0_1_range, :557-582)."""
import numpy as np


def row_shard(height, rank, world):
    """rows [r0, r1) of every frame and layer held by `rank`: H*g/G .. H*(g+1)/G"""
    return (height * rank) // world, (height * (rank + 1)) // world


def my_frames(nb_frames, rank, world):
    """frames whose statistics this rank computes (whole frames, frame-parallel)"""
    return list(range(rank, nb_frames, world))


def make_reducer(dist, device=None):
    """sgpu_reduce_fn over torch.distributed: an in-place all-reduce of a HOST buffer (NCCL: through
    a device copy; gloo: in place).  Unsigned counts travel as int64.  Keep the returned object alive
    for as long as the C side may call it."""
    import ctypes as C
    import torch
    from . import _lib
    ctypes_of = {0: C.c_float, 1: C.c_double, 2: C.c_uint32, 3: C.c_uint64}
    ops = {0: dist.ReduceOp.SUM, 1: dist.ReduceOp.MIN, 2: dist.ReduceOp.MAX}

    def cb(_user, buf, count, dtype, op):
        try:
            arr = np.ctypeslib.as_array(C.cast(buf, C.POINTER(ctypes_of[dtype])), shape=(count,))
            t = torch.from_numpy(arr.astype(np.int64) if dtype >= 2 else arr.copy())
            if device is not None:
                t = t.to(device)
            dist.all_reduce(t, op=ops[op])
            arr[:] = t.cpu().numpy().astype(arr.dtype)
            return 0
        except Exception as e:   # noqa: BLE001 — nothing may propagate through the C frames
            print(f"sgpu all-reduce callback failed: {e!r}", flush=True)
            return 1
    return _lib.REDUCE_FN(cb)


def merge_estimators(dist, arrays, mine, device=None):
    """all ranks end up with the full [C][N] estimator arrays; `arrays` hold valid entries only
    at frame indices `mine`.  Disjoint frame sets: a sum over zero-padded arrays is an all-gather."""
    import torch
    pack = torch.from_numpy(np.stack([np.asarray(a, dtype=np.float64) for a in arrays]))
    mask = torch.zeros_like(pack)
    mask[..., mine] = 1
    pack = pack * mask
    if device is not None:
        pack = pack.to(device)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(pack)
    return [x.cpu().numpy() for x in pack]


def gather_rows(dist, shard, dst=0):
    """gather the [C][rows][W] output shards to rank `dst` in image order -> [C][H][W] (or None)"""
    import torch
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return shard
    world, rank = dist.get_world_size(), dist.get_rank()
    u16 = shard.dtype == torch.uint16      # no 16-bit integer type in torch's NCCL binding: sent as byte views
    if u16:
        shard = shard.contiguous().view(torch.uint8)
    rows = torch.tensor([shard.shape[1]], dtype=torch.int64, device=shard.device)
    all_rows = [torch.zeros_like(rows) for _ in range(world)]
    dist.all_gather(all_rows, rows)
    counts = [int(r) for r in all_rows]
    most = max(counts)
    send = shard.contiguous()
    if send.shape[1] < most:      # dist.gather wants equal shapes: pad ragged shards, trim after
        pad = torch.zeros((shard.shape[0], most - send.shape[1], shard.shape[2]), dtype=shard.dtype, device=shard.device)
        send = torch.cat([send, pad], dim=1)
    parts = [torch.empty_like(send) for _ in range(world)] if rank == dst else None
    dist.gather(send, parts, dst=dst)
    if rank != dst:
        return None
    full = torch.cat([p[:, :c, :] for p, c in zip(parts, counts)], dim=1)
    return full.view(torch.uint16) if u16 else full


def reduce_irej(dist, irej):
    """sum of the per-shard rejection counters on every rank"""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(irej)
    return irej


def reduce_minmax(dist, mn, mx, device=None):
    """global min / max of the non-zero result pixels (output_norm)"""
    import torch
    t = torch.tensor([-float(mn), float(mx)], dtype=torch.float64, device=device)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return -float(t[0]), float(t[1])


class PeerImage:
    """The result image [C][H][W] float32 (or uint16) and its rejection maps, allocated once on rank `root` and mapped by
    every other rank (sgpu_ipc_export / sgpu_ipc_open, CUDA IPC): a rank hands sgpu_stack_block pointers to ITS rows
    inside the root's memory (`pointers(r0)`, `plane`), so the stacking kernels store over NVLink while they compute and
    no gather follows.  After the stack: ctx.synchronize() on every rank, then a barrier, then the root owns the image."""

    def __init__(self, dist, ctx, layers, height, width, n_maps, out32=True, root=0):
        import ctypes as C
        import torch
        from . import _lib
        self.L, self.ctx, self.dist, self.root = _lib.lib(), ctx, dist, root
        self.rank = dist.get_rank()
        self.shape, self.n_maps, self.out32 = (layers, height, width), n_maps, out32
        self.plane = height * width
        npx = layers * height * width
        al = lambda b: (b + 255) & ~255    # noqa: E731
        self.out_bytes, self.map_bytes = al(npx * (4 if out32 else 2)), al(npx * 2)
        self.base = C.c_void_p()
        handle = torch.zeros(64, dtype=torch.uint8)
        ok, why = 1, ""
        import os
        if os.environ.get("SGPU_NO_PEER") == str(self.rank):     # tests: make the mapping fail on one rank
            ok, why = 0, "SGPU_NO_PEER"
        if self.rank == root and ok:
            try:
                ctx.check(self.L.sgpu_device_alloc(ctx.h, self.out_bytes + n_maps * self.map_bytes, C.byref(self.base)))
                h = (C.c_ubyte * 64)()
                ctx.check(self.L.sgpu_ipc_export(ctx.h, self.base, h))
                handle = torch.tensor(list(h), dtype=torch.uint8)
                ctx.synchronize()
            except _lib.SgpuError as e:
                ok, why = 0, str(e)
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else None
        t = handle.to(dev) if dev is not None else handle
        dist.broadcast(t, src=root)
        if self.rank != root and ok:
            try:
                h = (C.c_ubyte * 64)(*t.cpu().tolist())
                ctx.check(self.L.sgpu_ipc_open(ctx.h, h, C.byref(self.base)))
            except _lib.SgpuError as e:
                ok, why = 0, str(e)
        # every rank learns whether every rank has the image: a mapping that fails on one GPU (no peer access, a
        # container without shared IPC) must not leave the others waiting
        flag = torch.tensor([ok], dtype=torch.int32)
        flag = flag.to(dev) if dev is not None else flag
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag.cpu()[0]) == 0:
            if self.base:
                if self.rank == root:
                    self.L.sgpu_device_free(ctx.h, self.base)
                elif ok:
                    self.L.sgpu_ipc_close(ctx.h, self.base)
            self.base = None
            raise RuntimeError("the root's image cannot be mapped on every rank" + (": " + why if why else ""))

    def pointers(self, row0):
        """device pointers to row `row0` of layer 0 of the image and of the maps (None where there is no map)"""
        b = self.base.value
        es = 4 if self.out32 else 2
        w = self.shape[2]
        maps = [b + self.out_bytes + k * self.map_bytes + row0 * w * 2 for k in range(self.n_maps)]
        return [b + row0 * w * es] + maps + [None] * (2 - self.n_maps)

    def tensors(self):
        """on the root: torch views of the image and the maps (no copy)"""
        import torch
        assert self.rank == self.root

        class _Arr:
            def __init__(self, ptr, shape, typestr):
                self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 2}
        b = self.base.value
        dev = torch.device("cuda", torch.cuda.current_device())
        out = torch.as_tensor(_Arr(b, self.shape, "<f4" if self.out32 else "<u2"), device=dev)
        maps = [torch.as_tensor(_Arr(b + self.out_bytes + k * self.map_bytes, self.shape, "<u2"), device=dev) for k in range(self.n_maps)]
        return out, maps

    def close(self):
        if self.base:
            if self.rank == self.root:
                self.dist.barrier()     # the peers unmap first
                self.L.sgpu_device_free(self.ctx.h, self.base)
            else:
                self.L.sgpu_ipc_close(self.ctx.h, self.base)
                self.dist.barrier()
            self.base = None


def frames_from_shards(dist, shard, rank=None, world=None):
    """Row shards -> whole frames, frame-parallel (first alternative of the statistics phase): every rank
    holds rows [H*g/G, H*(g+1)/G) of EVERY frame ([N][C][h_g][W], device resident) and ends up with ALL rows of the frames
    f = rank (mod G) ([N_g][C][H][W]): one all-to-all of pixel data over NVLink, (G-1)/G of the shard out and as much in.
    The per-frame statistics are whole-frame quantities (normalization.c:217-294): after this each frame's statistics
    are computed by one GPU with the single-GPU kernels — bit-identical to one GPU — and only the estimators travel
    (merge_estimators).  Returns (frames [N_g][C][H][W], indices of those frames)."""
    import torch
    world = dist.get_world_size() if world is None else world
    rank = dist.get_rank() if rank is None else rank
    n, c, h, w = shard.shape
    u16 = shard.dtype == torch.uint16
    if u16:     # (no 16-bit integer type in torch's NCCL binding: the samples travel as byte views, 2 W bytes per row)
        shard = shard.contiguous().view(torch.uint8)
        w = 2 * w
    mine = list(range(rank, n, world))
    rows = torch.tensor([h], dtype=torch.int64, device=shard.device)
    all_rows = [torch.zeros_like(rows) for _ in range(world)]
    dist.all_gather(all_rows, rows)
    heights = [int(r) for r in all_rows]
    send = [shard[r::world].contiguous() for r in range(world)]                       # my rows of rank r's frames
    recv = [torch.empty((len(mine), c, heights[r], w), dtype=shard.dtype, device=shard.device) for r in range(world)]
    dist.all_to_all(recv, send)
    del send
    whole = torch.cat(recv, dim=2)      # rank order is row order
    del recv
    return (whole.view(torch.uint16) if u16 else whole), mine
