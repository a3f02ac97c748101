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
