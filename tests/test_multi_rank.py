"""world_size-2 gloo test (CPU) of the row-shard plumbing bench.py uses on N GPUs
(siril_b200/sharding.py): shard bounds, estimator merge (frame-parallel -> all ranks), output-row
gather in image order, rejection-counter and min/max reductions.  The per-shard stacking itself is
the CPU oracle here (test infrastructure): stacking row shards separately and gathering must equal
stacking the whole image, which is the property the multi-GPU path relies on."""
import os
import socket
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from siril_b200 import sharding as sh
    from oracle import oracle as O
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(5)                      # same frames on every rank
        N, C, H, W = 12, 2, 11, 17
        fr = (0.1 + 0.01 * rng.standard_normal((N, C, H, W))).astype(np.float32)
        fr[rng.random(fr.shape) < 0.02] += 0.4
        orc = O.Oracle("port")
        # estimators: frame-parallel, merged
        mine = sh.my_frames(N, rank, world)
        loc = np.zeros((C, N)); sc = np.zeros((C, N))
        for f in mine:
            for c in range(C):
                st = orc.image_stats(fr[f, c])
                loc[c, f], sc[c, f] = st["location"], st["scale"]
        loc_all, sc_all = sh.merge_estimators(dist, [loc, sc], mine)
        full_loc = np.array([[orc.image_stats(fr[f, c])["location"] for f in range(N)] for c in range(C)])
        assert np.array_equal(loc_all, full_loc)
        assert np.all(sc_all > 0)
        # the all-reduce callback behind the statistics of row-sharded frames (sgpu_reduce_fn): every
        # element type and operator, and the property the sharded medians rest on — integer
        # histograms of the shards, summed, are the histogram of the frame
        import ctypes as ct
        r0, r1 = sh.row_shard(H, rank, world)
        red = sh.make_reducer(dist)
        fn = ct.cast(red, ct.CFUNCTYPE(ct.c_int, ct.c_void_p, ct.c_void_p, ct.c_size_t, ct.c_int, ct.c_int))
        plane = fr[3, 1]
        edges = np.linspace(plane.min(), plane.max() + 1e-6, 65)
        hist = np.histogram(plane[r0:r1], bins=edges)[0].astype(np.uint32)
        assert fn(None, hist.ctypes.data_as(ct.c_void_p), hist.size, 2, 0) == 0
        assert np.array_equal(hist, np.histogram(plane, bins=edges)[0].astype(np.uint32))
        cnt = np.array([np.count_nonzero(plane[r0:r1]), (r1 - r0) * W], dtype=np.uint64)
        assert fn(None, cnt.ctypes.data_as(ct.c_void_p), 2, 3, 0) == 0 and cnt.tolist() == [np.count_nonzero(plane), H * W]
        lo = np.array([plane[r0:r1].min()], dtype=np.float32); hi = np.array([plane[r0:r1].max()], dtype=np.float32)
        assert fn(None, lo.ctypes.data_as(ct.c_void_p), 1, 0, 1) == 0 and fn(None, hi.ctypes.data_as(ct.c_void_p), 1, 0, 2) == 0
        assert lo[0] == plane.min() and hi[0] == plane.max()
        sd = np.array([plane[r0:r1].astype(np.float64).sum()])
        assert fn(None, sd.ctypes.data_as(ct.c_void_p), 1, 1, 0) == 0 and abs(sd[0] - plane.astype(np.float64).sum()) <= 1e-12 * abs(sd[0])
        rows = np.zeros(H); rows[r0:r1] = np.arange(r0, r1) + 1.0       # the zero-padded sum that gathers per-row values
        assert fn(None, rows.ctypes.data_as(ct.c_void_p), H, 1, 0) == 0 and np.array_equal(rows, np.arange(H) + 1.0)
        # shard, stack, gather: every rank stacks ONLY its rows
        kw = dict(rejection=O.SIGMA, sig=(3.0, 3.0), create_rejmaps=True)
        part = orc.stack(np.ascontiguousarray(fr[:, :, r0:r1, :]), **kw)
        out = sh.gather_rows(dist, torch.from_numpy(part["out"]))
        irej = sh.reduce_irej(dist, torch.from_numpy(part["irej"].astype(np.int64)))
        mn, mx = sh.reduce_minmax(dist, float(part["out"][part["out"] != 0].min()), float(part["out"].max()))
        whole = orc.stack(fr, **kw)
        assert np.array_equal(irej.numpy(), whole["irej"].astype(np.int64))
        assert mn == float(whole["out"][whole["out"] != 0].min()) and mx == float(whole["out"].max())
        if rank == 0:
            assert out.shape == (C, H, W) and np.array_equal(out.numpy(), whole["out"])
        else:
            assert out is None
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, repr(e) + traceback.format_exc()[-600:]))
    finally:
        dist.destroy_process_group()


def test_row_shards_cover_image():
    sys.path.insert(0, ROOT)
    from siril_b200 import sharding as sh
    for H in (1, 7, 4176, 6388):
        for G in (1, 2, 3, 8):
            b = [sh.row_shard(H, g, G) for g in range(G)]
            assert b[0][0] == 0 and b[-1][1] == H and all(b[i][1] == b[i + 1][0] for i in range(G - 1))
    assert sh.my_frames(10, 1, 4) == [1, 5, 9]


def test_two_ranks_gloo():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, "ok"), (1, "ok")], res
