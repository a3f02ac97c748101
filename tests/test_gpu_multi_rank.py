"""Multi-GPU parity (needs >= 2 GPUs on the box; skipped otherwise): one process per GPU over NCCL,
every rank holding ONLY rows [H*g/G, H*(g+1)/G) of every frame.  Asserts what the strong-scaled
benchmark relies on: the normalization coefficients computed from the row shards (statistics
all-reduced, sgpu_image_stats_sharded) equal the single-GPU ones (counts / medians / MAD-derived
fields bit-equal, double sums to 1e-12), and the gathered output, both rejection maps and the
counters equal the single-GPU stack and the oracle."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _frames(dtype):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from test_gpu_parity import synth
    rng = np.random.default_rng(77)
    N, C, H, W = 24, 2, 90, 260
    fr = synth(rng, N, C, H, W, dtype, zeros=0.002, hot=0.004).astype(np.float64)
    # the signal depends on the row: a shard's own statistics differ from the frame's
    fr *= (1.0 + 0.5 * np.linspace(0, 1, H))[None, None, :, None]
    return (np.clip(fr, 0, 65535).astype(np.uint16) if dtype == np.uint16 else np.clip(fr, 0, 1).astype(np.float32))


def _worker(rank, world, port, q, dtype_name, normalize, rejection, sig):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    import siril_b200 as sb
    from siril_b200 import sharding as sh
    from siril_b200.stacking import compute_estimators, finalize_factors
    from test_gpu_parity import to_dev, to_np
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        fr = _frames(np.dtype(dtype_name).type)
        N, C, H, W = fr.shape
        r0, r1 = sh.row_shard(H, rank, world)
        mine = to_dev(np.ascontiguousarray(fr[:, :, r0:r1, :])).to(dev)
        args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=mine, type_of_rejection=rejection, sig=sig,
                               normalize=normalize, create_rejmaps=True)
        red = sh.make_reducer(dist, dev)
        off, mul, sc, bg = compute_estimators(args, shard=(red, r0, H))
        finalize_factors(args, off.copy(), mul.copy(), sc.copy())      # (works in place)
        assert args.method(args) == 0, getattr(args, "error", "")
        out = sh.gather_rows(dist, args.result.data)
        lo = sh.gather_rows(dist, args.result.rejmap_low)
        hi = sh.gather_rows(dist, args.result.rejmap_high)
        irej = sh.reduce_irej(dist, torch.from_numpy(args.result.irej.astype(np.int64)).to(dev)).cpu().numpy()
        # statistics the other way: one all-to-all turns the shards into whole frames, each GPU computes its frames'
        # statistics with the single-GPU kernels, the estimators are all-gathered (sharding.frames_from_shards)
        whole, idx = sh.frames_from_shards(dist, mine, rank, world)
        wargs = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=whole, normalize=normalize, ctx=args.ctx)
        part = [np.zeros((C, N)) for _ in range(4)]
        for arr, loc in zip(part, compute_estimators(wargs)):
            arr[:, idx] = loc
        est2 = sh.merge_estimators(dist, part, idx, dev)
        del whole, wargs
        # the same stack again, every rank's kernels storing its rows straight into rank 0's image over NVLink
        # (sgpu_ipc_export / sgpu_ipc_open + out_plane_stride): no gather
        import ctypes as Cc
        from siril_b200 import _lib
        L, ctx = _lib.lib(), args.ctx
        peer = sh.PeerImage(dist, ctx, C, H, W, 2)
        keep = []
        args.coeff = sb.NormCoeff(offset=args.coeff.offset, mul=args.coeff.mul, scale=args.coeff.scale)
        ca = sb.stacking._fill_cargs(args, N, C, True, keep)
        ca.out_plane_stride = H * W
        blk = Cc.c_void_p()
        desc = _lib.BlockDesc(N, C, W, r1 - r0, 1 if dtype_name == "float32" else 0)
        ctx.check(L.sgpu_block_wrap(ctx.h, Cc.byref(desc), Cc.c_void_p(mine.data_ptr()), Cc.byref(blk)))
        pirej = torch.zeros(6, dtype=torch.int64, device=dev)
        po, pl, ph = peer.pointers(r0)
        ctx.check(L.sgpu_stack_block(ctx.h, blk, Cc.byref(ca), Cc.c_void_p(po), Cc.c_void_p(pl), Cc.c_void_p(ph), Cc.c_void_p(pirej.data_ptr())))
        ctx.synchronize()
        dist.barrier()
        L.sgpu_block_destroy(blk)
        peer_res = None
        if rank == 0:
            pout, (plo, phi) = peer.tensors()
            peer_res = (to_np(pout.clone()), to_np(plo.clone()), to_np(phi.clone()))
        dist.all_reduce(pirej)
        peer.close()
        if rank == 0:
            q.put((rank, dict(off=off, mul=mul, sc=sc, bg=bg, coeff=(args.coeff.offset, args.coeff.mul, args.coeff.scale),
                              out=to_np(out), lo=to_np(lo), hi=to_np(hi), irej=irej, peer=peer_res, pirej=pirej.cpu().numpy(), est2=est2)))
        else:
            q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback
        q.put((rank, "ERR " + repr(e) + traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("dtype,normalize,rejection,sig", [
    (np.float32, 3, 2, (3.0, 3.0)),      # additive + scaling, sigma clip
    (np.uint16, 2, 5, (3.0, 3.0)),       # multiplicative, winsorized
], ids=["f32-addscale-sigma", "u16-mul-winsorized"])
def test_row_sharded_normalization_and_stack(orc, dtype, normalize, rejection, sig):
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs at least two GPUs")
    import torch.multiprocessing as mp
    import siril_b200 as sb
    from siril_b200.stacking import compute_estimators, finalize_factors
    from test_gpu_parity import to_dev, to_np
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q, np.dtype(dtype).name, normalize, rejection, sig)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=600) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    for r in range(1, world):
        assert res[r] == "ok", res[r]
    got = res[0]
    assert isinstance(got, dict), got
    # the same sequence on one GPU
    fr = _frames(dtype)
    args = sb.StackingArgs(method=sb.stack_mean_with_rejection, frames=to_dev(fr), type_of_rejection=rejection, sig=sig,
                           normalize=normalize, create_rejmaps=True)
    off, mul, sc, bg = compute_estimators(args)
    # estimators: medians / MAD-derived locations come from integer histograms and are bit-equal; IKSS scale
    # and the noise estimate are double sums
    for name, a, b in (("offset", got["off"], off), ("mul", got["mul"], mul)):
        assert np.array_equal(a, b), (name, np.abs(a - b).max())
    assert np.allclose(got["sc"], sc, rtol=1e-12, atol=0), np.abs(got["sc"] / sc - 1).max()
    assert np.allclose(got["bg"], bg, rtol=1e-12, atol=0), np.abs(got["bg"] / bg - 1).max()
    for a, b in zip(got["est2"], (off, mul, sc, bg)):     # frame-parallel after the all-to-all: one GPU's own numbers, bit for bit
        assert np.array_equal(a, b)
    finalize_factors(args, off.copy(), mul.copy(), sc.copy())
    for a, b in zip(got["coeff"], (args.coeff.offset, args.coeff.mul, args.coeff.scale)):
        assert np.allclose(a, b, rtol=1e-12, atol=1e-300)
    coeff = (args.coeff.offset.copy(), args.coeff.mul.copy(), args.coeff.scale.copy())
    assert args.method(args) == 0
    one = args.result
    # the sharded run used coefficients that may differ in the last bits: compare through the tolerance of the stack
    assert np.array_equal(got["irej"].reshape(3, 2), one.irej.astype(np.int64))
    assert np.array_equal(got["lo"], to_np(one.rejmap_low)) and np.array_equal(got["hi"], to_np(one.rejmap_high))
    a, b = got["out"], to_np(one.data)
    assert np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-30)) <= 1e-6
    # rows stored straight into rank 0's image by every rank's kernels: the same image, bit for bit, as the gathered one
    assert np.array_equal(got["peer"][0], got["out"], equal_nan=True)
    assert np.array_equal(got["peer"][1], got["lo"]) and np.array_equal(got["peer"][2], got["hi"])
    assert np.array_equal(got["pirej"], got["irej"].reshape(-1))
    ref = orc.stack(fr, rejection=rejection, sig=sig, normalize=normalize, offset=coeff[0], mul=coeff[1], scale=coeff[2],
                    create_rejmaps=True)
    assert np.array_equal(to_np(one.rejmap_low), ref["rej_low"]) and np.array_equal(one.irej, ref["irej"])
    assert np.max(np.abs(b - ref["out"]) / np.maximum(np.abs(ref["out"]), 1e-30)) <= 1e-6
