"""world_size-2 gloo tests (CPU) of the position-sharded index plumbing (deepreadmapper_b200/sharded.py): shard bounds,
label globalisation, the all-gather layout the merge kernel consumes, and the read split of the rerank stage.
The merge itself is a CUDA kernel (tests/test_gpu_parity.py::test_merge_topk); here its expectation is stated in numpy."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _local_lists(rank, n, k, lo, hi):
    """A rank's ascending (D, I) lists with shard-local labels, some rows partly unfilled."""
    rng = np.random.default_rng(100 + rank)
    D = np.sort(rng.integers(0, 40, (n, k)).astype(np.float32), axis=1)
    I = np.stack([rng.permutation(hi - lo)[:k] for _ in range(n)]).astype(np.int64)
    for r in range(n):
        o = np.lexsort((I[r], D[r]))
        D[r], I[r] = D[r][o], I[r][o]
    D[rank, k // 2:] = np.finfo(np.float32).max
    I[rank, k // 2:] = -1
    return D, I


def _worker(rank, world, port, n, k, n_windows, q):
    import torch
    import torch.distributed as dist
    from deepreadmapper_b200 import sharded
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        bounds = sharded.shard_bounds(n_windows, world)
        lo, hi = bounds[rank]
        D, I = _local_lists(rank, n, k, lo, hi)
        tD, tI = torch.from_numpy(D.copy()), torch.from_numpy(I.copy())
        sharded.globalize_(tI, lo)
        D_all, I_all = sharded.all_gather_topk(tD, tI)
        ok = tuple(D_all.shape) == (world, n, k) and tuple(I_all.shape) == (world, n, k)
        for g in range(world):
            Dg, Ig = _local_lists(g, n, k, *bounds[g])
            Ig = np.where(Ig >= 0, Ig + bounds[g][0], -1)
            ok = ok and np.array_equal(D_all[g].numpy(), Dg) and np.array_equal(I_all[g].numpy(), Ig)
        # the CUDA-only merge must refuse CPU tensors loudly
        try:
            sharded.merge_topk(D_all, I_all)
            ok = False
        except Exception as e:
            ok = ok and "CUDA" in str(e)
        q.put((rank, bool(ok), sharded.read_slice(1001, rank, world)))
    except Exception as e:  # report instead of leaving the parent waiting on the queue
        q.put((rank, repr(e), None))
    finally:
        dist.destroy_process_group()


def test_all_gather_layout_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port, world = _free_port(), 2
    ps = [ctx.Process(target=_worker, args=(r, world, port, 6, 16, 1702, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert [r[1] for r in res] == [True, True]
    assert res[0][2] == (0, 500) and res[1][2] == (500, 1001)  # contiguous read split, no gap, no overlap


def test_shard_bounds_cover_ids_and_keep_strand_pairs():
    from deepreadmapper_b200 import sharded
    for n_windows, world in [(1702, 2), (9283006, 8), (10, 4), (7, 3)]:
        b = sharded.shard_bounds(n_windows, world)
        assert b[0][0] == 0 and b[-1][1] == n_windows
        for (l0, h0), (l1, h1) in zip(b, b[1:]):
            assert h0 == l1 and l1 % 2 == 0
    assert sharded.read_chunks(10, 4) == [(0, 4), (4, 8), (8, 10)]
