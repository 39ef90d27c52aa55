"""One rank of a rank-per-process sharded search (run by tests/test_gpu_shards.py): opens its shard with
drm_shards_open_rank, runs every exchange mode through the host-buffer call and the device-step call, and checks its rows
of the result against the sharded oracle's (saved by the parent)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank, world, path, lo, work, k, ef = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], int(sys.argv[4]), sys.argv[5], int(sys.argv[6]), int(sys.argv[7])
    import deepreadmapper_b200 as drm
    idf = os.path.join(work, "nccl_id.bin")
    if rank == 0:
        uid = drm.Shards.unique_id()
        with open(idf + ".tmp", "wb") as f:
            f.write(uid)
        os.replace(idf + ".tmp", idf)
    else:
        t0 = time.time()
        while not os.path.exists(idf):
            assert time.time() - t0 < 120, "no ncclUniqueId from rank 0"
            time.sleep(0.05)
        uid = open(idf, "rb").read()
    sh = drm.Shards(path, lo, rank, rank=rank, world=world, unique_id=uid)
    assert (sh.world, sh.rank, sh.id_lo, sh.one_process) == (world, rank, lo, 0)
    Q, D0, I0 = np.load(os.path.join(work, "Q.npy")), np.load(os.path.join(work, "D0.npy")), np.load(os.path.join(work, "I0.npy"))
    n = len(Q)
    for exchange in (drm.XCHG_FUSED, drm.XCHG_ALLTOALL, drm.XCHG_ALLGATHER, drm.XCHG_FUSED):
        for chunk in (0, 250):
            D, I, _, _, st = sh.search_rerank(Q, k, ef, exchange=exchange, chunk=chunk)
            c = chunk or n
            rows = np.concatenate([np.arange(c0 + (min(c, n - c0) * rank) // world, c0 + (min(c, n - c0) * (rank + 1)) // world) for c0 in range(0, n, c)])
            assert np.array_equal(I[rows], I0[rows]) and np.array_equal(D[rows].view(np.uint32), D0[rows].view(np.uint32)), (exchange, chunk)
            others = np.setdiff1d(np.arange(n), rows)
            assert (I[others] == -2).all()  # only this rank's rows are written
    # the device-step call (what bench.py uses under torchrun)
    import torch
    torch.cuda.set_device(rank)
    dQ = torch.from_numpy(Q).cuda()
    for exchange in (drm.XCHG_FUSED, drm.XCHG_ALLTOALL):
        Dp, Ip, slo, cnt, stream = sh.step_device(dQ.data_ptr(), n, Q.shape[1], k, ef, exchange)
        sh.stats()  # synchronises the handle's stream
        out_D = torch.empty((cnt, k), dtype=torch.float32, device="cuda")
        out_I = torch.empty((cnt, k), dtype=torch.int64, device="cuda")
        from cuda import cudart  # cuda-python: a plain device-to-device copy out of the handle's buffers
        d2d = cudart.cudaMemcpyKind.cudaMemcpyDeviceToDevice
        assert cudart.cudaMemcpy(out_D.data_ptr(), Dp, cnt * k * 4, d2d)[0] == cudart.cudaError_t.cudaSuccess
        assert cudart.cudaMemcpy(out_I.data_ptr(), Ip, cnt * k * 8, d2d)[0] == cudart.cudaError_t.cudaSuccess
        assert np.array_equal(out_I.cpu().numpy(), I0[slo:slo + cnt]) and np.array_equal(out_D.cpu().numpy().view(np.uint32), D0[slo:slo + cnt].view(np.uint32))
    sh.close()
    print("rank ok", rank)


if __name__ == "__main__":
    main()
