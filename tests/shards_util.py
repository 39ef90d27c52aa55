"""Shared by tests/test_gpu_shards.py, tests/shards_rank_check.py and bench.py's sharded arm: builds position shards of a
genome and runs the CPU oracle the same sharded way (per-shard FAISS-rule search, labels made global, (distance, id) merge)."""
import os

import numpy as np


def build_shards(drm, synth, sharded, genome, world, out_dir, efc=64, seed=20, builder="cpu", ref_len=150):
    """One IHNp file per shard (labels local).  Returns (paths, id_lo list, n_windows)."""
    n_windows = 2 * (len(genome) - ref_len + 1)
    paths, los = [], []
    for rank, (lo, hi) in enumerate(sharded.shard_bounds(n_windows, world)):
        path = os.path.join(out_dir, f"shard{rank}of{world}.index")
        E = synth.embed_windows(genome[lo // 2: hi // 2 + ref_len - 1])
        assert E.shape[0] == hi - lo
        if builder == "gpu":
            drm.build_index_gpu(E, path, efc=efc, seed=seed + rank)
        else:
            drm.build_index(E, path, efc=efc, seed=seed + rank)
        paths.append(path)
        los.append(lo)
    return paths, los, n_windows


def oracle_sharded_search(ob, paths, los, Q, k, ef, threads=0):
    """The reference path run the sharded way on the CPU: (D, I) n x k, ordered by (distance, id), -1 / FLT_MAX padded."""
    Ds, Is = [], []
    for path, lo in zip(paths, los):
        oix = ob.OracleIndex(path)
        D0, I0, _ = oix.search(Q, k, ef, tie_rule=0, threads=threads)
        oix.close()  # one shard in host memory at a time (a chromosome-scale shard file is 11 GB)
        Ds.append(D0)
        Is.append(np.where(I0 >= 0, I0 + lo, -1))
    Da, Ia = np.concatenate(Ds, axis=1), np.concatenate(Is, axis=1)
    # unfilled slots sort last: (FLT_MAX, id -1 read as the largest unsigned)
    key_i = Ia.astype(np.uint64)
    order = np.lexsort((key_i, Da), axis=1)[:, :k]
    return np.take_along_axis(Da, order, axis=1), np.take_along_axis(Ia, order, axis=1)
