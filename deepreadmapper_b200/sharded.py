"""Position-sharded index over the GPUs of one box (BASELINE.json configs[4]).

One process per GPU (torch.distributed).  Rank g holds the HNSW-PQ index of the reference windows whose ids fall in
[id_lo_g, id_hi_g) - its own graph, labels local to the shard - and every rank sees every read of the chunk:

    local search  ->  labels + id_lo_g  ->  ONE all-gather of (D f32, I i64)[n, k] per rank  ->  per-read K-way merge
                                             (NCCL over NVLink)                                   (drm_merge_topk_device)

after which the SW rerank runs read-sharded again from the replicated genome.  The collective goes through
torch.distributed so that the same code runs on the gloo backend in the CPU tests; the merge itself exists only as a
CUDA kernel (no CPU product path).  A sharded search explores n_shards graphs with ef each, so it is NOT
result-identical to one big graph; its parity target is the oracle run the same sharded way (tests) and its recall is
reported against the exact ADC scan.
"""
from __future__ import annotations


def shard_bounds(n_windows: int, world: int):
    """Contiguous, strand-pair aligned id ranges: [(lo, hi)] with lo/hi even (ids 2p and 2p+1 stay together)."""
    n_pos = (n_windows + 1) // 2
    out = []
    for g in range(world):
        lo, hi = n_pos * g // world, n_pos * (g + 1) // world
        out.append((2 * lo, min(2 * hi, n_windows)))
    return out


def read_chunks(n_reads: int, chunk: int):
    return [(s, min(s + chunk, n_reads)) for s in range(0, n_reads, chunk)]


def read_slice(n_reads: int, rank: int, world: int):
    """The reads rank `rank` reranks after the merge (contiguous split, like the replicated-index mode)."""
    return n_reads * rank // world, n_reads * (rank + 1) // world


def globalize_(I_local, id_lo: int):
    """Shard-local labels -> global window ids, in place; unfilled slots stay -1."""
    I_local += (I_local >= 0).to(I_local.dtype) * id_lo
    return I_local


def all_gather_topk(D_local, I_global, group=None):
    """(D_all, I_all) of shape [world, n, k]: every rank's ascending top-k list for every read of the chunk.
    Works on any backend (nccl for CUDA tensors, gloo for CPU tensors)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return D_local.unsqueeze(0), I_global.unsqueeze(0)
    n = D_local.shape[0]
    D_cat = torch.empty((world * n,) + tuple(D_local.shape[1:]), dtype=D_local.dtype, device=D_local.device)
    I_cat = torch.empty((world * n,) + tuple(I_global.shape[1:]), dtype=I_global.dtype, device=I_global.device)
    dist.all_gather_into_tensor(D_cat, D_local.contiguous(), group=group)  # rank-major concatenation
    dist.all_gather_into_tensor(I_cat, I_global.contiguous(), group=group)
    D_all, I_all = D_cat.view((world,) + tuple(D_local.shape)), I_cat.view((world,) + tuple(I_global.shape))
    return D_all, I_all


def merge_topk(D_all, I_all, D_out=None, I_out=None, stream=None):
    """Global top-k per read, ordered by (distance, id): the CUDA merge kernel.  CUDA tensors only."""
    import torch
    from . import api
    if not D_all.is_cuda:
        raise api.DrmError(-3, "merge_topk needs CUDA tensors: the merge exists only as a CUDA kernel (no CPU fallback)")
    world, n, k = D_all.shape
    D_out = torch.empty((n, k), dtype=torch.float32, device=D_all.device) if D_out is None else D_out
    I_out = torch.empty((n, k), dtype=torch.int64, device=D_all.device) if I_out is None else I_out
    s = torch.cuda.current_stream(D_all.device) if stream is None else stream
    api.merge_topk_device(D_all.device.index, D_all.data_ptr(), I_all.data_ptr(), world, n, k, D_out.data_ptr(), I_out.data_ptr(), s.cuda_stream)
    return D_out, I_out


class ShardedSearcher:
    """Rank-local view of a position-sharded index: search every read against this rank's shard, exchange, merge."""

    def __init__(self, index, id_lo: int, device_index: int, group=None):
        self.index, self.id_lo, self.dev, self.group = index, int(id_lo), int(device_index), group

    def search_chunk(self, dQ, k: int, ef: int, counters_ptr: int = 0):
        """dQ: CUDA float32 (n, d).  Returns (D, I) of the GLOBAL top-k for every read of the chunk (on this rank's GPU)."""
        import torch
        n = dQ.shape[0]
        D = torch.empty((n, k), dtype=torch.float32, device=dQ.device)
        I = torch.empty((n, k), dtype=torch.int64, device=dQ.device)
        s = torch.cuda.current_stream(dQ.device)
        self.index.search_device(dQ.data_ptr(), n, k, ef, D.data_ptr(), I.data_ptr(), counters_ptr, s.cuda_stream)
        globalize_(I, self.id_lo)
        D_all, I_all = all_gather_topk(D, I, self.group)
        if D_all.shape[0] == 1:
            return D, I
        return merge_topk(D_all, I_all)
