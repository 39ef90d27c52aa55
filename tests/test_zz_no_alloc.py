"""drm_sw_pairs_device_noalloc (include/drm_b200.h): the generic SW pair kernel with its DP rows in a static device array,
for callers that may not allocate.  No counterpart in the reference; checked against the existing entry point drm_sw_pairs
(itself checked against the oracle's calc_sw_score in test_gpu_parity.py), bit for bit - the scores are integers.

Collected last on purpose (file name)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _pairs(seed, n, max_a, max_b):
    rng = np.random.default_rng(seed)
    alphabet = np.frombuffer(b"ACGTN", dtype=np.uint8)
    a_list, b_list = [], []
    for i in range(n):
        la, lb = int(rng.integers(0, max_a + 1)), int(rng.integers(0, max_b + 1))
        a = alphabet[rng.integers(0, 5, la)]
        if i % 3 == 0 and la and lb:  # related pairs, so that scores are not all tiny
            b = np.resize(a, lb).copy()
            flip = rng.random(lb) < 0.1
            b[flip] = alphabet[rng.integers(0, 4, int(flip.sum()))]
        else:
            b = alphabet[rng.integers(0, 5, lb)]
        a_list.append(a.tobytes()); b_list.append(b.tobytes())
    return a_list, b_list


def _device_args(torch, drm, a_list, b_list):
    from deepreadmapper_b200.api import flatten
    a_flat, a_off = flatten(a_list)
    b_flat, b_off = flatten(b_list)
    t = lambda buf, dt: torch.from_numpy(np.frombuffer(bytes(buf) + b"\0" * 16, dtype=dt).copy()).cuda()
    da, db = t(a_flat, np.uint8), t(b_flat, np.uint8)
    dao, dbo = torch.from_numpy(np.asarray(a_off, dtype=np.int64)).cuda(), torch.from_numpy(np.asarray(b_off, dtype=np.int64)).cuda()
    return da, dao, db, dbo


def _run(torch, drm, a_list, b_list, max_len_b, fill=-7):
    da, dao, db, dbo = _device_args(torch, drm, a_list, b_list)
    out = torch.full((max(len(a_list), 1),), fill, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    drm.sw_pairs_device_noalloc(0, da.data_ptr(), dao.data_ptr(), db.data_ptr(), dbo.data_ptr(), len(a_list), max_len_b, out.data_ptr(),
                                torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return out.cpu().numpy()[:len(a_list)]


def test_noalloc_pairs_match_existing_entry_point_two_calls_in_a_row():
    import torch
    import deepreadmapper_b200 as drm
    # call 1: many long pairs (every DP row of the static array gets dirty); call 2: fewer, shorter pairs with another
    # max_len_b, i.e. another row layout over the stale contents; call 3: the first again
    a1, b1 = _pairs(11, 3000, 200, 300)
    a2, b2 = _pairs(12, 257, 152, 150)
    want1, want2 = drm.sw_pairs(a1, b1), drm.sw_pairs(a2, b2)
    assert want1.max() > 50 and want2.max() > 50
    got1 = _run(torch, drm, a1, b1, 300)
    got2 = _run(torch, drm, a2, b2, 150)
    got3 = _run(torch, drm, a1, b1, 4095)
    assert np.array_equal(got1, want1)
    assert np.array_equal(got2, want2)
    assert np.array_equal(got3, want1)


def test_noalloc_pairs_limits_and_edge_cases():
    import torch
    import deepreadmapper_b200 as drm
    from deepreadmapper_b200 import api
    a, b = _pairs(13, 64, 100, 120)
    b[5] = b"ACGT" * 100  # longer than the max_len_b passed below
    want = drm.sw_pairs(a, b)
    got = _run(torch, drm, a, b, 120)
    assert got[5] == -1
    keep = np.arange(len(a)) != 5
    assert np.array_equal(got[keep], want[keep])
    # beyond the static scratch: a clear error, no launch
    with pytest.raises(drm.DrmError, match="above the limit of 4095"):
        _run(torch, drm, a, b, api.SW_PAIRS_NOALLOC_MAX_LEN_B + 1)
    # at the limit
    a2, b2 = [b"ACGT" * 1000, b""], [b"ACGT" * 1023 + b"ACG", b"ACGT"]
    assert np.array_equal(_run(torch, drm, a2, b2, api.SW_PAIRS_NOALLOC_MAX_LEN_B), drm.sw_pairs(a2, b2))
    # n == 0 is a no-op
    assert _run(torch, drm, [], [], 10).size == 0


def test_noalloc_pairs_inside_a_stream_capture():
    """A stream capture in global mode refuses cudaMalloc / cudaFree and the stream-ordered allocator's synchronous paths:
    the call goes through a capture and the replayed graph gives the same scores."""
    import torch
    import deepreadmapper_b200 as drm
    a, b = _pairs(14, 500, 152, 152)
    want = drm.sw_pairs(a, b)
    assert np.array_equal(_run(torch, drm, a, b, 152), want)  # set-up call: symbol address, module load
    da, dao, db, dbo = _device_args(torch, drm, a, b)
    out = torch.zeros(len(a), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, capture_error_mode="global"):
        drm.sw_pairs_device_noalloc(0, da.data_ptr(), dao.data_ptr(), db.data_ptr(), dbo.data_ptr(), len(a), 152, out.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream)
    for _ in range(2):
        out.fill_(-3)
        g.replay()
        torch.cuda.synchronize()
        assert np.array_equal(out.cpu().numpy(), want)
