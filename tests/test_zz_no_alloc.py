"""rt_sort_pairs_static: the sort entry point that allocates nothing (static device scratch,
no counterpart in the reference).

Collected last on purpose: its output is compared with the
existing entry point rt_sort_pairs on the same input, bit for bit, on calls in a row that share
the static scratch, through the C ABI.  Passes alone in a fresh process and after the others."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _pairs(n, seed, distinct_bits=30):
    rng = np.random.default_rng(seed)
    keys = rng.integers(0, 1 << distinct_bits, size=n, dtype=np.uint32)
    if distinct_bits < 30:
        keys = (keys * np.uint32(0x01041041 >> 2)) & np.uint32((1 << 30) - 1)   # spread over all digits
    vals = rng.permutation(n).astype(np.uint32)
    return keys, vals


def test_static_sort_equals_the_allocating_sort(rt):
    from gpu_raytracer_b200 import _lib
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    mx = _lib.RT_SORT_STATIC_MAX_PAIRS
    L = rt.lib()
    # calls in a row on the shared scratch: large, then smaller (stale contents of the first
    # must not leak), ragged sizes, few distinct keys (stability decides), the maximum
    for i, (n, bits_) in enumerate([(1_000_003, 30), (70_001, 30), (mx, 30), (300_000, 4), (1, 30), (7681, 30)]):
        keys, vals = _pairs(n, 100 + i, bits_)
        k_ref = torch.from_numpy(keys.view(np.int32)).cuda()
        v_ref = torch.from_numpy(vals.view(np.int32)).cuda()
        k_new, v_new = k_ref.clone(), v_ref.clone()
        rt.sort_pairs(k_ref, v_ref, 30)                      # the existing entry point
        rt.sort_pairs_static(k_new, v_new)                   # enqueued, not synchronised
        torch.cuda.synchronize()
        assert torch.equal(k_new, k_ref) and torch.equal(v_new, v_ref), (n, bits_)
        order = np.argsort(keys, kind="stable")
        assert np.array_equal(k_new.cpu().numpy().view(np.uint32), keys[order]), n
        assert np.array_equal(v_new.cpu().numpy().view(np.uint32), vals[order]), n
    # on a side stream, twice back to back on the same stream
    s = torch.cuda.Stream()
    keys, vals = _pairs(200_000, 7)
    a_k, a_v = torch.from_numpy(keys.view(np.int32)).cuda(), torch.from_numpy(vals.view(np.int32)).cuda()
    b_k, b_v = a_k.clone(), a_v.clone()
    torch.cuda.synchronize()
    rt.sort_pairs_static(a_k, a_v, stream=s.cuda_stream)
    rt.sort_pairs_static(b_k, b_v, stream=s.cuda_stream)
    s.synchronize()
    order = np.argsort(keys, kind="stable")
    for k, v in ((a_k, a_v), (b_k, b_v)):
        assert np.array_equal(k.cpu().numpy().view(np.uint32), keys[order])
        assert np.array_equal(v.cpu().numpy().view(np.uint32), vals[order])
    # beyond the maximum: a clear error, nothing touched
    big_k = torch.zeros(mx + 1, dtype=torch.int32, device="cuda")
    big_v = torch.zeros(mx + 1, dtype=torch.int32, device="cuda")
    rc = L.rt_sort_pairs_static(big_k.data_ptr(), big_v.data_ptr(), mx + 1, None)
    assert rc == _lib.RT_ERR_LIMIT and b"RT_SORT_STATIC_MAX_PAIRS" in L.bvh_last_error()
    # host memory is rejected, n = 0 is a no-op
    h = np.zeros(8, dtype=np.uint32)
    assert L.rt_sort_pairs_static(h.ctypes.data, h.ctypes.data, 8, None) == _lib.RT_ERR_INVALID
    assert L.rt_sort_pairs_static(None, None, 0, None) == _lib.RT_OK
