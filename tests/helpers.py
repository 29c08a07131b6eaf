"""Shared checks for the oracle and GPU tests (tree invariants, T3)."""
import numpy as np


def check_tree_invariants(n, left, right, parent, idx=None):
    """N-1 internal nodes, every leaf reached exactly once, parents consistent,
    leaf ranges nest (each internal node covers a contiguous range of leaves)."""
    if n == 0:
        return
    if n == 1:
        assert parent.shape == (1,) and parent[0] == 0xFFFFFFFF
        return
    leaf0 = n - 1
    assert left.shape == (n - 1,) and right.shape == (n - 1,) and parent.shape == (2 * n - 1,)
    assert parent[0] == 0xFFFFFFFF
    kids = np.concatenate([left, right])
    # every node except the root is a child exactly once
    assert np.array_equal(np.sort(kids), np.arange(1, 2 * n - 1, dtype=np.uint32))
    ar = np.arange(n - 1, dtype=np.uint32)
    assert np.array_equal(parent[left], ar) and np.array_equal(parent[right], ar)
    # ranges: iterative post-order from the root
    lo = np.full(2 * n - 1, -1, dtype=np.int64)
    hi = np.full(2 * n - 1, -1, dtype=np.int64)
    lo[leaf0:] = np.arange(n)
    hi[leaf0:] = np.arange(n)
    order = []
    stack = [0]
    while stack:
        v = stack.pop()
        order.append(v)
        for c in (int(left[v]), int(right[v])):
            if c < leaf0:
                stack.append(c)
    assert len(order) == n - 1
    for v in reversed(order):
        l, r = int(left[v]), int(right[v])
        assert hi[l] + 1 == lo[r], "children must cover adjacent leaf ranges"
        lo[v], hi[v] = lo[l], hi[r]
    assert lo[0] == 0 and hi[0] == n - 1
    if idx is not None:
        assert np.array_equal(np.sort(idx), np.arange(n, dtype=np.uint32))


def check_boxes(n, left, right, boxes, tris, idx):
    """Leaf boxes bound their triangle exactly; internal boxes are the union."""
    if n == 0:
        return
    leaf0 = n - 1
    t = tris.reshape(n, 3, 4)[:, :, :3][idx]
    assert np.array_equal(boxes[leaf0:, :3], t.min(axis=1))
    assert np.array_equal(boxes[leaf0:, 3:], t.max(axis=1))
    if n > 1:
        assert np.array_equal(boxes[:leaf0, :3], np.minimum(boxes[left, :3], boxes[right, :3]))
        assert np.array_equal(boxes[:leaf0, 3:], np.maximum(boxes[left, 3:], boxes[right, 3:]))


def compare_hits(h_a, h_b, rtol=1e-5):
    """The BASELINE.json:5 bar: identical hit/miss and primitive id; t within
    rtol relative; u, v within rtol (they live in [0, 1])."""
    pa = np.ascontiguousarray(h_a[:, 3]).view(np.int32)
    pb = np.ascontiguousarray(h_b[:, 3]).view(np.int32)
    assert np.array_equal(pa, pb), f"{(pa != pb).sum()} prim ids differ"
    m = (pa >= 0) & (pa != 0x7FFFFFFF)
    a, b = h_a[m].astype(np.float64), h_b[m].astype(np.float64)
    assert (np.abs(a[:, 0] - b[:, 0]) <= rtol * np.abs(b[:, 0])).all(), "t beyond tolerance"
    assert (np.abs(a[:, 1:3] - b[:, 1:3]) <= rtol).all(), "u/v beyond tolerance"
