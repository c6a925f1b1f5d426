"""Host scene builder (csrc/rt_scene.cpp) through rt_host_build_bvh -- no GPU needed.

The tree is the product's own structure (the reference has none: OpenCLRaytracer.cpp:13-33 only marshals the objects),
so these are invariants, not reference vectors: every live object sits in exactly one leaf, every child box contains
what is below it, the result does not depend on the number of builder threads.
"""
import numpy as np
import pytest

from pyrt import scenes as S
from pyrt import types as T


def _object_points(objs):
    """View-space points whose convex hull contains each object: (n, k, 3)."""
    n = len(objs)
    mv = objs["mv"].reshape(n, 4, 4).astype(np.float64)  # m[col][row]
    M = np.transpose(mv[:, :3, :3], (0, 2, 1))  # (n, row, col)
    t = mv[:, 3, :3]
    corners = np.array([[x, y, z] for x in (-1, 1) for y in (-1, 1) for z in (-1, 1)], dtype=np.float64)
    pts = np.zeros((n, 8, 3))
    for i, typ in enumerate(objs["type"]):
        if typ == T.SPHERE:  # unit sphere: the bounding cube [-1, 1]^3 of it is NOT inside the object's box in general,
            # so use the exact extent per axis: centre +- row norms
            r = np.sqrt((M[i] ** 2).sum(axis=1))
            pts[i] = t[i] + corners * r
        elif typ == T.BOX:
            pts[i] = t[i] + (0.5 * corners) @ M[i].T
        else:  # triangle (-.5,-.5,0) (.5,-.5,0) (-.5,.5,0)
            v = np.array([[-0.5, -0.5, 0], [0.5, -0.5, 0], [-0.5, 0.5, 0]] + [[-0.5, -0.5, 0]] * 5, dtype=np.float64)
            pts[i] = t[i] + v @ M[i].T
    return pts


def _walk(info, nodes):
    """(inner node index, side, child ref, box lo, box hi) for every child record reachable from the root."""
    out = []
    if info.root < 0:
        return out
    todo = [info.root]
    seen = set()
    while todo:
        i = todo.pop()
        assert i not in seen, "node %d reached twice" % i
        seen.add(i)
        nd = nodes[i]
        for side in (0, 1):
            c = int(nd["c%d" % side])
            ctr = nd["ctr%d" % side].astype(np.float64)
            ext = nd["ext%d" % side].astype(np.float64)
            out.append((i, side, c, ctr - ext, ctr + ext))
            if c >= 0:
                todo.append(c)
    assert len(seen) == len(nodes), "unreachable nodes"
    return out


@pytest.mark.parametrize("second_type", [T.BOX, T.TRIANGLE])
@pytest.mark.parametrize("leaf_size", [1, 4])
def test_tree_covers_every_object_and_boxes_nest(rt, second_type, leaf_size):
    objs, lights = S.synthetic_scene(3000, 200, second_type=second_type)
    objs["type"][7] = 9  # unknown type: never hit, not in the tree (the reference's switch has no default)
    info, nodes, leaf = rt.host_build_bvh(objs, lights, leaf_size=leaf_size)
    live = np.flatnonzero(objs["type"] <= 2)
    assert info.n_leaf_objs == len(live) and sorted(leaf.tolist()) == live.tolist()
    assert info.depth <= 64  # RT_BVH_STACK
    if leaf_size == 1:
        assert info.n_nodes == len(live) - 1
    pts = _object_points(objs)
    recs = _walk(info, nodes)
    boxes = {}  # inner node -> hull of its two child boxes
    in_leaves = 0
    for i, side, c, lo, hi in recs:
        assert np.all(hi >= lo)
        if c < 0:
            code = ~c
            first, cnt = code >> 4, (code & 15) + 1
            assert cnt <= leaf_size
            in_leaves += cnt
            for o in leaf[first:first + cnt]:
                assert np.all(pts[o] >= lo - 1e-9) and np.all(pts[o] <= hi + 1e-9), "object %d sticks out of its leaf box" % o
        b = boxes.setdefault(i, [np.full(3, np.inf), np.full(3, -np.inf)])
        b[0] = np.minimum(b[0], lo)
        b[1] = np.maximum(b[1], hi)
    assert in_leaves == len(live)
    for i, side, c, lo, hi in recs:
        if c >= 0:  # the record of an inner child contains both boxes stored in that child
            tol = 1e-6 * np.maximum(1.0, np.abs(hi))
            assert np.all(boxes[c][0] >= lo - tol) and np.all(boxes[c][1] <= hi + tol), "node %d sticks out of its parent's record" % c
    f = np.concatenate([nodes["f0"], nodes["f1"]])
    assert np.all(f > 0) and np.all(f < 0.1)


def test_tree_does_not_depend_on_the_thread_count(rt):
    """Subtrees are built on worker threads and spliced; the numbering is a preorder walk, so the bytes are the same."""
    objs, lights = S.synthetic_scene(20000, 800)
    ref = None
    for threads in (1, 2, 5, 0):
        info, nodes, leaf = rt.host_build_bvh(objs, lights, threads=threads)
        cur = (info.n_nodes, info.depth, info.root, nodes.tobytes(), leaf.tobytes())
        if ref is None:
            ref = cur
        assert cur == ref, "tree differs with %d threads" % threads
    # preorder layout: a node's left child, when inner, is the next node
    inner_left = nodes["c0"] >= 0
    assert np.array_equal(nodes["c0"][inner_left], np.flatnonzero(inner_left) + 1)


def test_builder_edge_cases_and_errors(rt):
    objs, lights = S.synthetic_scene(5, 3)
    info, nodes, leaf = rt.host_build_bvh(objs[:1], lights)  # one object: the root is a leaf
    assert info.root < 0 and info.n_leaf_objs == 1
    info, nodes, leaf = rt.host_build_bvh(objs[:0], lights)  # empty scene: a placeholder leaf that nothing can hit
    assert info.root == -1 and info.n_leaf_objs == 1 and info.n_nodes == 1
    unknown = objs.copy()
    unknown["type"] = 5  # objects of unknown type only: same placeholder
    info, nodes, leaf = rt.host_build_bvh(unknown, lights)
    assert info.root == -1 and info.n_leaf_objs == 1
    same = np.repeat(objs[:1], 40)  # coincident centres: median splits
    info, nodes, leaf = rt.host_build_bvh(same, lights)
    assert info.n_leaf_objs == 40 and sorted(leaf.tolist()) == list(range(40))
    bad = objs.copy()
    bad["mvInverse"][2][:] = 0.0
    with pytest.raises(rt.RtError) as e:
        rt.host_build_bvh(bad, lights)
    assert "object 2" in str(e.value)
    with pytest.raises(rt.RtError):
        rt.host_build_bvh(objs, lights, leaf_size=17)


def test_readback_band_plan_partitions_the_frame(rt):
    """rt_plan_bands: the split rt_render uses for its readback bands -- contiguous, complete, largest band first."""
    for height, br, ns in ((4320, 1, 1), (4320, 16, 8), (2069, 16, 2), (37, 16, 3), (5, 16, 1)):
        for shard in range(ns):
            n_blocks = -(-height // br)
            units = height if ns == 1 else len(range(shard, n_blocks, ns))
            for nb in (1, 2, 3, 16):
                for ratio in (1.25, 1.85, 7.0, 16.0):
                    b = rt.plan_bands(height, br, shard, ns, nb, ratio)
                    assert b[0] == 0 and b[-1] == units and np.all(np.diff(b.astype(np.int64)) >= 0)
                    sizes = np.diff(b.astype(np.int64))
                    if units >= 64 * nb and nb <= 3:
                        assert np.all(sizes[:-1] >= sizes[1:]) and sizes[-1] > 0
                    if nb == 2 and units >= 64:
                        assert abs(sizes[0] / max(sizes[1], 1) - ratio) < 0.1 * ratio + 0.6
    with pytest.raises(rt.RtError):
        rt.plan_bands(100, 16, 3, 2, 2, 7.0)


def test_launch_sequence_plan_of_the_default_readback(rt):
    """rt_plan_sequences: how a default rt_render cuts its frame into launch sequences (each a readback band) -- contiguous,
    complete, never growing; shrinking by 0.7 where the copy is fast, equal and more numerous where it is slow; fewer for
    small shards; at least four units per sequence."""
    W = 7680
    for height, br, ns in ((4320, 1, 1), (4320, 16, 2), (4320, 16, 8), (2069, 16, 2), (37, 16, 3), (5, 16, 1), (180, 1, 1)):
        for shard in range(ns):
            n_blocks = -(-height // br)
            units = height if ns == 1 else len(range(shard, n_blocks, ns))
            for ratio in (0.3, 0.9, 1.25, 2.5, 8.0):
                for k in (0, 1, 2, 5, 16):
                    b = rt.plan_sequences(W, height, br, shard, ns, k, 0.0, ratio).astype(np.int64)
                    sizes = np.diff(b)
                    assert b[0] == 0 and b[-1] == units and np.all(sizes >= 0)
                    n = len(sizes)
                    assert 1 <= n <= 16 and (k == 0 or n <= max(k, 1)) and (n == 1 or units >= 4 * n)
                    if n > 1:
                        assert np.all(sizes[:-1] + 1 >= sizes[1:]) and sizes[-1] > 0  # rounding may move one unit
    # one GPU, idle link: six sequences, each 0.7 of the one before
    s = np.diff(rt.plan_sequences(W, 4320, kernel_copy_ratio=2.5).astype(np.int64))
    assert len(s) == 6 and np.allclose(s[1:] / s[:-1], 0.7, atol=0.02)
    # one of eight ranks behind a shared host: (nearly) equal sequences, more of them, none below 0.4 Mpixel
    s = np.diff(rt.plan_sequences(W, 4320, 16, 3, 8, kernel_copy_ratio=1.0).astype(np.int64))
    assert len(s) >= 7 and s.max() - s.min() <= 1 and s.min() * 16 * W >= 400000
    # a small frame is not cut into crumbs
    assert len(rt.plan_sequences(320, 180, kernel_copy_ratio=2.5)) - 1 <= 2
    # explicit ratio
    s = np.diff(rt.plan_sequences(W, 4320, n_sequences=4, q=0.5).astype(np.int64))
    assert len(s) == 4 and np.allclose(s[1:] / s[:-1], 0.5, atol=0.02)
    for bad in (dict(n_sequences=17), dict(q=1.5), dict(kernel_copy_ratio=0.0), dict(shard=2, n_shards=2)):
        with pytest.raises(rt.RtError):
            rt.plan_sequences(W, 4320, **bad)
