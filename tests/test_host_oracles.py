"""CPU oracles of hw2/hw3/hw4 in the native host library vs plain NumPy restatements."""
import numpy as np
import pytest

from cme213_sp18_b200.ops import host
from cme213_sp18_b200.utils import heat as heat_utils


def test_shift_bytes_wraps():
    d = np.arange(256, dtype=np.uint8)
    assert np.array_equal(host.shift_bytes(d, 25), (d.astype(np.int32) + 25).astype(np.uint8))


def test_pagerank_graph_shape_matches_reference_rule():
    # ref hw2code/main_q2.cu:87-120: node i has i/(N/(2*avg-1)+1)+1 edges
    n, avg = 1000, 3
    rp, edges, inv = host.pagerank_graph(n, avg, seed=5)
    per_block = n // (2 * avg - 1) + 1
    deg = np.arange(n) // per_block + 1
    assert np.array_equal(np.diff(rp), deg)
    assert np.allclose(inv, 1.0 / deg.astype(np.float32))
    assert edges.max() < n and len(edges) == rp[-1]
    rp2, edges2, _ = host.pagerank_graph(n, avg, seed=5)
    assert np.array_equal(edges, edges2)          # reproducible (the reference was not, Q16)


def test_pagerank_matches_numpy():
    n, avg, iters = 3000, 4, 6
    rp, edges, inv = host.pagerank_graph(n, avg, seed=11)
    r0 = np.full(n, 1.0 / n, dtype=np.float32)
    got = host.pagerank(rp, edges, r0, inv, iters)
    cur = r0.copy()
    for _ in range(iters):
        nxt = np.empty_like(cur)
        for i in range(n):
            s = np.float32(0)
            for j in range(rp[i], rp[i + 1]):
                s = np.float32(s + np.float32(cur[edges[j]] * inv[edges[j]]))
            nxt[i] = np.float32(np.float32(0.5) / np.float32(n) + np.float32(0.5) * s)
        cur = nxt
    ulp, _ = host.max_ulp(got, cur)
    assert ulp <= 2


def _np_heat(grid, order, iters, xcfl, ycfl, s, semantics):
    b = order // 2
    coef = {2: [1, -2, 1], 4: [-1, 16, -30, 16, -1],
            8: [-9, 128, -1008, 8064, -14350, 8064, -1008, 128, -9]}[order]
    cur = grid.astype(np.float64).copy()
    nxt = cur.copy()
    gy, gx = cur.shape
    mask = np.ones_like(cur, dtype=bool)
    mask[b:gy - b, b:gx - b] = False
    for _ in range(iters):
        if semantics == "cpu":
            cur[mask] = nxt[mask] * s
        else:
            nxt[mask] = cur[mask] * s
        c = cur[b:gy - b, b:gx - b]
        xs = sum(coef[k + b] * cur[b:gy - b, b + k:gx - b + k] for k in range(-b, b + 1))
        ys = sum(coef[k + b] * cur[b + k:gy - b + k, b:gx - b] for k in range(-b, b + 1))
        nxt[b:gy - b, b:gx - b] = c + xcfl * xs + ycfl * ys
        cur, nxt = nxt, cur
    return cur


@pytest.mark.parametrize("order", [2, 4, 8])
@pytest.mark.parametrize("semantics", ["gpu", "cpu"])
def test_heat_cpu_against_float64_numpy(order, semantics):
    p = heat_utils.SimParams(nx=96, ny=80, lx=1.0, ly=1.0, iters=30, order=order)
    g0 = host.heat_init(p.gx, p.gy, p.dx, p.dy)
    got = host.heat_cpu(g0, order, p.iters, p.xcfl, p.ycfl, p.bc_scale, semantics)
    want = _np_heat(g0, order, p.iters, float(np.float32(p.xcfl)), float(np.float32(p.ycfl)),
                    float(np.float32(p.bc_scale)), semantics)
    err = np.abs(got - want).max() / np.abs(want).max()
    assert err < 5e-5, err


def test_heat_init_is_sin_sin():
    p = heat_utils.SimParams(nx=20, ny=10, lx=1.0, ly=2.0, iters=1, order=4)
    g = host.heat_init(p.gx, p.gy, p.dx, p.dy)
    i, j = np.meshgrid(np.arange(p.gx), np.arange(p.gy))
    assert np.array_equal(g, (np.sin(i * p.dx) * np.sin(j * p.dy)).astype(np.float32))


def test_sim_params_match_reference_defaults():
    # ref hw3code/params.in + simParams.cpp:53-77
    p = heat_utils.SimParams(nx=4096, ny=4096, lx=1.0, ly=1.0, iters=400, order=8)
    assert (p.gx, p.border) == (4104, 4)
    assert p.calc_bytes() == 400 * 4096 * 4096 * 18 * 4
    assert np.float32(p.bc_scale) == np.float32(1.0)     # border is constant at this size


def test_max_ulp():
    a = np.array([1.0, -1.0, 0.0, 1e-30], dtype=np.float32)
    b = np.nextafter(a, np.float32(np.inf)).astype(np.float32)
    assert host.max_ulp(a, b)[0] == 1
    assert host.max_ulp(a, a)[0] == 0
    assert host.max_ulp(np.array([0.0], np.float32), np.array([-0.0], np.float32))[0] == 0
    c = np.array([1.0, 2.0], dtype=np.float64)
    d = np.nextafter(np.nextafter(c, np.inf), np.inf)
    assert host.max_ulp(c, d)[0] == 2
    assert host.max_ulp(np.array([np.nan], np.float32), np.array([1.0], np.float32))[0] > 2**62


def test_letter_histogram():
    t = np.frombuffer(b"abcabcaaz", dtype=np.uint8)
    h = host.letter_histogram(t)
    assert h[ord("a")] == 4 and h[ord("z")] == 1 and h.sum() == 9
    h2 = host.letter_histogram(t, stride=3, offset=1)
    assert h2[ord("b")] == 2 and h2[ord("a")] == 1 and h2.sum() == 3


def test_ulp_helpers_and_strided_range():
    from cme213_sp18_b200.utils import strided, ulp

    assert ulp.almost_equal_ulps(1.0, np.nextafter(1.0, 2.0), max_ulps=1)
    assert not ulp.almost_equal_ulps(1.0, 1.0 + 1e-9, max_ulps=512)
    m = np.arange(6, dtype=np.float64).reshape(2, 3)
    assert ulp.almost_equal_matrix(m, m.T.ravel(), column_major=True)
    assert ulp.almost_equal_matrix(m, m.ravel(), column_major=False)
    assert not ulp.almost_equal_matrix(m, m.ravel(), column_major=True)
    assert ulp.max_ulps_f32([1.0, 2.0], [1.0, np.nextafter(np.float32(2.0), np.float32(3.0))]) == 1
    assert list(strided.strided_range(list(range(10)), 1, 3)) == [1, 4, 7]
    assert strided.strided_range(np.arange(10), 2, 4, 9).tolist() == [2, 6]


def test_vector_file_roundtrip(tmp_path):
    from cme213_sp18_b200.utils import vecio

    v = np.array([3, 0, 4294967295, 17], dtype=np.uint32)
    vecio.write_vector_to_file(tmp_path / "vec", v)
    assert np.array_equal(vecio.read_vector_from_file(tmp_path / "vec"), v)


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_edge_balanced_row_partition(world):
    """Row partition of the multi-GPU PageRank: contiguous, complete, and balanced in edges to
    within one row's degree (node degree grows with the index, ref hw2code/main_q2.cu:100-106)."""
    from cme213_sp18_b200.parallel.pagerank_multi import edge_balanced_cuts

    row_ptr, _, _ = host.pagerank_graph(50_000, 6, 1, with_edges=False)
    cuts = edge_balanced_cuts(row_ptr, world)
    assert cuts[0] == 0 and cuts[-1] == 50_000 and len(cuts) == world + 1
    assert all(a <= b for a, b in zip(cuts, cuts[1:]))
    m = int(row_ptr[-1])
    max_deg = int(np.diff(row_ptr).max())
    for r in range(world):
        edges = int(row_ptr[cuts[r + 1]] - row_ptr[cuts[r]])
        assert abs(edges - m / world) <= max_deg + 1, (r, edges, m / world)


def test_column_blocked_csr_regroups_edges_by_source_block():
    """The set-up step of the column-blocked PageRank: every (block, row) segment holds exactly the
    row's edges that point into the block, in their original order."""
    import torch

    from cme213_sp18_b200.parallel.pagerank_multi import column_blocked_csr

    n, avg = 5000, 5
    rp, edges, _ = host.pagerank_graph(n, avg, 3)
    lo, hi = 1200, 4100                                   # a slice of rows, like one rank's shard
    local_ptr = torch.from_numpy((rp[lo: hi + 1] - rp[lo]).astype(np.int64))
    e = torch.from_numpy(edges[rp[lo]: rp[hi]].astype(np.int64).astype(np.int32))
    block = 700
    blk_ptr, edges_b, nb = column_blocked_csr(local_ptr, e, n, block)
    assert nb == (n + block - 1) // block and blk_ptr.numel() == nb * (hi - lo) + 1
    assert int(blk_ptr[-1]) == e.numel()
    bp, eb = blk_ptr.numpy(), edges_b.numpy()
    for i in (0, 1, 57, 1500, hi - lo - 1):
        row = edges[rp[lo + i]: rp[lo + i + 1]]
        for b in range(nb):
            seg = eb[bp[b * (hi - lo) + i]: bp[b * (hi - lo) + i + 1]]
            want = row[(row // block) == b]
            assert np.array_equal(seg.astype(np.int64), want.astype(np.int64))
