"""hw1 host workloads: parity sums and radix sort (ref hw1code/main_q1.cpp, main_q2.cpp, tests_q2.cpp)."""
import ctypes

import numpy as np
import pytest

from cme213_sp18_b200.ops import host


def glibc_rand(n):
    """First n values of glibc rand() with the default seed: how the reference's golden
    'input' file was produced (ref hw1code/test_files/input starts 1804289383)."""
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(1)
    return np.array([libc.rand() for _ in range(n)], dtype=np.uint32)


def np_stage_oracle(keys, num_blocks, bits, start):
    block = len(keys) // num_blocks
    nb = (len(keys) + block - 1) // block
    buckets = 1 << bits
    dig = (keys >> start) & (buckets - 1)
    bh = np.zeros((nb, buckets), dtype=np.uint32)
    for b in range(nb):
        bh[b] = np.bincount(dig[b * block:(b + 1) * block], minlength=buckets)
    gh = bh.sum(axis=0, dtype=np.uint32)
    gx = np.concatenate([[0], np.cumsum(gh)[:-1]]).astype(np.uint32)
    bx = gx[None, :] + np.concatenate([np.zeros((1, buckets), np.uint32), np.cumsum(bh, axis=0)[:-1]]).astype(np.uint32)
    order = np.argsort(dig, kind="stable")
    return block, nb, bh.ravel(), gh, gx, bx.ravel().astype(np.uint32), keys[order]


def test_parity_sum_matches_numpy():
    rng = np.random.default_rng(0)
    v = rng.integers(0, 101, size=1_000_003, dtype=np.uint32)
    exact, mod32 = host.parity_sum(v, parallel=True)
    assert exact == (int(v[v % 2 == 0].sum(dtype=np.uint64)), int(v[v % 2 == 1].sum(dtype=np.uint64)))
    assert host.parity_sum(v, parallel=False)[0] == exact
    assert mod32 == (exact[0] % 2**32, exact[1] % 2**32)
    for t in (1, 2, 3):
        assert host.parity_sum(v, parallel=True, threads=t)[0] == exact


def test_parity_sum_wraps_like_uint():
    v = np.full(3, 0xFFFFFFFE, dtype=np.uint32)
    exact, mod32 = host.parity_sum(v)
    assert exact[0] == 3 * 0xFFFFFFFE and mod32[0] == (3 * 0xFFFFFFFE) % 2**32


@pytest.mark.parametrize("bits,start", [(8, 0), (8, 8), (4, 12)])
def test_radix_stages_against_numpy(bits, start):
    keys = glibc_rand(40_000)
    block, nb, bh, gh, gx, bx, out = np_stage_oracle(keys, 8, bits, start)
    got_bh = host.radix_block_histograms(keys, nb, bits, start, block)
    assert np.array_equal(got_bh, bh)
    got_gh = host.radix_reduce_histograms(got_bh, nb, 1 << bits)
    assert np.array_equal(got_gh, gh)
    got_gx = host.radix_scan_global(got_gh)
    assert np.array_equal(got_gx, gx)
    got_bx = host.radix_block_exscan(got_gx, got_bh, nb, 1 << bits)
    assert np.array_equal(got_bx, bx)
    got = host.radix_populate(got_bx, nb, bits, start, block, keys)
    assert np.array_equal(got, out)


def test_radix_stages_against_reference_golden_files(reference_dir):
    """Replays Test1..Test5 of the reference (ref hw1code/tests_q2.cpp:80-157) when its golden
    files are mounted."""
    if reference_dir is None:
        pytest.skip("reference tree not mounted")
    d = reference_dir / "hw1code" / "test_files"
    load = lambda n: np.loadtxt(d / n, dtype=np.uint64).astype(np.uint32)  # noqa: E731
    keys = load("input")
    assert np.array_equal(keys, glibc_rand(len(keys)))
    block = len(keys) // 8
    nb = (len(keys) + block - 1) // block
    bh = host.radix_block_histograms(keys, nb, 8, 0, block)
    assert np.array_equal(bh, load("blockhistograms"))
    gh = host.radix_reduce_histograms(load("blockhistograms"), nb, 256)
    assert np.array_equal(gh, load("globalhisto"))
    gx = host.radix_scan_global(load("globalhisto"))
    assert np.array_equal(gx, load("globalhistoexscan"))
    bx = host.radix_block_exscan(load("globalhistoexscan"), load("blockhistograms"), nb, 256)
    assert np.array_equal(bx, load("blockexscan"))
    out = host.radix_populate(load("blockexscan"), nb, 8, 0, block, keys)
    assert np.array_equal(out, load("sorted"))


@pytest.mark.parametrize("n", [0, 1, 7, 4096, 100_003])
@pytest.mark.parametrize("bits", [4, 8, 16])
def test_radix_sort_full(n, bits):
    rng = np.random.default_rng(n + bits)
    keys = rng.integers(0, 2**32, size=n, dtype=np.uint64).astype(np.uint32)
    want = np.sort(keys)
    assert np.array_equal(host.radix_sort(keys, bits, parallel=False), want)
    for blocks in (1, 3, 8):
        assert np.array_equal(host.radix_sort(keys, bits, num_blocks=blocks), want)
    assert np.array_equal(host.std_sort(keys), want)


def test_radix_sort_rejects_bad_digit_width():
    with pytest.raises(ValueError):
        host.radix_sort(np.arange(4, dtype=np.uint32), num_bits=5)
