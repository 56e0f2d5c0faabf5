"""hw1 q2 driver (ref hw1code/main_q2.cpp:229-312): std::sort vs serial radix vs block-parallel
radix on 4 M random keys (16-bit digits, 8 blocks), the five stage tests, and with --sweep the
threads x blocks timing grid of the reference's QUESTION6 build.  --gpu adds the device sort."""
from __future__ import annotations

import argparse
import time

import numpy as np

from ..ops import host

SIZE = 4_000_000
GRID = [1, 2, 4, 8, 12, 16, 24, 32, 40, 48]


def stage_tests(log=print) -> bool:
    """Test1..Test5 (ref hw1code/tests_q2.cpp:80-157) against a NumPy restatement."""
    rng = np.random.default_rng(0)
    keys = rng.integers(0, 2**31, size=40_000, dtype=np.uint64).astype(np.uint32)
    block, buckets = len(keys) // 8, 256
    nb = (len(keys) + block - 1) // block
    dig = keys & 0xFF
    bh = np.stack([np.bincount(dig[b * block:(b + 1) * block], minlength=buckets) for b in range(nb)]).astype(np.uint32)
    gh = bh.sum(0, dtype=np.uint32)
    gx = np.concatenate([[0], np.cumsum(gh)[:-1]]).astype(np.uint32)
    bx = (gx[None] + np.concatenate([np.zeros((1, buckets), np.uint32), np.cumsum(bh, 0)[:-1]])).astype(np.uint32)
    out = keys[np.argsort(dig, kind="stable")]
    checks = [
        ("Test1", np.array_equal(host.radix_block_histograms(keys, nb, 8, 0, block), bh.ravel())),
        ("Test2", np.array_equal(host.radix_reduce_histograms(bh.ravel(), nb, buckets), gh)),
        ("Test3", np.array_equal(host.radix_scan_global(gh), gx)),
        ("Test4", np.array_equal(host.radix_block_exscan(gx, bh.ravel(), nb, buckets), bx.ravel())),
        ("Test5", np.array_equal(host.radix_populate(bx.ravel(), nb, 8, 0, block, keys), out)),
    ]
    for name, good in checks:
        log(f"{name}: {'PASSED' if good else 'FAILED'}")
    return all(g for _, g in checks)


def timed(fn):
    t0 = time.perf_counter()
    out = fn()
    return out, time.perf_counter() - t0


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="radix", description=__doc__)
    ap.add_argument("--size", type=int, default=SIZE)
    ap.add_argument("--bits", type=int, default=16)
    ap.add_argument("--blocks", type=int, default=8)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--gpu", action="store_true")
    args = ap.parse_args(argv)
    ok = stage_tests()
    keys = np.random.default_rng(1).integers(0, 2**31 + 1, size=args.size, dtype=np.uint64).astype(np.uint32)
    want, t = timed(lambda: host.std_sort(keys))
    print(f"stl: {t}")
    got, t = timed(lambda: host.radix_sort(keys, args.bits, parallel=False))
    ok &= np.array_equal(got, want)
    print(f"serial radix: {t}")
    got, t = timed(lambda: host.radix_sort(keys, args.bits, args.blocks, True, args.threads))
    ok &= np.array_equal(got, want)
    print(f"parallel radix: {t}")
    if args.sweep:
        print("threads\\blocks " + " ".join(f"{b:>8}" for b in GRID))
        for th in GRID:
            row = [timed(lambda: host.radix_sort(keys, args.bits, b, True, th))[1] for b in GRID]
            print(f"{th:>14} " + " ".join(f"{v:>8.4f}" for v in row))
    if args.gpu:
        import torch

        from ..ops import device as dev
        from ..utils.timing import time_cuda_ms

        d = torch.from_numpy(keys.view(np.int32)).cuda()
        out = dev.radix_sort(d)
        ok &= np.array_equal(out.cpu().numpy().view(np.uint32), want)
        print(f"device radix: {time_cuda_ms(lambda: dev.radix_sort(d), 2, 5) * 1e-3}")
    print("all sorts correct" if ok else "MISMATCH")
    return 0 if ok else 1


if __name__ == "__main__":
    raise SystemExit(main())
