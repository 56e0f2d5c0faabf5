"""hw1 q1 driver (ref hw1code/main_q1.cpp:59-87): even/odd sums of 30 M random values in [0,100],
serial vs OpenMP, printed as the reference prints them (uint wrap-around)."""
from __future__ import annotations

import argparse
import time

import numpy as np

from ..ops import host


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="parity", description=__doc__)
    ap.add_argument("--size", type=int, default=30_000_000)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--gpu", action="store_true")
    args = ap.parse_args(argv)
    v = np.random.default_rng(0).integers(0, 101, size=args.size, dtype=np.uint32)
    t0 = time.perf_counter()
    exact_p, mod_p = host.parity_sum(v, True, args.threads)
    tp = time.perf_counter() - t0
    print(f"Parallel\nSum Even: {mod_p[0]}\nSum Odd: {mod_p[1]}\nTime: {tp}")
    t0 = time.perf_counter()
    exact_s, mod_s = host.parity_sum(v, False)
    ts = time.perf_counter() - t0
    print(f"Serial\nSum Even: {mod_s[0]}\nSum Odd: {mod_s[1]}\nTime: {ts}")
    ok = exact_p == exact_s
    if args.gpu:
        import torch

        from ..ops import device as dev

        out = dev.parity_sum(torch.from_numpy(v.view(np.int32)).cuda()).cpu().numpy()
        ok &= (int(out[0]), int(out[1])) == exact_s
        print(f"Device\nSum Even: {int(out[0]) % 2**32}\nSum Odd: {int(out[1]) % 2**32}")
    return 0 if ok else 1


if __name__ == "__main__":
    raise SystemExit(main())
