"""hw2 q1 driver (ref hw2code/main_q1.cu:74-224): `shift <doublings> [textfile]`.
Doubles the text `doublings` times, times the host->device copy and the host shift, then runs
every kernel variant at every size, checks the bytes exactly and prints GB/s = 2*bytes/time."""
from __future__ import annotations

import argparse
import time

import numpy as np
import torch

from ..ops import device as dev
from ..ops import host
from ..utils.text import load_text
from ..utils.timing import time_cuda_ms


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="shift", description=__doc__)
    ap.add_argument("doublings", type=int, nargs="?", default=8)
    ap.add_argument("text", nargs="?", default=None)
    ap.add_argument("--seed", type=int, default=1)
    args = ap.parse_args(argv)
    base = np.frombuffer(load_text(args.text), dtype=np.uint8)
    data = np.tile(base, 1 << args.doublings)
    shift = int(np.random.default_rng(args.seed).integers(1, 26))
    pinned = torch.from_numpy(data.copy()).pin_memory()
    d_in = torch.empty(data.size, dtype=torch.uint8, device="cuda")
    ms = time_cuda_ms(lambda: d_in.copy_(pinned, non_blocking=True), warmup=1, iters=3)
    print(f"Size (B): {data.size}")
    print(f"Host to device bandwidth: {data.size / ms / 1e6:.3f} GB/sec")
    t0 = time.perf_counter()
    want = host.shift_bytes(data, shift)
    print(f"Host shift bandwidth: {2 * data.size / (time.perf_counter() - t0) / 1e9:.3f} GB/sec")
    d_out = torch.empty_like(d_in)
    print(f"\n{'Size (B)':>14} {'char':>10} {'uint':>10} {'uint2':>10} {'uint4':>10}   (GB/sec)")
    ok = True
    for k in range(args.doublings + 1):
        n = base.size << k
        row = []
        for variant in ("char", "uint", "uint2", "uint4"):
            d_out.zero_()
            src, dst = d_in[:n], d_out[:n]
            ms = time_cuda_ms(lambda: dev.shift_bytes(src, shift, variant, dst), warmup=2, iters=5)
            good = np.array_equal(dst.cpu().numpy(), want[:n])
            ok &= good
            row.append(f"{2 * n / ms / 1e6:>10.2f}" if good else f"{'WRONG':>10}")
        print(f"{n:>14} " + " ".join(row))
    return 0 if ok else 1


if __name__ == "__main__":
    raise SystemExit(main())
