"""hw2 q2 driver (ref hw2code/main_q2.cu:151-233): sweep N in 2^15..2^20 x avg_edges in 2..19,
6 iterations each, GPU vs host at 100 ULP, table of GB/s with the reference's byte model.
`--nodes N --avg E` runs a single configuration (e.g. the 64 M-node target); under torchrun the
row-partitioned multi-GPU version with the fused rank all-gather is used."""
from __future__ import annotations

import argparse

import numpy as np
import torch

from ..models.pagerank import pagerank_bytes, pagerank_device
from ..ops import host
from ..parallel.dist import init_distributed, shutdown
from ..utils.timing import EventTimer

MAX_ULPS = 100


def run_one(n, avg, iters, seed, check=True):
    rp, edges, inv = host.pagerank_graph(n, avg, seed)
    r0 = np.full(n, 1.0 / n, dtype=np.float32)
    d = [torch.from_numpy(a).cuda() for a in (rp, edges.view(np.int32), r0, inv)]
    from ..ops import device as dev
    from ..utils.timing import time_cuda_ms

    ca, cb, out = (torch.empty_like(d[2]) for _ in range(3))
    # timed region = the kernels only, like the reference (ref hw2code/pagerank.cu:100-120)
    call = lambda: dev.pagerank_iterate(d[0], d[1], d[2], d[3], ca, cb, out, iters)   # noqa: E731
    ms = time_cuda_ms(call, warmup=2, iters=5)
    if ms < 0.2:
        # a run this short is shorter than the interpreter's own cost of issuing it, and the GPU would
        # be timed waiting for Python: time ten runs issued back to back and take the mean
        def ten():
            for _ in range(10):
                call()
        ms = time_cuda_ms(ten, warmup=1, iters=5) / 10
    ulp = host.max_ulp(out.cpu().numpy(), host.pagerank(rp, edges, r0, inv, iters))[0] if check else 0
    return ms, ulp


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="pagerank", description=__doc__)
    ap.add_argument("--nodes", type=int, default=None)
    ap.add_argument("--avg", type=int, default=8)
    ap.add_argument("--iters", type=int, default=6)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--no-check", action="store_true")
    ap.add_argument("--block-nodes", type=int, default=None,
                    help="source nodes per block of the column-blocked propagation (default: 16 M once the "
                         "graph has more than 24 M nodes; 0 = plain CSR)")
    ap.add_argument("--l2-fetch", type=int, default=0, choices=[0, 32, 64, 128],
                    help="L2 fetch granularity in bytes for the run (0 = leave the driver's default of 64); "
                         "measured: 32 changes nothing at 64 M nodes (51.79 vs 51.83 ms), HBM fetches 64 bytes anyway")
    args = ap.parse_args(argv)
    info = init_distributed()
    if args.l2_fetch and torch.cuda.is_available():
        from ..ops import device as dev

        was = dev.l2_fetch_granularity(args.l2_fetch)
        if info.is_main:
            print(f"L2 fetch granularity {was} -> {args.l2_fetch} bytes")
    ok = True
    if info.world > 1 or args.nodes is not None:
        from ..parallel.pagerank_multi import PageRankShards

        n = args.nodes or (1 << 20)
        sh = PageRankShards(n, args.avg, info, args.seed, block_nodes=args.block_nodes)
        sh.run(1)
        t = EventTimer().start()
        sh.run(args.iters)
        ms = t.stop().elapsed_ms()
        from ..utils.timing import max_over_ranks_ms

        ms = max_over_ranks_ms(ms)
        full = sh.gather_rank()
        if info.is_main:
            gbs = pagerank_bytes(n, args.avg, args.iters) / ms / 1e6
            msg = f"N={n} avg_edges={args.avg} ranks={info.world}: {ms:.3f} ms, {gbs:.2f} GB/s"
            if sh.blocked is not None:
                msg += f" ({sh.blocked[2]} column blocks of {sh.block_nodes} nodes)"
            if not args.no_check and n <= (1 << 24):
                rp, edges, inv = host.pagerank_graph(n, args.avg, args.seed)
                want = host.pagerank(rp, edges, np.full(n, 1.0 / n, np.float32), inv, args.iters)
                ulp = host.max_ulp(full.cpu().numpy(), want)[0]
                ok = ulp <= MAX_ULPS
                msg += f", max ULP {ulp} ({'ok' if ok else 'FAILED'})"
            print(msg)
    elif info.is_main:
        sizes = [1 << k for k in range(15, 21)]
        print("avg_edges " + " ".join(f"{n:>10}" for n in sizes) + "   (GB/s)")
        for avg in range(2, 20):
            row = []
            for n in sizes:
                ms, ulp = run_one(n, avg, args.iters, args.seed, not args.no_check)
                ok &= ulp <= MAX_ULPS
                row.append(f"{pagerank_bytes(n, avg, args.iters) / ms / 1e6:>10.2f}" if ulp <= MAX_ULPS else f"{'WRONG':>10}")
            print(f"{avg:>9} " + " ".join(row))
    shutdown()
    return 0 if ok else 1


if __name__ == "__main__":
    raise SystemExit(main())
