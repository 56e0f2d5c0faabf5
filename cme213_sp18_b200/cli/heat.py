"""hw3 driver (ref hw3code/main.cu:295-412): `heat [-g] [-b] [-s] [params.in]`.

  -g global-memory kernel   -b block/loop kernel   -s shared-memory tiled kernel
Runs the CPU computation, then each selected GPU variant, prints time (ms) and GBytes/sec with
the reference's byte model, checks every grid value (border included) at 512 ULP and prints
L2Ref / LInf / L2Err.  Under torchrun (--nproc-per-node N) `-m` runs the multi-GPU slab version
with the halo exchange fused into the sweep.  With no params file the reference default
(4096 4096 / 1 1 / 400 / 8) is used."""
from __future__ import annotations

import argparse
import time
from pathlib import Path

import numpy as np
import torch

from ..models.heat import heat_device
from ..ops import host
from ..parallel.dist import init_distributed, shutdown
from ..utils.heat import SimParams
from ..utils.timing import EventTimer

ULP = 512


def report(name, got, ref, ms, p, errfile):
    ulp, _ = host.max_ulp(got, ref)
    bad = int((np.abs(got.view(np.int32).astype(np.int64) - ref.view(np.int32).astype(np.int64)) > ULP).sum()) \
        if ulp > ULP else 0
    l2ref = float(np.sqrt((ref.astype(np.float64) ** 2).sum()))
    linf = float(np.abs(got.astype(np.float64) - ref).max())
    l2err = float(np.sqrt(((got.astype(np.float64) - ref) ** 2).sum())) / l2ref
    print(f"{name:>12} {ms:>12.3f} {p.calc_bytes() / ms / 1e6:>14.2f} {p.real_bytes() / ms / 1e6:>14.2f}")
    print(f"{'':>12} L2Ref {l2ref:.6e}  LInf {linf:.3e}  L2Err {l2err:.3e}  max ULP {ulp}"
          + ("" if ulp <= ULP else f"  ** {bad} values beyond {ULP} ULP, see {errfile} **"))
    if ulp > ULP:
        idx = np.argwhere(np.abs(got.view(np.int32).astype(np.int64) - ref.view(np.int32).astype(np.int64)) > ULP)[:1000]
        Path(errfile).write_text("".join(f"Mismatch at ({i},{j}): cpu={ref[j, i]} gpu={got[j, i]}\n" for j, i in idx))
    return ulp <= ULP


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="heat", description=__doc__,
                                 formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("-g", dest="do_global", action="store_true")
    ap.add_argument("-b", dest="do_block", action="store_true")
    ap.add_argument("-s", dest="do_shared", action="store_true")
    ap.add_argument("-m", dest="do_multi", action="store_true", help="multi-GPU slabs (under torchrun)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU oracle (large grids)")
    ap.add_argument("params", nargs="?", default=None)
    for k in ("nx", "ny", "iters", "order"):
        ap.add_argument(f"--{k}", type=int, default=None)
    args = ap.parse_args(argv)
    p = SimParams.from_file(args.params) if args.params and Path(args.params).exists() else SimParams()
    for k in ("nx", "ny", "iters", "order"):
        if getattr(args, k) is not None:
            setattr(p, k, getattr(args, k))
    info = init_distributed()
    rank0 = info.is_main
    need_host_grid = not args.no_cpu or args.do_global or args.do_block or args.do_shared
    g0 = host.heat_init(p.gx, p.gy, p.dx, p.dy) if (need_host_grid and (rank0 or not args.no_cpu)) else None
    ok = True
    ref = None
    if rank0:
        print(f"Order: {p.order}, {p.nx}x{p.ny}, iters {p.iters}")
        print(f"{'':>12} {'time (ms)':>12} {'GBytes/sec':>14} {'real GB/s':>14}")
    if not args.no_cpu and rank0:
        t0 = time.perf_counter()
        ref = host.heat_cpu(g0, p.order, p.iters, p.xcfl, p.ycfl, p.bc_scale, "gpu")
        ms = (time.perf_counter() - t0) * 1e3
        print(f"{'CPU':>12} {ms:>12.3f} {p.calc_bytes() / ms / 1e6:>14.2f} {p.real_bytes() / ms / 1e6:>14.2f}")
    variants = [(args.do_global, "global", "Global"), (args.do_block, "loop", "Block"),
                (args.do_shared, "shared", "Shared")]
    if rank0 and any(on for on, _, _ in variants):
        d0 = torch.from_numpy(g0).cuda()
        for on, variant, name in variants:
            if not on:
                continue
            heat_device(d0, p, variant, iters=2)                      # warm-up
            t = EventTimer().start()
            out = heat_device(d0, p, variant)
            ms = t.stop().elapsed_ms()
            got = out.cpu().numpy()
            if ref is not None:
                ok &= report(name, got, ref, ms, p, f"{variant}Errors.txt")
            else:
                print(f"{name:>12} {ms:>12.3f} {p.calc_bytes() / ms / 1e6:>14.2f} {p.real_bytes() / ms / 1e6:>14.2f}")
    if args.do_multi:
        from ..parallel.heat_multi import HeatSlabs

        from ..utils.timing import max_over_ranks_ms

        slabs = HeatSlabs(p, info)
        reload = (lambda: slabs.load(g0)) if g0 is not None and not args.no_cpu else slabs.load_initial
        reload()
        slabs.run(2)
        reload()
        t = EventTimer().start()
        slabs.run()
        ms = max_over_ranks_ms(t.stop().elapsed_ms())
        full = slabs.gather() if not args.no_cpu else None      # collective: same decision on every rank
        if rank0:
            name = f"Slabs x{info.world}"
            if ref is not None and full is not None:
                ok &= report(name, full.cpu().numpy(), ref, ms, p, "slabErrors.txt")
            else:
                print(f"{name:>12} {ms:>12.3f} {p.calc_bytes() / ms / 1e6:>14.2f} {p.real_bytes() / ms / 1e6:>14.2f}")
    shutdown()
    return 0 if ok else 1


if __name__ == "__main__":
    raise SystemExit(main())
