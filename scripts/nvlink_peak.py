"""NVLink peer bandwidth as this repo's kernels see it (2 ranks under torchrun):
rank 1 pulls 256 MiB from rank 0's symmetric buffer with the copy kernel (full grid / 64 blocks)
and with a copy engine; both ranks pull from each other at once for the bidirectional figure.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/nvlink_peak.py
"""
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from cme213_sp18_b200.ops import device as dev  # noqa: E402
from cme213_sp18_b200.parallel.dist import SymmetricTensor, barrier, init_distributed, shutdown  # noqa: E402
from cme213_sp18_b200.utils.timing import max_over_ranks_ms, time_cuda_ms  # noqa: E402

info = init_distributed()
n = 256 << 20
sym = SymmetricTensor(n, torch.uint8, info, zero=False)
sym.tensor.fill_(info.rank + 1)
dst = torch.empty(n, dtype=torch.uint8, device=info.device)
torch.cuda.synchronize()
barrier(info)
peer = sym.peer_ptrs[(info.rank + 1) % info.world]
res = {}


def timed(name, fn, both):
    barrier(info)
    ms = time_cuda_ms(fn, warmup=2, iters=5) if (both or info.rank == 1) else 0.0
    torch.cuda.synchronize()
    barrier(info)
    ms = max_over_ranks_ms(ms)
    res[name] = {"ms": ms, "GB/s per direction": n / ms / 1e6}


for both in (False, True):
    tag = "both ranks pull at once" if both else "rank 1 pulls alone"
    timed(f"copy kernel, full grid ({tag})", lambda: dev.peer_copy(peer, dst, n), both)
    timed(f"copy kernel, 64 blocks ({tag})", lambda: dev.peer_copy(peer, dst, n, max_blocks=64), both)
    timed(f"copy engine ({tag})", lambda: dev.peer_memcpy(peer, dst, n), both)
assert int(dst[12345]) == (info.rank + 1) % info.world + 1
local = torch.empty(n, dtype=torch.uint8, device=info.device)
timed("local HBM copy with the same kernel (both ranks)", lambda: dev.peer_copy(sym.tensor.data_ptr(), local, n), True)
if info.is_main:
    print(json.dumps(res, indent=1))
    out = ROOT / "gpurun_out" / "final" / "nvlink_peak.json"
    out.parent.mkdir(parents=True, exist_ok=True)
    out.write_text(json.dumps(res, indent=1))
shutdown()
