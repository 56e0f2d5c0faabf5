"""Timing of the hw4 cipher pipeline and the hw1 GPU kernels (CUDA events), next to the
reference's Thrust programs rebuilt for sm_100 (baseline/_ref/hw4code, whole-program wall time:
the reference has no timers, ref hw4code/create_cipher.cu, solve_cipher.cu).

    python scripts/time_hw14.py [--out profiles/evidence_r2/hw14_timing.json]
"""
import argparse
import json
import subprocess
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from cme213_sp18_b200.models import cipher as Cm  # noqa: E402
from cme213_sp18_b200.ops import device as dev  # noqa: E402
from cme213_sp18_b200.utils.text import synthetic_english  # noqa: E402
from cme213_sp18_b200.utils.timing import time_cuda_ms  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--out", default=str(ROOT / "profiles" / "evidence_r2" / "hw14_timing.json"))
args = ap.parse_args()
res = {}

# ---- hw4: the text the reference ships when it is on the box, else synthetic English of that size
moby = ROOT / "baseline" / "_ref" / "hw4code" / "mobydick.txt"
raw = moby.read_bytes() if moby.exists() else synthetic_english(1_235_150, seed=7).encode()
res["hw4_text_bytes"] = len(raw)
created = Cm.create_cipher(raw, 8)
torch.cuda.synchronize()
res["hw4_create_ms"] = time_cuda_ms(lambda: Cm.create_cipher(raw, 8), warmup=2, iters=5)
res["hw4_solve_ms"] = time_cuda_ms(lambda: Cm.solve_cipher(created.cipher), warmup=2, iters=5)
# the individual kernels on a 64 MiB text (bandwidth)
big = torch.from_numpy(np.frombuffer((raw * (64 * 2**20 // len(raw) + 1))[: 64 * 2**20], dtype=np.uint8).copy()).cuda()
n = big.numel()
for name, fn, bytes_moved in (
    ("sanitize_compact", lambda: dev.sanitize_compact(big), 2 * n),
    ("byte_histogram", lambda: dev.byte_histogram(big), n),
    ("strided_histograms_p8", lambda: dev.strided_histograms(big, 8), n),
    ("lag_match_64", lambda: dev.lag_match_counts(big, 4, 64), 65 * n),
):
    try:
        ms = time_cuda_ms(fn, warmup=2, iters=5)
        res[f"hw4_{name}_ms"] = ms
        res[f"hw4_{name}_gbs"] = bytes_moved / ms / 1e6
    except Exception as e:  # noqa: BLE001
        res[f"hw4_{name}_error"] = str(e)[:200]

# ---- the reference's programs: whole-process wall time (context creation included), best of 3
ref4 = ROOT / "baseline" / "_ref" / "hw4code"
if (ref4 / "create_cipher").exists():
    for prog, argv in (("create_cipher", ["./create_cipher", "mobydick.txt", "8"]),
                       ("solve_cipher", ["./solve_cipher", "cipher_text.txt"])):
        best = None
        for _ in range(3):
            t0 = time.perf_counter()
            r = subprocess.run(argv, cwd=ref4, capture_output=True, text=True)
            dt = (time.perf_counter() - t0) * 1e3
            if r.returncode == 0:
                best = dt if best is None else min(best, dt)
        res[f"ref_hw4_{prog}_process_wall_ms"] = best

# ---- hw1 on the GPU
g = torch.Generator().manual_seed(3)
v = torch.randint(0, 101, (30_000_000,), generator=g, dtype=torch.int32).cuda()
ms = time_cuda_ms(lambda: dev.parity_sum(v), warmup=2, iters=5)
res["hw1_parity_sum_30M_ms"] = ms
res["hw1_parity_sum_gbs"] = v.numel() * 4 / ms / 1e6
keys = torch.randint(0, 2**31 - 1, (4_000_000,), generator=g, dtype=torch.int32).cuda()
res["hw1_radix_sort_4M_ms"] = time_cuda_ms(lambda: dev.radix_sort(keys), warmup=2, iters=5)
res["torch_sort_4M_ms"] = time_cuda_ms(lambda: torch.sort(keys), warmup=2, iters=5)
print(json.dumps(res, indent=1))
Path(args.out).parent.mkdir(parents=True, exist_ok=True)
Path(args.out).write_text(json.dumps(res, indent=1))
