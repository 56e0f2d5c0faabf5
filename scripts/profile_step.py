"""Per-kernel device times of one bf16 training step (CUDA events, median of repeats)."""
import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cme213_sp18_b200.models import mlp as M  # noqa: E402
from cme213_sp18_b200.models.mlp_engine import MLPEngineBF16  # noqa: E402
from cme213_sp18_b200.ops import device as dev  # noqa: E402
from cme213_sp18_b200.parallel.dist import DistInfo  # noqa: E402
from cme213_sp18_b200.utils.timing import time_cuda_ms  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--hidden", type=int, default=8192)
ap.add_argument("--rows", type=int, default=65536)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--out", default=None)
args = ap.parse_args()

H = [784, args.hidden, 10]
rows = args.rows
info = DistInfo(device=torch.device("cuda", 0))
eng = MLPEngineBF16(H, rows, info)
eng.load_params(M.NeuralNetwork(H))
x = torch.randint(0, 256, (rows, 784), dtype=torch.uint8, device="cuda")
lab = torch.randint(0, 10, (rows,), dtype=torch.int32, device="cuda")
_, w2, b2 = eng._views(eng.params.tensor)
gw1, gw2, gb2 = eng._views(eng.grads.tensor)

steps = {
    "stage_u8_to_bf16": lambda: eng.stage(x.data_ptr(), rows),
    "zero_grads": lambda: eng.grads.tensor.zero_(),
    "fwd1_tcgen05": lambda: dev.mlp_fwd1(eng.xb, eng.w1c.tensor, eng.a1, rows, eng.K, eng.Hh),
    "fwd2_softmax_simt": lambda: dev.mlp_fwd2_softmax(eng.a1, w2, b2, lab, eng.dz2, None, eng.pred, gb2,
                                                 1.0 / rows, rows, eng.Hh),
    "fwd2_tc_softmax": lambda: dev.mlp_fwd2_tc(eng.a1, eng.w2c, b2, lab, eng.dz2, eng.dz2b, None, eng.pred, gb2,
                                               1.0 / rows, rows, eng.Hh),
    "bwd_elem_dz1_only": lambda: dev.mlp_bwd_elem(eng.a1, eng.dz2, w2, eng.dz1, None, rows, eng.Hh),
    "bwd_dz1_tc": lambda: dev.mlp_bwd_dz1_tc(eng.a1, eng.dz2b, eng.w2t, eng.dz1, rows, eng.Hh),
    "bwd_tail_fused_dz1_dw2": lambda: dev.mlp_bwd_tail_tc(eng.a1, eng.dz2b, eng.w2t, eng.dz1, gw2, eng.Hh, rows, eng.Hh),
    "dw2_tc": lambda: dev.mlp_dw2_tc(eng.a1, eng.dz2b, gw2, rows, eng.Hh),
    "recast_w2": lambda: eng._recast_w2(),
    "bwd_elem_simt": lambda: dev.mlp_bwd_elem(eng.a1, eng.dz2, w2, eng.dz1, gw2, rows, eng.Hh),
    "dw1_tcgen05_streamk": lambda: dev.mlp_dw1(eng.dz1, eng.xb, gw1, rows, eng.K, eng.Hh, eng.ldx),
    "allreduce_sgd": lambda: eng.update(0.001, 1e-4),
    "full_step": lambda: eng.step(x.data_ptr(), lab, rows, 0.001, 1e-4, rows),
}
res = {}
for name, fn in steps.items():
    res[name] = time_cuda_ms(fn, warmup=2, iters=args.iters)
gemm_flops = 2.0 * rows * eng.K * eng.Hh
res["fwd1_tflops"] = gemm_flops / res["fwd1_tcgen05"] / 1e9
res["dw1_tflops"] = gemm_flops / res["dw1_tcgen05_streamk"] / 1e9
# cuBLAS comparison on the same shapes (library call, comparison only)
a = eng.xb[:, : eng.ldx]
w = eng.w1c.tensor.view(eng.Hp, eng.ldx)
res["cublas_fwd1_ms"] = time_cuda_ms(lambda: torch.matmul(a, w.T), warmup=2, iters=args.iters)
res["cublas_dw1_ms"] = time_cuda_ms(lambda: torch.matmul(eng.dz1.T, a), warmup=2, iters=args.iters)
res["config"] = {"H": H, "rows": rows}
print(json.dumps(res, indent=1))
if args.out:
    Path(args.out).parent.mkdir(parents=True, exist_ok=True)
    Path(args.out).write_text(json.dumps(res, indent=1))
