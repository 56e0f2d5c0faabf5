"""Per-kernel GPU time of the small-config step (graph replay of 20 calls each, CUDA events)."""
import os, sys, json
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from cme213_sp18_b200.ops import device as dev
from cme213_sp18_b200.models import mlp as M
from cme213_sp18_b200.models.mlp_engine import MLPEngineBF16

rows = int(os.environ.get("ROWS", 3200))
H = [784, int(os.environ.get("HID", 100)), 10]
eng = MLPEngineBF16(H, rows)
eng.load_params(M.NeuralNetwork(H))
x = torch.randint(0, 256, (rows, 784), dtype=torch.uint8, device="cuda")
lab = torch.randint(0, 10, (rows,), dtype=torch.int32, device="cuda")
eng.step(x.data_ptr(), lab, rows, 0.01, 1e-4, rows, graph=False)
torch.cuda.synchronize()

def timed(fn, reps=20):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn(); torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps):
                fn()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps

G = torch.zeros((eng.Hp, eng.ldx), device="cuda")
res = {}
res["dw1_plain"] = timed(lambda: dev.mlp_dw1(eng.dz1, eng.xb, G, rows, eng.K, eng.Hh, eng.ldx))
res["fwd1"] = timed(lambda: dev.mlp_fwd1(eng.xb, eng.w1c.tensor, eng.a1, rows, eng.K, eng.Hh))
if eng.fwd_scratch is not None:
    res["fwd_fused+tail"] = timed(lambda: dev.mlp_fwd_fused(eng.xb, eng.w1c.tensor, eng.w2c, eng.params.tensor[eng.o_b2:], lab, eng.a1, eng.dz2, eng.dz2b, eng.loss, eng.pred, eng.grads.tensor[eng.o_b2:], eng.fwd_scratch, eng.max_rows, 1.0 / rows, rows, eng.K, eng.Hh, eng.C))
res["stage"] = timed(lambda: eng.stage(x.data_ptr(), rows))
res["step_graph"] = timed(lambda: eng.step(x.data_ptr(), lab, rows, 0.01, 1e-4, rows, graph=False), reps=10)
print(json.dumps({k: round(v, 2) for k, v in res.items()}))
