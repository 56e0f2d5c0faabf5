"""Where does the end-to-end step time of the small config go?  Host issue time of the native
loop, H2D copy rate alone, and the loop with and without the loss read-back."""
import sys, time, json
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from cme213_sp18_b200.models import mlp as M
from cme213_sp18_b200.models.mlp_engine import MLPEngineBF16
from cme213_sp18_b200.parallel.dist import DistInfo
from cme213_sp18_b200.parallel.dp_trainer import NativeFeeder

H, bs, pool, steps = [784, 100, 10], 3200, 8, 2000
info = DistInfo(device=torch.device("cuda", 0))
eng = MLPEngineBF16(H, bs, info)
eng.load_params(M.NeuralNetwork(H))
x = torch.randint(0, 256, (pool * bs, 784), dtype=torch.uint8).pin_memory()
lab = torch.randint(0, 10, (pool * bs,), dtype=torch.int32).pin_memory()
loss = torch.zeros(steps, dtype=torch.float32).pin_memory()
f = NativeFeeder(eng, 784)
batches = [((i % pool) * bs, bs) for i in range(steps)]
out = {}
for name, lo in (("with_loss", loss), ("no_loss", None)):
    f.run(x, lab, batches[:50], 0.001, 1e-4, 1, 0, lo)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    f.run(x, lab, batches, 0.001, 1e-4, 1, 0, lo)
    t1 = time.perf_counter()
    e1.record()
    torch.cuda.synchronize()
    out[name] = {"host_issue_us_per_step": (t1 - t0) / steps * 1e6, "device_us_per_step": e0.elapsed_time(e1) / steps * 1e3}
# H2D alone
d = torch.empty((bs, 784), dtype=torch.uint8, device="cuda")
for nbytes_rows in (bs, bs * 8):
    dd = torch.empty((nbytes_rows, 784), dtype=torch.uint8, device="cuda")
    src = x[:nbytes_rows]
    for _ in range(5):
        dd.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(200):
        dd.copy_(src, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 200 * 1e3
    out[f"h2d_{nbytes_rows * 784}_bytes"] = {"us": us, "GBps": nbytes_rows * 784 / us / 1e3}
print(json.dumps(out))
