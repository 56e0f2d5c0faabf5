"""Scratch: stress the overlapped step (one GPU)."""
import sys, os
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
import torch
from cme213_sp18_b200.models import mlp as M
from cme213_sp18_b200.models.mlp_engine import MLPEngineBF16
from cme213_sp18_b200.ops import device as dev
from cme213_sp18_b200.parallel.dist import DistInfo

mode, rows, steps = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
H = [784, 8192, 10]
info = DistInfo(device=torch.device("cuda", 0))
g = torch.Generator().manual_seed(1)
x = torch.randint(0, 2, (rows, 784), generator=g, dtype=torch.uint8)
lab = torch.randint(0, 10, (rows,), generator=g, dtype=torch.int32)
xd, ld = x.cuda(), lab.cuda()
eng = MLPEngineBF16(H, rows, info)
eng.load_params(M.NeuralNetwork(H))
try:
    for i in range(steps):
        eng.step(xd.data_ptr(), ld, rows, 0.001, 1e-4, rows, want_loss="loss" in mode, graph="graph" in mode)
        if i % 50 == 0:
            torch.cuda.synchronize()
            st = dev.mlp_status(eng.plan)
            if st:
                print(mode, "status", hex(st), "at step", i, flush=True)
                break
    torch.cuda.synchronize()
    print(mode, rows, "DONE status", hex(dev.mlp_status(eng.plan)), flush=True)
except Exception as e:
    print(mode, rows, "FAILED", str(e).splitlines()[0], flush=True)
