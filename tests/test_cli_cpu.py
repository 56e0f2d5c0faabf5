"""Command-line drivers that run without a GPU."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def run_module(mod, *args, cwd=None, timeout=300):
    env = dict(os.environ, PYTHONPATH=str(ROOT), OMP_NUM_THREADS="4")
    return subprocess.run([sys.executable, "-m", mod, *map(str, args)], capture_output=True, text=True,
                          cwd=cwd, env=env, timeout=timeout)


def test_mlp_cli_grade3_on_cpu(tmp_path):
    """Grade preset 3 end to end on the CPU schedule: serial + parallel + checkNNErrors."""
    res = run_module("cme213_sp18_b200.cli.mlp", "-g", "3", "--device", "cpu", "--train-size", "2000",
                     "--test-size", "50", "--data-dir", tmp_path / "none", "--outdir", tmp_path / "Outputs",
                     cwd=tmp_path)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "num_neuron=100, reg=0.0001, learning_rate=0.025, num_epochs=1, batch_size=800" in res.stdout
    assert "Precision on validation set for parallel training" in res.stdout
    assert "Correctness test failed" not in res.stdout
    out = tmp_path / "Outputs"
    assert (out / "NNErrors.txt").exists() and (out / "CpuGpuDiff.txt").exists()
    assert (out / "CPUmats" / "SequentialW0-0.mat").exists()
    assert len((out / "Pred_testset.txt").read_text()) == 50


def test_mlp_cli_flags_parse():
    from cme213_sp18_b200.cli.mlp import PRESETS, build_parser

    a = build_parser().parse_args("-n 64 -r 0.5 -l 0.1 -e 3 -b 16 -g 0 -p 2 -s -d".split())
    assert (a.neurons, a.reg, a.lr, a.epochs, a.batch, a.print_every, a.run_seq, a.debug) == \
        (64, 0.5, 0.1, 3, 16, 2, True, True)
    d = build_parser().parse_args([])
    assert (d.neurons, d.reg, d.lr, d.epochs, d.batch) == (1000, 1e-4, 0.001, 20, 800)   # ref main.cpp:58-67
    assert PRESETS[1]["epochs"] == 40 and PRESETS[2]["lr"] == 0.01 and PRESETS[3]["print_every"] == 1


def test_radix_and_parity_cli():
    r = run_module("cme213_sp18_b200.cli.radix", "--size", "200000")
    assert r.returncode == 0 and "Test5: PASSED" in r.stdout and "all sorts correct" in r.stdout
    p = run_module("cme213_sp18_b200.cli.parity", "--size", "1000000")
    assert p.returncode == 0 and p.stdout.count("Sum Even") == 2


def test_sim_params_file_roundtrip(tmp_path):
    from cme213_sp18_b200.utils.heat import SimParams

    f = tmp_path / "params.in"
    f.write_text("4096 4096\n1 1 \n400\n8\n")                 # the reference's params.in verbatim
    p = SimParams.from_file(f)
    assert (p.nx, p.ny, p.lx, p.ly, p.iters, p.order) == (4096, 4096, 1.0, 1.0, 400, 8)
    p.to_file(tmp_path / "p2.in")
    assert SimParams.from_file(tmp_path / "p2.in") == p
    with pytest.raises(ValueError):
        SimParams(order=3)


def test_tracer_and_cli_trace_file(tmp_path):
    """Span tracer (CPU clock here, CUDA events on a GPU) and the mlp CLI's --trace output."""
    import json
    import time

    from cme213_sp18_b200.utils.tracing import Tracer

    tr = Tracer(rank=3, device="cpu")
    with tr.span("outer", step=1):
        with tr.span("inner"):
            time.sleep(0.01)
    tot = tr.totals_ms()
    assert tot["outer"] >= tot["inner"] >= 9.0
    doc = json.loads(tr.export_chrome_trace(tmp_path / "t.json").read_text())
    ev = [e for e in doc["traceEvents"] if e["ph"] == "X"]
    assert {e["name"] for e in ev} == {"outer", "inner"} and all(e["pid"] == 3 for e in ev)
    assert Tracer(enabled=False, device="cpu").totals_ms() == {}

    r = run_module("cme213_sp18_b200.cli.mlp", "-n", "16", "-e", "2", "-b", "100", "-s", "--device", "cpu",
                   "--train-size", "400", "--test-size", "20", "--data-dir", tmp_path / "none",
                   "--outdir", tmp_path / "Outputs", "--trace", tmp_path / "run.json", cwd=tmp_path)
    assert r.returncode == 0, r.stdout + r.stderr
    names = [e["name"] for e in json.loads((tmp_path / "run.json").read_text())["traceEvents"]]
    assert names.count("serial epoch") == 2 and "parallel_train" in names


def test_bench_contract_on_a_box_without_a_gpu():
    """bench.py's reference arm must print ONE JSON line and exit 0 even where it cannot run; the
    "ours" arm must fail loudly (no silent CPU fallback for a GPU benchmark)."""
    import json
    import subprocess
    import sys

    import torch

    if torch.cuda.is_available():
        pytest.skip("this is the no-GPU contract")
    root = Path(__file__).resolve().parent.parent
    r = subprocess.run([sys.executable, str(root / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and "unavailable" in line
    r = subprocess.run([sys.executable, str(root / "bench.py"), "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode != 0                       # no GPU: refuse, do not pretend
