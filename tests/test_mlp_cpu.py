"""MLP oracle and data-parallel semantics on the CPU (ref fpcode/neural_network.cpp)."""
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

from cme213_sp18_b200.models import mlp as M
from cme213_sp18_b200.parallel.dp_trainer import parallel_train_cpu
from cme213_sp18_b200.utils import checkpoint, grading, mnist

ROOT = Path(__file__).resolve().parent.parent


def small_problem(n=240, d=20, h=7, c=10, seed=0):
    x, lab = mnist.synthetic_mnist(n, seed=seed, image_size=d, num_classes=c, pixel_max=1)
    nn = M.NeuralNetwork([d, h, c])
    return nn, x, lab


def test_init_matches_reference_rule():
    nn = M.NeuralNetwork([784, 100, 10])
    assert nn.W[0].shape == (100, 784) and nn.W[1].shape == (10, 100)
    assert float(nn.b[0].abs().max()) == 0.0
    assert 0.009 < float(nn.W[0].std()) < 0.011
    again = M.NeuralNetwork([784, 100, 10])
    assert torch.equal(nn.W[0], again.W[0])           # same seed on every rank: no broadcast needed


def test_gradcheck_real_threshold():
    nn, x, lab = small_problem(n=12, d=6, h=5, c=10)
    nn.W[0] *= 30
    nn.W[1] *= 30
    y = M.label_to_y(lab, 10)
    cache = M.feedforward(nn, x)
    g = M.backprop(nn, y, 1e-3, cache)
    num = M.numgrad(nn, x.double(), y, 1e-3)
    assert M.gradcheck(g, num, threshold=1e-6)
    g.dW[0][0, 0] += 1e-3
    assert not M.gradcheck(g, num, threshold=1e-6)


def test_loss_and_softmax():
    nn, x, lab = small_problem()
    c = M.feedforward(nn, x)
    assert torch.allclose(c.yc.sum(dim=1), torch.ones(x.shape[0], dtype=M.F64))
    big = torch.tensor([[1000.0, 0.0]], dtype=M.F64)
    assert torch.isfinite(M.softmax(big)).all()        # max-subtracted
    val = M.loss(nn, c.yc, M.label_to_y(lab, 10), 0.5)
    ce = torch.nn.functional.cross_entropy(torch.log(c.yc), lab.long()).item()
    l2 = sum((w * w).sum().item() for w in nn.W)
    assert abs(val - (ce + 0.25 * l2)) < 1e-12


def test_serial_train_learns_and_handles_short_last_batch():
    nn, x, lab = small_problem(n=250, d=30, h=16)
    before = M.precision(M.predict(nn, x), lab)
    out = M.train(nn, x, lab, learning_rate=2.0, reg=1e-4, epochs=60, batch_size=80)
    assert out["iterations"] == 60 * 4                 # 250/80 -> 3 full + 1 short batch
    after = M.precision(M.predict(nn, x), lab)
    assert after > max(0.9, before)


@pytest.mark.parametrize("world", [1, 2, 4])
def test_fake_ranks_equal_serial_when_batch_divides(world):
    nn_s, x, lab = small_problem(n=240)
    nn_p = nn_s.clone()
    M.train(nn_s, x, lab, learning_rate=0.1, reg=1e-3, epochs=2, batch_size=80)
    parallel_train_cpu(nn_p, x, lab, 0.1, 1e-3, 2, 80, fake_ranks=world)
    rep = grading.check_nn_errors(nn_s, nn_p, filename="/tmp/cme_nnerr.txt", log=lambda *_: None)
    assert rep["passed"], rep


def test_remainder_drop_rule():
    # batch 50 over 4 ranks: in_proc = 12, two samples of every minibatch are never used (Q1)
    nn_p, x, lab = small_problem(n=100)
    nn_ref = nn_p.clone()
    parallel_train_cpu(nn_p, x, lab, 0.1, 0.0, 1, 50, fake_ranks=4)
    y = M.label_to_y(lab, 10)
    for start in (0, 50):
        used = slice(start, start + 48)
        g = M.backprop(nn_ref, y[used], 0.0, M.feedforward(nn_ref, x[used]))
        M.sgd_reference_step(nn_ref, g, 0.1)
    assert grading.check_nn_errors(nn_ref, nn_p, filename="/tmp/cme_nnerr.txt", log=lambda *_: None)["passed"]
    assert M.dp_in_proc(800, 54000 - 53600, 8) == 50 and M.dp_in_proc(50, 1000, 4) == 12


def test_raw_ascii_roundtrip_and_checkpoint(tmp_path):
    nn, _, _ = small_problem()
    checkpoint.save_raw_ascii(nn.W[0], tmp_path / "W0.mat")
    back = checkpoint.load_raw_ascii(tmp_path / "W0.mat")
    assert torch.equal(back, nn.W[0])                  # %.16e round-trips fp64 exactly
    first = (tmp_path / "W0.mat").read_text().splitlines()[0].split()
    assert len(first) == nn.W[0].shape[1]              # one matrix row per line, like raw_ascii
    checkpoint.save_checkpoint(tmp_path / "ck.pt", nn, epoch=3, iteration=17)
    other = M.NeuralNetwork(nn.H)
    other.W[0] += 1
    meta = checkpoint.load_checkpoint(tmp_path / "ck.pt", other)
    assert meta["epoch"] == 3 and meta["iteration"] == 17 and torch.equal(other.W[0], nn.W[0])
    assert len(checkpoint.colmajor_bytes(nn.W[1])) == nn.W[1].numel() * 8


def test_debug_dump_and_diff_log(tmp_path):
    nn, x, lab = small_problem(n=160)
    par = nn.clone()
    M.train(nn, x, lab, 0.1, 1e-3, 1, 80, print_every=1, debug=True, outdir=str(tmp_path / "CPUmats"),
            log=lambda *_: None)
    assert (tmp_path / "CPUmats" / "SequentialW0-1.mat").exists()
    dl = grading.CpuGpuDiffLog(tmp_path / "CpuGpuDiff.txt", tmp_path / "CPUmats")
    parallel_train_cpu(par, x, lab, 0.1, 1e-3, 1, 80, fake_ranks=2, debug_log=dl, print_every=1)
    dl.close()
    lines = (tmp_path / "CpuGpuDiff.txt").read_text().splitlines()
    assert lines[0].startswith("Iteration") and len(lines) == 3
    assert all(v < 1e-12 for _, vals in dl.rows for v in vals)


def test_idx_roundtrip(tmp_path):
    x, lab = mnist.synthetic_mnist(50, seed=3)
    mnist.write_idx_images(tmp_path / "train-images-idx3-ubyte", x)
    mnist.write_idx_labels(tmp_path / "train-labels-idx1-ubyte", lab)
    x2, lab2, xt, src = mnist.load_or_synthesize(tmp_path, n_train=50, n_test=10)
    assert src == "idx" and torch.equal(x, x2) and torch.equal(lab, lab2) and xt.shape == (10, 784)
    raw = (tmp_path / "train-images-idx3-ubyte").read_bytes()
    assert raw[:4] == b"\x00\x00\x08\x03"              # big-endian magic, ref mnist.cpp:9-18
    # the native reader: partial reads, bad files
    from cme213_sp18_b200.ops import host

    assert host.idx_info(tmp_path / "train-images-idx3-ubyte") == (2051, (50, 28, 28))
    part = host.idx_read(tmp_path / "train-images-idx3-ubyte", 7, 3, 784)
    assert torch.equal(torch.from_numpy(part).view(3, 784), x[7:10])
    assert torch.equal(mnist.read_idx_images(tmp_path / "train-images-idx3-ubyte", 5), x[:5])
    with pytest.raises(ValueError):
        host.idx_read(tmp_path / "train-images-idx3-ubyte", 48, 3, 784)      # past the end
    (tmp_path / "junk").write_bytes(b"not an idx file")
    with pytest.raises(ValueError):
        host.idx_info(tmp_path / "junk")
    with pytest.raises(ValueError):
        mnist.read_idx_labels(tmp_path / "train-images-idx3-ubyte")           # wrong magic


def test_save_label_format(tmp_path):
    grading.save_label(tmp_path / "Pred.txt", torch.tensor([3, 1, 4]))
    assert (tmp_path / "Pred.txt").read_text() == "314"          # no separator (Q17)
    assert (tmp_path / "Pred.txt.lines").read_text() == "3\n1\n4\n"


@pytest.mark.timeout(180)
def test_gloo_two_process_matches_serial(tmp_path):
    """Real multi-process run of the CPU schedule over gloo (world_size 2)."""
    script = tmp_path / "worker.py"
    script.write_text(f"""
import sys, torch
sys.path.insert(0, {str(ROOT)!r})
from cme213_sp18_b200.models import mlp as M
from cme213_sp18_b200.parallel.dist import init_distributed, shutdown
from cme213_sp18_b200.parallel.dp_trainer import parallel_train
from cme213_sp18_b200.utils import mnist, grading
info = init_distributed(device="cpu")
x, lab = mnist.synthetic_mnist(240, seed=0, image_size=20, pixel_max=1)
nn = M.NeuralNetwork([20, 7, 10])
ser = nn.clone()
parallel_train(nn, x, lab, 0.1, 1e-3, 2, 80, info=info)
if info.rank == 0:
    M.train(ser, x, lab, 0.1, 1e-3, 2, 80)
    rep = grading.check_nn_errors(ser, nn, filename={str(tmp_path / 'err.txt')!r}, log=lambda *_: None)
    assert rep["passed"], rep
    print("GLOO_OK")
shutdown()
""")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29533", OMP_NUM_THREADS="1")
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533", str(script)],
                         capture_output=True, text=True, env=env, timeout=170)
    assert res.returncode == 0, res.stdout + res.stderr
    assert "GLOO_OK" in res.stdout


def test_relu_and_cli_checkpoint_resume(tmp_path):
    assert torch.equal(M.relu(torch.tensor([-1.0, 0.0, 2.0])), torch.tensor([0.0, 0.0, 2.0]))
    env = dict(os.environ, PYTHONPATH=str(ROOT), OMP_NUM_THREADS="2")
    base = [sys.executable, "-m", "cme213_sp18_b200.cli.mlp", "--device", "cpu", "-n", "16", "-e", "1", "-b", "200",
            "-l", "0.1", "--train-size", "1000", "--test-size", "20", "--pixel-max", "1",
            "--data-dir", str(tmp_path / "none"), "--outdir", str(tmp_path / "Outputs")]
    r1 = subprocess.run(base + ["--save-checkpoint", str(tmp_path / "ck.pt")], capture_output=True, text=True,
                        cwd=tmp_path, env=env, timeout=200)
    assert r1.returncode == 0 and (tmp_path / "ck.pt").exists(), r1.stdout + r1.stderr
    r2 = subprocess.run(base + ["--resume", str(tmp_path / "ck.pt")], capture_output=True, text=True, cwd=tmp_path,
                        env=env, timeout=200)
    assert r2.returncode == 0 and "Resumed from" in r2.stdout, r2.stdout + r2.stderr


def test_native_cpp_serial_trainer_matches_the_torch_oracle():
    """csrc/host/mlp_cpu.cpp (C++/OpenMP, no BLAS) against models/mlp.train: same schedule, short
    last batch included, weights and the printed losses equal to rounding
    (ref fpcode/neural_network.cpp:219-279)."""
    from cme213_sp18_b200.ops import host

    x, lab = mnist.synthetic_mnist(1100, seed=3, pixel_max=1)          # 800 + 300: a short last batch
    a, b = M.NeuralNetwork([784, 37, 10]), M.NeuralNetwork([784, 37, 10])
    ra = M.train(a, x, lab, 0.3, 1e-3, 3, 800, print_every=2, log=lambda *_: None)
    rb = M.train_native(b, x, lab, 0.3, 1e-3, 3, 800, print_every=2, log=lambda *_: None)
    assert ra["iterations"] == rb["iterations"] == 6
    for p, q in zip(a.W + a.b, b.W + b.b):
        assert float((p - q).abs().max() / p.abs().max()) < 1e-12
    assert len(ra["losses"]) == len(rb["losses"]) == 3
    for (i, u), (j, v) in zip(ra["losses"], rb["losses"]):
        assert i == j and abs(u - v) < 1e-12 * max(1.0, abs(u))
    probs, pred = host.mlp_forward(x.double().numpy(), b.W[0].numpy(), b.b[0].numpy(), b.W[1].numpy(), b.b[1].numpy())
    assert torch.equal(torch.from_numpy(pred).long(), M.predict(b, x))
    assert abs(float(probs.sum(axis=1).max()) - 1.0) < 1e-12
    with pytest.raises(ValueError):
        host.mlp_train(x.double().numpy(), (lab + 20).numpy(), b.W[0].numpy().copy(), b.b[0].numpy().copy(),
                       b.W[1].numpy().copy(), b.b[1].numpy().copy(), 0.1, 0.0, 1, 800)


def test_cpp_cpu_example_trains_on_the_host_library_alone():
    """examples/train_mlp_cpu.cpp: the serial trainer from a C++ program that links only
    libcme213b_host.so (no GPU, no Python)."""
    from cme213_sp18_b200 import build

    exe = build.build()["example_cpu"]
    res = subprocess.run([str(exe), "48", "4000", "20", "1.0"], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0 and "TRAIN_CPU_OK" in res.stdout, res.stdout + res.stderr
