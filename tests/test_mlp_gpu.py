"""GPU engines of the MLP against the fp64 CPU oracle."""
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

from cme213_sp18_b200.models import mlp as M
from cme213_sp18_b200.utils import grading, mnist

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _info():
    from cme213_sp18_b200.parallel.dist import DistInfo

    return DistInfo(device=torch.device("cuda", 0))


@pytest.mark.parametrize("H,rows", [([784, 100, 10], 800), ([784, 1000, 10], 333), ([50, 64, 10], 64)])
def test_bf16_single_step_matches_oracle(H, rows):
    from cme213_sp18_b200.models.mlp_engine import MLPEngineBF16

    x, lab = mnist.synthetic_mnist(rows, seed=1, image_size=H[0], pixel_max=1)
    nn = M.NeuralNetwork(H)
    nn.W[0] *= 10
    nn.W[1] *= 10
    # make the oracle see exactly the bf16-rounded weights the tensor cores see
    nn.W[0] = nn.W[0].bfloat16().double()
    nn.b[0] = nn.b[0].bfloat16().double()
    ref = nn.clone()
    lr, reg = 0.5, 1e-3
    g = M.backprop(ref, M.label_to_y(lab, 10), reg, M.feedforward(ref, x))
    M.sgd_reference_step(ref, g, lr)

    eng = MLPEngineBF16(H, rows, _info())
    eng.load_params(nn)
    xd = x.cuda()
    eng.step(xd.data_ptr(), lab.cuda(), rows, lr, reg, rows, want_loss=True)
    got = nn.clone()
    eng.store_params(got)
    # compare the UPDATE (lr * gradient), which is what the kernels computed
    for name, w0, w_ref, w_got in (("W0", nn.W[0], ref.W[0], got.W[0]), ("b0", nn.b[0], ref.b[0], got.b[0]),
                                   ("W1", nn.W[1], ref.W[1], got.W[1]), ("b1", nn.b[1], ref.b[1], got.b[1])):
        du_ref, du_got = w_ref - w0, w_got - w0
        err = (du_ref - du_got).abs().max() / du_ref.abs().max()
        assert err < 3e-2, (name, float(err))
    want_loss = -(M.label_to_y(lab, 10) * torch.log(M.feedforward(nn, x).yc)).sum().item()
    assert abs(float(eng.loss.item()) - want_loss) < 2e-2 * want_loss
    # the bf16 compute copy tracks the fp32 master
    w1 = eng.params.tensor[: eng.n_w1]
    assert torch.equal(eng.w1c.tensor.float(), w1.bfloat16().float())


def test_bf16_training_learns():
    from cme213_sp18_b200.parallel.dp_trainer import parallel_train

    x, lab = mnist.synthetic_mnist(4000, seed=2, pixel_max=1)
    nn = M.NeuralNetwork([784, 128, 10])
    out = parallel_train(nn, x, lab, 0.5, 1e-4, 6, 400, info=_info(), print_every=5, log=lambda *_: None)
    assert out["iterations"] == 60
    acc = M.precision(M.predict(nn, x), lab)
    assert acc > 0.95, acc
    assert out["losses"][0][1] > out["losses"][-1][1]


def test_bf16_predict_matches_cpu():
    from cme213_sp18_b200.models.mlp_engine import MLPEngineBF16

    x, lab = mnist.synthetic_mnist(1000, seed=4, pixel_max=1)
    nn = M.NeuralNetwork([784, 100, 10])
    M.train(nn, x, lab, 0.5, 0.0, 3, 200)
    eng = MLPEngineBF16(nn.H, 300, _info())
    eng.load_params(nn)
    got = eng.predict(x).cpu().long()
    want = M.predict(nn, x)
    assert (got == want).double().mean() > 0.99


@pytest.mark.parametrize("hidden", [100, 37])
def test_fp64_engine_grade_tolerance(hidden, tmp_path):
    """Grade-mode contract: parallel (GPU fp64) vs serial within 1e-7 relative inf-norm
    (ref fpcode/utils/tests.cpp:51-75, grade preset 3: lr 0.025, batch 800, 1 epoch)."""
    from cme213_sp18_b200.parallel.dp_trainer import parallel_train

    x, lab = mnist.synthetic_mnist(2400 + 123, seed=5)           # raw 0..255 pixels, short last batch
    seq = M.NeuralNetwork([784, hidden, 10])
    par = seq.clone()
    M.train(seq, x, lab, 0.025, 1e-4, 1, 800)
    parallel_train(par, x, lab, 0.025, 1e-4, 1, 800, info=_info(), precision="fp64")
    rep = grading.check_nn_errors(seq, par, filename=tmp_path / "NNErrors.txt", log=lambda *_: None)
    assert rep["passed"], rep


def test_native_step_executor_matches_python_engine():
    """cme_mlp_step (one C call per step: the product path, update kernel next to the dW1 GEMM)
    against the launch sequence issued piece by piece from Python on one stream."""
    from cme213_sp18_b200.models.mlp_engine import MLPEngineBF16

    H, rows = [784, 300, 10], 1000
    x, lab = mnist.synthetic_mnist(rows, seed=8, pixel_max=1)
    nn = M.NeuralNetwork(H)
    xd, ld = x.cuda(), lab.cuda()
    a, b = MLPEngineBF16(H, rows, _info()), MLPEngineBF16(H, rows, _info())
    a.load_params(nn)
    b.load_params(nn)
    for _ in range(3):
        a._step_body(xd.data_ptr(), ld, rows, 0.5, 1e-4, rows, True)
        b.step(xd.data_ptr(), ld, rows, 0.5, 1e-4, rows, want_loss=True)
    a.check_health()
    b.check_health()
    # dW2 / db2 are still summed with atomics: compare with a tolerance
    d = (a.params.tensor - b.params.tensor).abs().max() / a.params.tensor.abs().max()
    assert float(d) < 1e-4, float(d)
    assert abs(float(a.loss.item()) - float(b.loss.item())) < 1e-3 * float(a.loss.item())


@pytest.mark.parametrize("H,rows", [([784, 1024, 10], 2048), ([784, 600, 10], 1500), ([784, 128, 10], 900)])
def test_graph_replay_serves_every_minibatch_and_learning_rate(H, rows):
    """One recorded graph per row count: the inputs, labels, lr, reg and scale are device-side
    step arguments.  Replays over different minibatches and a changing learning rate must match
    eager steps, and the graph cache must not grow."""
    from cme213_sp18_b200.models.mlp_engine import MLPEngineBF16

    x, lab = mnist.synthetic_mnist(3 * rows, seed=21, pixel_max=1)
    nn = M.NeuralNetwork(H)
    xd, ld = x.cuda(), lab.cuda()
    a, b = MLPEngineBF16(H, rows, _info()), MLPEngineBF16(H, rows, _info())
    a.load_params(nn)
    b.load_params(nn)
    for i in range(6):
        lo, lr = (i % 3) * rows, 0.5 / (1 + i)
        a.step(xd[lo:lo + rows].data_ptr(), ld[lo:lo + rows], rows, lr, 1e-4, rows, want_loss=True, graph=False)
        b.step(xd[lo:lo + rows].data_ptr(), ld[lo:lo + rows], rows, lr, 1e-4, rows, want_loss=True, graph=True)
        assert abs(float(a.loss.item()) - float(b.loss.item())) < 1e-3 * float(a.loss.item())
    b.check_health()
    assert len(b._graphs) == 1
    d = (a.params.tensor - b.params.tensor).abs().max() / a.params.tensor.abs().max()
    assert float(d) < 1e-4, float(d)
    assert torch.equal(b.w1c.tensor.float(), b.params.tensor[: b.n_w1].bfloat16().float())


def test_dw1_is_bit_reproducible_and_needs_no_zero_fill():
    """Ordered split-K: the same inputs give the same bits on every call, whatever the output
    buffer held before (the first k-chunk of every tile stores)."""
    from cme213_sp18_b200.ops import device as dev

    rows, K, H, ldx = 8192, 785, 1024, 832
    g = torch.Generator().manual_seed(5)
    dz = (torch.randn((rows, H), generator=g) * 0.1).bfloat16().cuda()
    xb = torch.zeros((rows, ldx), dtype=torch.bfloat16)
    xb[:, :K] = torch.rand((rows, K), generator=g).bfloat16()
    xb = xb.cuda()
    outs = []
    for fill in (0.0, 123.0, float("nan")):
        G = torch.full((H, ldx), fill, dtype=torch.float32, device="cuda")
        dev.mlp_dw1(dz, xb, G, rows, K, H, ldx)
        torch.cuda.synchronize()
        outs.append(G)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    want = dz.float().T @ xb.float()
    err = (outs[0] - want).abs().max() / want.abs().max()
    assert float(err) < 2e-3, float(err)
    assert float(outs[0][:, K:].abs().max()) == 0.0           # pad columns are written as zeros


def _torchrun(script: Path, nproc: int, port: int, timeout: int = 300):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    return subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                           f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
                           "--master-port", str(port), str(script)],
                          capture_output=True, text=True, env=env, timeout=timeout)


def _need_gpus(n):
    if torch.cuda.device_count() < n:
        pytest.skip(f"needs {n} CUDA devices")


@pytest.mark.multigpu
@pytest.mark.timeout(600)
@pytest.mark.parametrize("world", [2, 4, 8])
def test_multi_gpu_fused_allreduce_matches_single_gpu_and_nccl(tmp_path, world):
    """The reference's only distributed test is an N-rank run diffed against the serial run
    (ref fpcode/utils/tests.cpp:51-75, fpcode/run.sh:38-39): the same here at 2, 4 and 8 ranks,
    for the fused kernel (peer loads and NVSwitch multicast), NCCL and the fp64 engine."""
    _need_gpus(world)
    script = tmp_path / "worker.py"
    script.write_text(f"""
import sys, torch
sys.path.insert(0, {str(ROOT)!r})
from cme213_sp18_b200.models import mlp as M
from cme213_sp18_b200.parallel.dist import init_distributed, shutdown, DistInfo
from cme213_sp18_b200.parallel.dp_trainer import parallel_train
from cme213_sp18_b200.utils import mnist, grading
info = init_distributed()
x, lab = mnist.synthetic_mnist(3200, seed=0, pixel_max=1)
base = M.NeuralNetwork([784, 128, 10])
fused, nccl, fp64 = base.clone(), base.clone(), base.clone()
p2p, mc = base.clone(), base.clone()
parallel_train(fused, x if info.rank == 0 else None, lab if info.rank == 0 else None, 0.5, 1e-4, 2, 800,
               info=info, comm="fused", data_mode="owner")
parallel_train(p2p, x, lab, 0.5, 1e-4, 2, 800, info=info, comm="p2p", data_mode="sharded")
parallel_train(mc, x, lab, 0.5, 1e-4, 2, 800, info=info, comm="mc", data_mode="sharded")
parallel_train(nccl, x, lab, 0.5, 1e-4, 2, 800, info=info, comm="nccl", data_mode="sharded")
parallel_train(fp64, x, lab, 0.5, 1e-4, 2, 800, info=info, precision="fp64", data_mode="sharded")
fp64n = base.clone()          # the same engine on the library all-reduce
parallel_train(fp64n, x, lab, 0.5, 1e-4, 2, 800, info=info, precision="fp64", comm="nccl", data_mode="sharded")
if info.rank == 0:
    ser = base.clone()
    M.train(ser, x, lab, 0.5, 1e-4, 2, 800)
    r1 = grading.check_nn_errors(ser, fp64, filename={str(tmp_path / 'e1.txt')!r}, log=lambda *_: None)
    assert r1["passed"], r1
    for a, b in zip(fp64.W + fp64.b, fp64n.W + fp64n.b):
        assert float((a - b).abs().max() / b.abs().max()) < 1e-11
    for name, got in (("fused", fused), ("p2p", p2p), ("mc", mc)):
        for a, b in zip(got.W + got.b, nccl.W + nccl.b):
            d = (a - b).abs().max() / b.abs().max()
            assert d < 2e-3, (name, float(d))
        for a, b in zip(got.W + got.b, ser.W + ser.b):
            d = (a - b).abs().max() / b.abs().max()
            assert d < 5e-2, (name, float(d))
    print("MULTI_GPU_OK")
shutdown()
""")
    res = _torchrun(script, world, 29541 + world)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "MULTI_GPU_OK" in res.stdout


@pytest.mark.multigpu
@pytest.mark.timeout(600)
@pytest.mark.parametrize("world", [2, 4, 8])
def test_multi_gpu_native_runtime_matches_python_engine(tmp_path, world):
    """The C++ loop on 2 ranks (update kernel next to the dW1 GEMM at H = 1024; peer-load and
    multicast reductions) against the Python-driven resident trainer, through parallel_train."""
    script = tmp_path / "worker.py"
    script.write_text(f"""
import sys, torch
sys.path.insert(0, {str(ROOT)!r})
from cme213_sp18_b200.models import mlp as M
from cme213_sp18_b200.parallel.dist import init_distributed, shutdown
from cme213_sp18_b200.parallel.dp_trainer import parallel_train
from cme213_sp18_b200.utils import mnist
info = init_distributed()
x, lab = mnist.synthetic_mnist(2900, seed=0, pixel_max=1)       # short last batch: 500 of 800
base = M.NeuralNetwork([784, 1024, 10])
out = {{}}
for name, kw in (("resident", dict(comm="fused", data_mode="sharded")),
                 ("stream_p2p", dict(comm="p2p", data_mode="stream")),
                 ("stream_mc", dict(comm="mc", data_mode="stream"))):
    nn = base.clone()
    r = parallel_train(nn, x, lab, 0.1, 1e-4, 2, 800, info=info, print_every=3, log=lambda *_: None, **kw)
    out[name] = (nn, r)
ref = out["resident"][0]
for name in ("stream_p2p", "stream_mc"):
    nn, r = out[name]
    for a, b in zip(nn.W + nn.b, ref.W + ref.b):
        d = (a - b).abs().max() / b.abs().max()
        assert d < 2e-3, (name, float(d))
    assert r["iterations"] == 8 and [it for it, _ in r["losses"]] == [0, 3, 6], r
    # rank-local loss of the streamed run against the global mean of the resident run
    for (_, a), (_, b) in zip(r["losses"], out["resident"][1]["losses"]):
        assert a > 0 and abs(a - b) < 0.25 * b, (name, r["losses"], out["resident"][1]["losses"])
if info.rank == 0:
    print("TWO_GPU_NATIVE_OK")
shutdown()
""")
    _need_gpus(world)
    res = _torchrun(script, world, 29561 + world)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "TWO_GPU_NATIVE_OK" in res.stdout


@pytest.mark.multigpu
@pytest.mark.timeout(300)
def test_dead_peer_is_reported_not_waited_for_forever(tmp_path):
    """A rank that skips a step must not hang the others: every cross-GPU wait of the update is
    bounded, the kernel records which wait gave up, and the host raises (the reference fails fast
    through MPI_SAFE_CALL, ref fpcode/main.cpp:26-32)."""
    _need_gpus(2)
    script = tmp_path / "worker.py"
    script.write_text(f"""
import os, sys, time, torch
sys.path.insert(0, {str(ROOT)!r})
from cme213_sp18_b200.models import mlp as M
from cme213_sp18_b200.models.mlp_engine import MLPEngineBF16
from cme213_sp18_b200.ops._lib import NativeError
from cme213_sp18_b200.parallel.dist import init_distributed
from cme213_sp18_b200.utils import mnist
info = init_distributed()
H, rows = [784, 128, 10], 256
x, lab = mnist.synthetic_mnist(rows, seed=3, pixel_max=1)
eng = MLPEngineBF16(H, rows, info, comm="p2p", timeout_ms=400)
eng.load_params(M.NeuralNetwork(H))
xd, ld = x.to(info.device), lab.to(info.device)
eng.step(xd.data_ptr(), ld, rows, 0.1, 1e-4, 2 * rows)
eng.check_health()                        # a step every rank takes part in is healthy
if info.rank == 0:
    t0 = time.time()
    eng.step(xd.data_ptr(), ld, rows, 0.1, 1e-4, 2 * rows)     # rank 1 never issues this one
    try:
        eng.check_health()
        print("NO_ERROR_RAISED")
    except NativeError as e:
        print("DEAD_PEER_REPORTED after %.2f s: %s" % (time.time() - t0, e))
else:
    time.sleep(4)
sys.stdout.flush()
os._exit(0)
""")
    res = _torchrun(script, 2, 29549)
    assert "DEAD_PEER_REPORTED" in res.stdout, res.stdout[-3000:] + res.stderr[-3000:]


def test_streaming_feeder_matches_resident_training():
    """The end-to-end path (pinned host -> device every step, loss read back every step) must
    train exactly like the device-resident path, including under CUDA-graph replay."""
    from cme213_sp18_b200.models.mlp_engine import MLPEngineBF16
    from cme213_sp18_b200.parallel.dp_trainer import StreamingFeeder

    H, bs, steps = [784, 128, 10], 500, 6
    x, lab = mnist.synthetic_mnist(bs * 3, seed=11, pixel_max=1)
    nn = M.NeuralNetwork(H)
    a, b = MLPEngineBF16(H, bs, _info()), MLPEngineBF16(H, bs, _info())
    a.load_params(nn)
    b.load_params(nn)
    xd, ld = x.cuda(), lab.cuda()
    batches = [((i % 3) * bs, bs) for i in range(steps)]
    for start, n in batches:
        a.step(xd[start:start + n].data_ptr(), ld[start:start + n], n, 0.5, 1e-4, n)
    feeder = StreamingFeeder(b, 784, graph=True)
    losses = torch.zeros(steps, dtype=torch.float32).pin_memory()
    feeder.run(x.pin_memory(), lab.pin_memory(), batches, 0.5, 1e-4, 1, 0, losses)
    torch.cuda.synchronize()
    d = (a.params.tensor - b.params.tensor).abs().max() / a.params.tensor.abs().max()
    assert float(d) < 1e-4, float(d)
    assert feeder.h2d_bytes == steps * bs * (784 + 4) and feeder.d2h_bytes == 4 * steps
    assert float(losses[0]) > float(losses[-1]) > 0


@pytest.mark.parametrize("graphs", [True, False])
def test_native_runtime_matches_streaming_feeder(graphs):
    """cme_mlp_runtime_run (the C++ minibatch loop: copy stream, graph per step, loss read-back)
    against the Python-driven feeder, with a short last batch."""
    from cme213_sp18_b200.models.mlp_engine import MLPEngineBF16
    from cme213_sp18_b200.parallel.dp_trainer import NativeFeeder, StreamingFeeder

    H, bs = [784, 128, 10], 500
    x, lab = mnist.synthetic_mnist(bs * 3 - 100, seed=12, pixel_max=1)
    nn = M.NeuralNetwork(H)
    a, b = MLPEngineBF16(H, bs, _info()), MLPEngineBF16(H, bs, _info())
    a.load_params(nn)
    b.load_params(nn)
    batches = [(0, bs), (bs, bs), (2 * bs, bs - 100)] * 2
    xp, lp = x.pin_memory(), lab.pin_memory()
    la = torch.zeros(len(batches), dtype=torch.float32).pin_memory()
    lb = torch.zeros(len(batches), dtype=torch.float32).pin_memory()
    StreamingFeeder(a, 784, graph=graphs).run(xp, lp, batches, 0.5, 1e-4, 1, 0, la)
    nf = NativeFeeder(b, 784, graph=graphs)
    nf.run(xp, lp, batches, 0.5, 1e-4, 1, 0, lb)
    torch.cuda.synchronize()
    d = (a.params.tensor - b.params.tensor).abs().max() / a.params.tensor.abs().max()
    assert float(d) < 1e-4, float(d)
    assert torch.allclose(la, lb, rtol=1e-3), (la, lb)
    rows = sum(n for _, n in batches)
    assert nf.h2d_bytes == rows * (784 + 4) and nf.d2h_bytes == 4 * len(batches)
    st = nf.rt.stats()
    assert st["steps"] == len(batches) and st["graphs"] == (2 if graphs else 0)      # one per row count
    # a second run on the same runtime reuses the recorded graphs
    nf.run(xp, lp, batches[:3], 0.5, 1e-4, 1, 0, lb)
    torch.cuda.synchronize()
    assert nf.rt.stats()["graphs"] == (2 if graphs else 0)
    nf.rt.close()


def test_stream_mode_of_parallel_train_learns():
    from cme213_sp18_b200.parallel.dist import DistInfo
    from cme213_sp18_b200.parallel.dp_trainer import parallel_train

    x, lab = mnist.synthetic_mnist(4000, seed=3, pixel_max=1)
    nn = M.NeuralNetwork([784, 64, 10])
    out = parallel_train(nn, x, lab, learning_rate=0.5, reg=1e-4, epochs=6, batch_size=500,
                         info=DistInfo(device=torch.device("cuda", 0)), data_mode="stream", print_every=8,
                         log=lambda *_: None)
    assert out["iterations"] == 48 and out["losses"][0][1] > out["losses"][-1][1]
    acc = M.precision(M.predict(nn, x.double()), lab)
    assert acc > 0.9, acc


def test_warmup_kernel_runs():
    from cme213_sp18_b200.ops import device as dev

    assert dev.warmup() == 1
