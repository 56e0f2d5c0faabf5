#!/usr/bin/env python
"""Headline benchmark: MLP data-parallel training throughput (samples/s).

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torchrun)
    python bench.py --impl reference --gpus N ...            (the reference arm, see baseline/)

Metric and config follow BASELINE.json: "MLP train samples/s", wide MLP 784-8192-10 with a
global batch of 65536 (strong scaling: the global batch is fixed as N grows, like the
reference's `-b`), bf16 tensor-core compute with fp32 master weights, synthetic MNIST-shape data,
random-init weights.  `--config small` selects the 784-100-10 / batch 3200 config.

value  = device-timed steps on data already resident in HBM (uint8), CUDA events, max over ranks
e2e    = the same steps through the public streaming path: every step copies its inputs from
         pinned host memory (H2D) and reads the step's loss back (D2H)
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

CONFIGS = {
    "wide": {"H": [784, 8192, 10], "global_batch": 65536},
    "small": {"H": [784, 100, 10], "global_batch": 3200},
    "grade": {"H": [784, 1000, 10], "global_batch": 800},      # the reference's default run (-n 1000 -b 800)
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="wide", choices=sorted(CONFIGS))
    ap.add_argument("--comm", default="fused", choices=["fused", "p2p", "mc", "nccl"],
                    help="fused = in-kernel all-reduce, multimem from 4 ranks up else peer loads/stores; "
                         "p2p / mc force one variant; nccl = library baseline")
    ap.add_argument("--data-mode", default="sharded", choices=["sharded", "owner"],
                    help="sharded: every rank holds its slices in its own HBM; owner: rank 0 holds the "
                         "minibatches and every rank's staging kernel pulls its slice over NVLink "
                         "(the reference's MPI_Scatter, fused)")
    ap.add_argument("--no-prefetch", action="store_true",
                    help="owner mode: pull every slice inside its own step's staging kernel instead of one step ahead")
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp64"],
                    help="fp64 = the grade-mode parity engine (the reference's precision)")
    ap.add_argument("--global-batch", type=int, default=0, help="override the config's global batch "
                    "(profiling only: the recorded metric uses the config's)")
    ap.add_argument("--lr", type=float, default=0.001)
    ap.add_argument("--reg", type=float, default=1e-4)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-check", action="store_true", help="skip the numerics self-check after the timed runs")
    ap.add_argument("--feeder", default="auto", choices=["auto", "native", "python"],
                    help="end-to-end input path: C++ runtime or the Python-driven feeder")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel separately")
    ap.add_argument("--breakdown", action="store_true",
                    help="also time the phases of one step separately (eager, serial, rank 0)")
    return ap.parse_args()


def _timed(fn_step, W, K, info, sampler=None):
    """W untimed + K timed steps, barrier + synchronise on both sides, CUDA events, max over ranks."""
    import torch

    from cme213_sp18_b200.parallel.dist import barrier
    from cme213_sp18_b200.utils.timing import max_over_ranks_ms

    for i in range(W):
        fn_step(i)
    torch.cuda.synchronize()
    barrier(info)
    torch.cuda.synchronize()
    if sampler:
        sampler.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        fn_step(W + i)
    e1.record()
    torch.cuda.synchronize()
    barrier(info)
    torch.cuda.synchronize()
    return max_over_ranks_ms(e0.elapsed_time(e1))


def self_check(args, H, gb, info, trainer_comm) -> dict:
    """Numerics of the distributed step, outside every timed region: learnable synthetic data
    (utils/mnist.synthetic_mnist, not random labels), a few steps at the benchmark's shape, then
      - the loss must fall,
      - every rank's replica (bf16 W1, fp32 W2 / b2) must be bit-identical,
      - the N-rank weights must match a ONE-rank run of the same global batch on rank 0
        (at N = 1: the graph-replayed overlapped step against the eager single-stream step),
      - a second run from the same start must reproduce the first bit for bit.
    The reference checks its 4-rank run against the serial run the same way
    (ref fpcode/utils/tests.cpp:51-75, fpcode/main.cpp:268-272)."""
    import torch
    import torch.distributed as dist

    from cme213_sp18_b200.models import mlp as M
    from cme213_sp18_b200.models.mlp_engine import MLPEngineBF16
    from cme213_sp18_b200.parallel.dist import DistInfo, barrier
    from cme213_sp18_b200.utils import mnist

    world, rank = info.world, info.rank
    in_proc = gb // world
    # a step size at which plain SGD from the reference's 0.01 * randn start descends at this width
    steps, lr = 10, (0.005 if H[1] >= 4096 else 0.05 if H[1] >= 512 else 0.2)
    x, lab = mnist.synthetic_mnist(gb, seed=77, pixel_max=1)
    nn0 = M.NeuralNetwork(H)
    xs = x[rank * in_proc:(rank + 1) * in_proc].to(info.device)
    ls = lab[rank * in_proc:(rank + 1) * in_proc].to(info.device)

    def run(eng, xd, ld, rows, norm, graph):
        eng.load_params(nn0)
        losses = []
        for _ in range(steps):
            eng.step(xd.data_ptr(), ld, rows, lr, args.reg, norm, want_loss=True, graph=graph)
            losses.append(eng.loss.clone())
        eng.check_health()
        return [float(v.item()) for v in losses]

    def digest(eng):
        a = eng.w1c.tensor.view(torch.int16).to(torch.int64)
        t = eng.params.tensor[eng.o_w2:].view(torch.int32).to(torch.int64)
        w = torch.arange(1, a.numel() + 1, device=a.device, dtype=torch.int64) % 1000003
        return torch.stack([a.sum(), (a * w).sum(), t.sum(), (t * w[: t.numel()]).sum()])

    eng = MLPEngineBF16(H, in_proc, info, comm=trainer_comm)
    l1 = run(eng, xs, ls, in_proc, gb // world * world, graph=not args.no_graph)
    d1 = digest(eng)
    tail1 = eng.params.tensor[eng.o_w2:].clone()
    snap = M.NeuralNetwork(H)
    eng.store_params(snap)
    run(eng, xs, ls, in_proc, gb // world * world, graph=not args.no_graph)
    d2 = digest(eng)
    reproducible = bool(torch.equal(d1, d2))
    # the one-kernel forward + slot-wise dW2 path stores every gradient in a fixed order; narrow
    # models (hidden width not a multiple of 256) still add dW2 / db2 with atomics
    deterministic_path = eng.fwd_scratch is not None and eng.dw2_scratch is not None
    rerun_diff = float((eng.params.tensor[eng.o_w2:] - tail1).abs().max() / tail1.abs().max())
    if world > 1:
        all_d = [torch.zeros_like(d1) for _ in range(world)]
        dist.all_gather(all_d, d1)
        identical = all(bool(torch.equal(all_d[0], t)) for t in all_d)
        tl = torch.tensor(l1, dtype=torch.float64, device=info.device)
        dist.all_reduce(tl)
        l1 = [float(v) for v in tl]
    else:
        identical = True
    del eng
    out = {"steps": steps, "lr": lr, "loss_first": l1[0] / gb, "loss_last": l1[-1] / gb,
           "loss_falls": l1[-1] < l1[0] - 0.02 * gb, "replicas_bit_identical": identical,
           "rerun_bit_identical": reproducible, "rerun_max_rel_diff_w2": rerun_diff,
           "fixed_order_path": deterministic_path}
    # one-rank oracle on rank 0: same global batch, eager, single stream
    if rank == 0:
        torch.cuda.empty_cache()
        solo = MLPEngineBF16(H, gb, DistInfo(device=info.device))
        solo.overlap = False
        solo.plan.overlap = 0
        run(solo, x.to(info.device), lab.to(info.device), gb, gb, graph=False)
        ref = M.NeuralNetwork(H)
        solo.store_params(ref)
        errs = {}
        for name, a, b in (("W1", snap.W[0], ref.W[0]), ("b1", snap.b[0], ref.b[0]),
                           ("W2", snap.W[1], ref.W[1]), ("b2", snap.b[1], ref.b[1])):
            # compare the training displacement, which is what the step computed
            base = {"W1": nn0.W[0], "b1": nn0.b[0], "W2": nn0.W[1], "b2": nn0.b[1]}[name]
            den = float((b - base).abs().max()) or 1.0
            errs[name] = float((a - b).abs().max()) / den
        out["vs_one_rank_rel_err"] = errs
        out["matches_one_rank"] = max(errs.values()) < 2e-2
        del solo
        torch.cuda.empty_cache()
    barrier(info)
    # bit-identical reruns are required where every sum has a fixed order; elsewhere the two runs
    # may differ by the rounding of the atomics' arrival order and no more
    rerun_ok = reproducible if deterministic_path else rerun_diff < 1e-5
    out["ok"] = bool(out["loss_falls"] and identical and rerun_ok and out.get("matches_one_rank", True))
    return out


def run_ours(args) -> dict:
    import torch
    import torch.distributed as dist

    from cme213_sp18_b200.models import mlp as M
    from cme213_sp18_b200.ops import device as dev
    from cme213_sp18_b200.parallel.dist import SymmetricTensor, barrier, init_distributed
    from cme213_sp18_b200.parallel.dp_trainer import DataParallelTrainer, NativeFeeder, StreamingFeeder
    from cme213_sp18_b200.utils.clocks import ClockSampler
    from cme213_sp18_b200.utils.timing import max_over_ranks_ms

    info = init_distributed()
    if info.world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={info.world}: launch with torchrun")
    if args.precision == "fp64":
        return run_ours_fp64(args, info)
    cfg = CONFIGS[args.config]
    H, gb = cfg["H"], args.global_batch or cfg["global_batch"]
    world, rank = info.world, info.rank
    in_proc = gb // world
    D = H[0]
    W, K = args.warmup, args.steps

    # synthetic MNIST-shape pool: `pool` distinct minibatches, revisited round-robin
    # wide: one step's activations alone exceed L2; small: enough distinct minibatches that the
    # inputs (64 x 2.5 MB = 160 MB of uint8) do not stay in the 126 MB L2 between visits
    pool = 4 if args.config == "wide" else 64
    g = torch.Generator().manual_seed(1234)
    x_host = torch.randint(0, 256, (pool * gb, D), generator=g, dtype=torch.uint8)
    lab_host = torch.randint(0, H[2], (pool * gb,), generator=g, dtype=torch.int32)

    trainer = DataParallelTrainer(H, gb, info, precision="bf16", comm=args.comm, data_mode="sharded")
    eng = trainer.engine
    nn = M.NeuralNetwork(H)
    eng.load_params(nn)
    ls = torch.stack([lab_host[p * gb + rank * in_proc: p * gb + (rank + 1) * in_proc] for p in range(pool)])
    l_dev = ls.to(info.device)
    owner = args.data_mode == "owner" and world > 1
    if owner:
        # rank 0 holds every pooled minibatch in a symmetric buffer; every rank's staging kernel reads
        # its slice from there through the peer mapping (NVLink), one byte per pixel
        sym = SymmetricTensor(pool * gb * D, torch.uint8, info, zero=False)
        if rank == 0:
            sym.tensor.copy_(x_host.reshape(-1).to(info.device))
        torch.cuda.synchronize()
        barrier(info)
        x_ptrs = [sym.peer_ptrs[0] + (p * gb + rank * in_proc) * D for p in range(pool)]
    else:
        # this rank's slices of every pooled minibatch, resident on the device as uint8
        xs = torch.stack([x_host[p * gb + rank * in_proc: p * gb + (rank + 1) * in_proc] for p in range(pool)])
        x_dev = xs.to(info.device)
        x_ptrs = [x_dev[p].data_ptr() for p in range(pool)]

    def one_step(i):
        p = i % pool
        eng.step(x_ptrs[p], l_dev[p], in_proc, args.lr, args.reg, in_proc * world, graph=not args.no_graph)
        if owner and not args.no_prefetch:
            # the next minibatch's slice starts crossing NVLink now, under this step's kernels
            eng.prefetch(x_ptrs[(i + 1) % pool], in_proc, local=rank == 0)

    one_step(0)                        # set-up: the first call records the step's graph
    sampler = ClockSampler(info.local_rank, period_ms=5).start() if rank == 0 else None
    launches0 = dev.launches()
    ms = _timed(one_step, W, K, info, sampler)
    launches = (dev.launches() - launches0) * K // (W + K)
    eng.check_health()
    value = K * gb / (ms * 1e-3)

    # ---- end to end: pinned host -> device every step, loss back every step ----
    e2e = None
    if not args.no_e2e:
        xp, lp = x_host.pin_memory(), lab_host.pin_memory()
        # the C++ loop issues the same step (streaming update next to the dW1 GEMM) without
        # Python on the per-step path; --feeder python keeps the Python-driven feeder for comparison
        native = args.feeder in ("native", "auto") and eng.comm != "nccl"
        feeder = (NativeFeeder if native else StreamingFeeder)(eng, D, graph=not args.no_graph)
        loss_pinned = torch.zeros(max(W, K), dtype=torch.float32).pin_memory()
        batches = lambda n, off: [(((off + i) % pool) * gb, in_proc) for i in range(n)]  # noqa: E731
        feeder.run(xp, lp, batches(W, 0), args.lr, args.reg, world, rank, loss_pinned)
        torch.cuda.synchronize()
        barrier(info)
        torch.cuda.synchronize()
        feeder.h2d_bytes = feeder.d2h_bytes = 0
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        feeder.run(xp, lp, batches(K, W), args.lr, args.reg, world, rank, loss_pinned)
        t1.record()
        torch.cuda.synchronize()
        barrier(info)
        torch.cuda.synchronize()
        ms_e2e = max_over_ranks_ms(t0.elapsed_time(t1))
        eng.check_health()
        e2e = {"value": K * gb / (ms_e2e * 1e-3), "unit": "samples/s", "ms_per_step": ms_e2e / K,
               "h2d_bytes_per_step": feeder.h2d_bytes * world // K,
               "d2h_bytes_per_step": feeder.d2h_bytes * world // K,
               "last_loss_per_sample": float(loss_pinned[K - 1]) / in_proc,
               "feeder": "native (cme_mlp_runtime_run)" if native else "python (StreamingFeeder)"}
    clocks = sampler.stop() if sampler else None

    # ---- exposed communication: the same per-rank work on an engine that has no peers ----
    exposed = None
    if world > 1:
        from cme213_sp18_b200.models.mlp_engine import MLPEngineBF16
        from cme213_sp18_b200.parallel.dist import DistInfo

        solo = MLPEngineBF16(H, in_proc, DistInfo(device=info.device))
        solo.load_params(nn)
        xl = x_host[rank * in_proc:(rank + 1) * in_proc].to(info.device)

        def solo_step(i):
            solo.step(xl.data_ptr(), l_dev[i % pool], in_proc, args.lr, args.reg, in_proc * world,
                      graph=not args.no_graph)

        solo_step(0)
        ms_solo = _timed(solo_step, W, K, info)
        exposed = {"ms_per_step_without_peers": ms_solo / K,
                   "exposed_allreduce_ms_per_step": max(0.0, (ms - ms_solo) / K)}
        del solo, xl

    breakdown = None
    if args.breakdown:
        names = ["stage", "zero_grad_tail", "fwd1", "fwd2_softmax", "backward(+publish)", "update(+recast)"]
        acc = {n: 0.0 for n in names}
        reps = 5
        for r in range(reps + 2):
            p = r % pool
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(names) + 1)]
            barrier(info)
            ev[0].record()
            eng.stage(x_ptrs[p], in_proc); ev[1].record()
            eng.grads.tensor[eng.o_w2:].zero_(); ev[2].record()
            dev.mlp_fwd1(eng.xb, eng.w1c.tensor, eng.a1, in_proc, eng.K, eng.Hh); ev[3].record()
            eng.forward(in_proc, l_dev[p], 1.0 / (in_proc * world), skip_fwd1=True); ev[4].record()
            eng.backward(in_proc, publish=True); ev[5].record()
            eng.update(args.lr, args.reg); ev[6].record()
            torch.cuda.synchronize()
            if r >= 2:
                for i, n in enumerate(names):
                    acc[n] += ev[i].elapsed_time(ev[i + 1]) / reps
        breakdown = {n: round(v, 4) for n, v in acc.items()}

    comm_name = ("nccl" if eng.comm == "nccl" else "fused-multimem" if eng.use_mc else "fused-p2p") if world > 1 else "local"
    trainer_comm = args.comm
    del trainer, eng
    torch.cuda.empty_cache()
    check = None if args.no_check else self_check(args, H, gb, info, trainer_comm)

    out = {
        "metric": "MLP train samples/s", "value": value, "unit": "samples/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "impl": "ours",
        "config": {"model": f"mlp-{H[0]}-{H[1]}-{H[2]}", "global_batch": gb, "seq_len": 1,
                   "parallelism": f"dp{world}", "comm": comm_name,
                   "data_mode": ("owner (rank 0 holds the minibatch; every rank pulls its slice over NVLink "
                                 + ("inside the staging kernel)" if args.no_prefetch else "one step ahead on a copy stream)"))
                   if owner else "sharded (every rank's slice resident in its own HBM)",
                   "l2": "per-step working set (activations) exceeds L2; inputs rotate over "
                         f"{pool} distinct minibatches" if args.config == "wide"
                         else f"inputs rotate over {pool} distinct minibatches ({pool * gb * D / 1e6:.0f} MB, larger than L2)",
                   "optimizer": "sgd+l2 fused with the gradient all-reduce (streaming update kernel)",
                   "launch": "eager" if args.no_graph else "one CUDA graph per step (indirect step arguments)"},
        "gpu_launches": launches, "clocks": clocks, "e2e": e2e,
        "tflops": value * M.flops_per_sample(H) / 1e12,
    }
    if exposed:
        out["communication"] = exposed
    if breakdown:
        out["breakdown_ms"] = breakdown
    if check:
        out["self_check"] = check
    if dist.is_initialized():
        dist.destroy_process_group()
    return out if rank == 0 else {}


def run_ours_fp64(args, info) -> dict:
    """The reference's own precision: every GEMM, the softmax and the update in fp64 on this repo's
    SIMT kernels (MLPEngineFP64), the gradient all-reduce + SGD as ONE kernel over peer memory.
    Same metric, same timing rules; `dtype` says fp64.  This is the arm to hold against
    `--impl reference` when precision must match (ref fpcode/neural_network.cpp:373-560)."""
    import torch
    import torch.distributed as dist

    from cme213_sp18_b200.models import mlp as M
    from cme213_sp18_b200.models.mlp_engine import MLPEngineFP64
    from cme213_sp18_b200.ops import device as dev
    from cme213_sp18_b200.parallel.dist import DistInfo, barrier
    from cme213_sp18_b200.utils import mnist
    from cme213_sp18_b200.utils.clocks import ClockSampler
    from cme213_sp18_b200.utils.timing import max_over_ranks_ms

    cfg = CONFIGS[args.config]
    H, gb = cfg["H"], args.global_batch or cfg["global_batch"]
    world, rank = info.world, info.rank
    in_proc = gb // world
    D, W, K = H[0], args.warmup, args.steps
    # enough distinct minibatches that the inputs do not sit in the 126 MB L2 between visits
    pool = max(4, min(64, (160 << 20) // (gb * D) + 1))
    g = torch.Generator().manual_seed(1234)
    x_host = torch.randint(0, 256, (pool * gb, D), generator=g, dtype=torch.uint8)
    lab_host = torch.randint(0, H[2], (pool * gb,), generator=g, dtype=torch.int32)
    sl = lambda t, p: t[p * gb + rank * in_proc: p * gb + (rank + 1) * in_proc]  # noqa: E731
    x_dev = torch.stack([sl(x_host, p) for p in range(pool)]).to(info.device)
    l_dev = torch.stack([sl(lab_host, p) for p in range(pool)]).to(info.device)

    comm = "nccl" if args.comm == "nccl" else "fused"
    eng = MLPEngineFP64(H, in_proc, info, comm=comm)
    nn = M.NeuralNetwork(H)
    eng.load_params(nn)

    def one_step(i):
        p = i % pool
        eng.step(x_dev[p], l_dev[p], in_proc, args.lr, args.reg, in_proc * world)

    one_step(0)
    sampler = ClockSampler(info.local_rank, period_ms=5).start() if rank == 0 else None
    launches0 = dev.launches()
    ms = _timed(one_step, W, K, info, sampler)
    launches = (dev.launches() - launches0) * K // (W + K)
    eng.check_health()
    value = K * gb / (ms * 1e-3)

    e2e = None
    if not args.no_e2e:
        xp, lp = x_host.pin_memory(), lab_host.pin_memory()
        xs = torch.empty((in_proc, D), dtype=torch.uint8, device=info.device)
        ls = torch.empty(in_proc, dtype=torch.int32, device=info.device)
        loss_pinned = torch.zeros(max(W, K), dtype=torch.float64).pin_memory()

        def e2e_step(i, j):
            p = i % pool
            lo = p * gb + rank * in_proc
            xs.copy_(xp[lo: lo + in_proc], non_blocking=True)
            ls.copy_(lp[lo: lo + in_proc], non_blocking=True)
            eng.loss.zero_()
            eng.step(xs, ls, in_proc, args.lr, args.reg, in_proc * world, want_loss=True)
            loss_pinned[j: j + 1].copy_(eng.loss, non_blocking=True)

        for i in range(W):
            e2e_step(i, i)
        torch.cuda.synchronize()
        barrier(info)
        torch.cuda.synchronize()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for i in range(K):
            e2e_step(W + i, i)
        t1.record()
        torch.cuda.synchronize()
        barrier(info)
        torch.cuda.synchronize()
        ms_e2e = max_over_ranks_ms(t0.elapsed_time(t1))
        eng.check_health()
        e2e = {"value": K * gb / (ms_e2e * 1e-3), "unit": "samples/s", "ms_per_step": ms_e2e / K,
               "h2d_bytes_per_step": (in_proc * D + in_proc * 4) * world, "d2h_bytes_per_step": 8 * world,
               "last_loss_per_sample": float(loss_pinned[K - 1]) / in_proc, "feeder": "python (torch copies)"}
    clocks = sampler.stop() if sampler else None

    check = None
    if not args.no_check:
        # N ranks against ONE rank on the same global batch, both in fp64: the reference's own
        # criterion (max-norm relative error below 1e-7, ref fpcode/utils/tests.cpp:51-75)
        steps, lr = 10, (0.005 if H[1] >= 4096 else 0.05 if H[1] >= 512 else 0.2)
        x, lab = mnist.synthetic_mnist(gb, seed=77, pixel_max=1)
        nn0 = M.NeuralNetwork(H)
        xr = x[rank * in_proc:(rank + 1) * in_proc].to(info.device)
        lr_ = lab[rank * in_proc:(rank + 1) * in_proc].to(info.device)

        def run(e, xd, ld, rows):
            e.load_params(nn0)
            out = []
            for _ in range(steps):
                e.loss.zero_()
                e.step(xd, ld, rows, lr, args.reg, gb // world * world, want_loss=True)
                out.append(e.loss.clone())
            e.check_health()
            return [float(v.item()) for v in out]

        l1 = run(eng, xr, lr_, in_proc)
        p1 = eng.params.clone()
        run(eng, xr, lr_, in_proc)
        rerun_same = bool(torch.equal(p1, eng.params))
        identical = True
        if world > 1:
            dg = torch.stack([p1.view(torch.int64).sum(), (p1.view(torch.int64) >> 7).sum()])
            all_d = [torch.zeros_like(dg) for _ in range(world)]
            dist.all_gather(all_d, dg)
            identical = all(bool(torch.equal(all_d[0], t)) for t in all_d)
            tl = torch.tensor(l1, dtype=torch.float64, device=info.device)
            dist.all_reduce(tl)
            l1 = [float(v) for v in tl]
        check = {"steps": steps, "lr": lr, "loss_first": l1[0] / gb, "loss_last": l1[-1] / gb,
                 "loss_falls": l1[-1] < l1[0] - 0.02 * gb, "replicas_bit_identical": identical,
                 "rerun_bit_identical": rerun_same}
        if rank == 0:
            solo = MLPEngineFP64(H, gb, DistInfo(device=info.device))
            run(solo, x.to(info.device), lab.to(info.device), gb)
            n = sum(eng.sizes)
            err = float((p1[:n] - solo.params[:n]).abs().max() / solo.params[:n].abs().max())
            check["vs_one_rank_max_rel_err"] = err
            check["matches_one_rank"] = err < 1e-7
            del solo
        barrier(info)
        check["ok"] = bool(check["loss_falls"] and identical and rerun_same and check.get("matches_one_rank", True))

    out = {
        "metric": "MLP train samples/s", "value": value, "unit": "samples/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "fp64", "data": "synthetic", "impl": "ours",
        "config": {"model": f"mlp-{H[0]}-{H[1]}-{H[2]}", "global_batch": gb, "seq_len": 1,
                   "parallelism": f"dp{world}",
                   "comm": "local" if world == 1 else ("nccl" if comm == "nccl" else "fused-p2p (one kernel: all-reduce + L2 + SGD)"),
                   "data_mode": "sharded (every rank's slice resident in its own HBM as uint8)",
                   "l2": f"inputs rotate over {pool} distinct minibatches ({pool * gb * D / 1e6:.0f} MB, larger than L2)",
                   "launch": "eager"},
        "gpu_launches": launches, "clocks": clocks, "e2e": e2e,
        "tflops": value * M.flops_per_sample(H) / 1e12,
    }
    if check:
        out["self_check"] = check
    if dist.is_initialized():
        dist.destroy_process_group()
    return out if rank == 0 else {}


def run_reference(args) -> dict:
    try:
        from baseline.reference_arm import run as ref_run
    except Exception as e:  # noqa: BLE001
        return {"impl": "reference", "unavailable": f"reference harness not importable: {e}"}
    return ref_run(args, CONFIGS[args.config])


def main() -> int:
    args = parse()
    out = run_reference(args) if args.impl == "reference" else run_ours(args)
    if out:
        print(json.dumps(out))
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
