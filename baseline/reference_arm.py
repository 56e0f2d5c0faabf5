"""bench.py --impl reference: the reference's own parallel_train() on the same metric and config.

The reference sources run unmodified (see build_ref.py for how they are built here).  Every
minibatch goes through the reference's stock path: MPI_Scatter of fp64 pixels and one-hot labels
from rank 0, H2D of the batch and of all weights, 13 synchronised kernel launches, D2H of the four
gradients, four MPI_Allreduce calls on host memory, SGD on the host
(ref fpcode/neural_network.cpp:401-575).  One minibatch of synthetic MNIST-shape data is reused for
every step (epochs = steps, one batch per epoch), which is the reference's own loop structure.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np


def run(args, cfg) -> dict:
    from .build_ref import build

    lib_path = build()
    if lib_path is None:
        return {"impl": "reference", "unavailable": "reference sources or nvcc not present to build baseline/_ref/libref_mlp.so"}
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        return {"impl": "reference", "unavailable": f"--gpus {args.gpus} but WORLD_SIZE={world}"} if rank == 0 else {}
    lib = C.CDLL(str(lib_path))
    sig = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_int,
           C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.ref_parallel_train.argtypes = sig
    lib.ref_init.argtypes = [C.c_int]
    if lib.ref_init(local) != 0:
        return {"impl": "reference", "unavailable": "reference harness could not initialise MPI shim / GPU"}
    H, gb = cfg["H"], cfg["global_batch"]
    D, hidden, ncls = H
    rng = np.random.default_rng(1234)
    x = rng.integers(0, 256, size=(gb, D), dtype=np.uint8)
    lab = rng.integers(0, ncls, size=gb, dtype=np.int32)
    n_here = gb if rank == 0 else 0
    sec, chk = C.c_double(0), C.c_double(0)

    def call(epochs):
        rc = lib.ref_parallel_train(x.ctypes.data, lab.ctypes.data, n_here, D, hidden, ncls, args.lr, args.reg,
                                    epochs, gb, C.byref(sec), C.byref(chk))
        if rc != 0:
            raise RuntimeError("ref_parallel_train failed")
        return sec.value

    if args.warmup > 0:
        call(args.warmup)
    seconds = call(args.steps)
    # max over ranks of the wall time (the reference times with std::chrono, ref main.cpp:239-251)
    if world > 1:
        import torch
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("gloo", rank=rank, world_size=world)
        t = torch.tensor([seconds], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        seconds = float(t.item())
        dist.destroy_process_group()
    lib.ref_finalize()
    if rank != 0:
        return {}
    in_proc = gb // world
    nparams = hidden * D + hidden + ncls * hidden + ncls
    value = args.steps * in_proc * world / seconds
    h2d = world * (in_proc * D * 8 + in_proc * ncls * 8 + nparams * 8)
    d2h = world * nparams * 8
    return {
        "metric": "MLP train samples/s", "value": value, "unit": "samples/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": seconds / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "fp64",
        "data": "synthetic", "impl": "reference",
        "config": {"model": f"mlp-{D}-{hidden}-{ncls}", "global_batch": gb, "seq_len": 1,
                   "parallelism": f"dp{world}",
                   "how": "reference sources unmodified (neural_network.cpp, gpu_func.cu, utils/common.cpp) "
                          "built for sm_100 against local stand-ins for Armadillo, MPI (shared memory) and "
                          "helper_cuda.h; timed with std::chrono around parallel_train() like fpcode/main.cpp"},
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "note": "the reference has one path only: every step stages inputs and weights through the host"},
        "gpu_launches": None, "weights_checksum": chk.value,
    }
