"""Data-parallel MLP training (ref fpcode/neural_network.cpp:401-575 parallel_train).

Semantics kept from the reference: full weight replica per rank, each minibatch split evenly
with `in_proc = min(batch, N - start) // P` (the remainder is dropped), gradients
normalised by in_proc*P and summed over ranks, L2 term applied once, synchronous SGD.

What is different: nothing goes through the host.  The dataset is uploaded once (uint8) into
the owner's symmetric buffer; every step each rank's staging kernel pulls its own slice from the
owner over NVLink (the MPI_Scatter), weights stay resident on the device, and the gradient
all-reduce + SGD is one kernel over peer memory.  `device="cpu"` runs the same schedule on
torch CPU fp64 tensors with a gloo (or in-process "fake rank") all-reduce so the data-parallel
semantics are testable without GPUs.
"""
from __future__ import annotations

import time

import torch
import torch.distributed as dist

from ..models import mlp as M
from .dist import DistInfo, SymmetricTensor


# ----------------------------------------------------------------------------- CPU path
def _cpu_rank_grads(nn, Xs, ys, norm, reg, world):
    cache = M.feedforward(nn, Xs)
    return M.backprop(nn, ys, reg, cache, norm=norm, reg_scale=1.0 / world)


def parallel_train_cpu(nn, X, labels, lr, reg, epochs, batch_size, info: DistInfo | None = None,
                       fake_ranks: int | None = None, debug_log=None, print_every: int = -1) -> dict:
    """fp64 CPU schedule of parallel_train.  Either a real gloo group (`info.world > 1`, this
    process computes its own slice) or `fake_ranks` logical ranks inside one process."""
    info = info or DistInfo()
    world = fake_ranks or info.world
    N = X.shape[0]
    if info.world > 1:
        t = torch.tensor([N], dtype=torch.int64)
        dist.broadcast(t, src=0)                       # ref MPI_Bcast(N), neural_network.cpp:411
        N = int(t.item())
    Xd = X.to(M.F64)
    y = M.label_to_y(labels, nn.H[2])
    it = 0
    num_batches = (N + batch_size - 1) // batch_size
    for _epoch in range(epochs):
        for batch in range(num_batches):
            start = batch * batch_size
            in_proc = M.dp_in_proc(batch_size, N - start, world)
            if in_proc > 0:
                norm = in_proc * world
                ranks = range(world) if fake_ranks else [info.rank]
                total = None
                for r in ranks:
                    lo = start + r * in_proc
                    g = _cpu_rank_grads(nn, Xd[lo: lo + in_proc], y[lo: lo + in_proc], norm, reg, world)
                    flat = torch.cat([t.reshape(-1) for t in (g.dW[0], g.db[0], g.dW[1], g.db[1])])
                    total = flat if total is None else total + flat
                if not fake_ranks and info.world > 1:
                    dist.all_reduce(total)             # ref MPI_Allreduce x4, :501-536
                o = 0
                for p in (nn.W[0], nn.b[0], nn.W[1], nn.b[1]):
                    n = p.numel()
                    p -= lr * total[o: o + n].view_as(p)
                    o += n
            if debug_log is not None and info.is_main and M.print_flag_for(batch, it, print_every):
                debug_log.write(nn, it)
            it += 1
    return {"iterations": it}


# ----------------------------------------------------------------------------- GPU path
def as_pixels_u8(X: torch.Tensor) -> torch.Tensor:
    """The device engines train on uint8 pixels (the IDX payload).  A float tensor is accepted
    only if every value already IS an integer in 0..255; anything else (normalised inputs, say)
    would silently train on different data than the CPU path, so it is refused."""
    if X.dtype == torch.uint8:
        return X
    r = X.round()
    if not bool(((r == X) & (X >= 0) & (X <= 255)).all()):
        raise ValueError("the GPU engines take uint8 pixels (or float tensors holding exact integers in "
                         "0..255); rescale the inputs, or train on device='cpu' (fp64, any real-valued X)")
    return r.to(torch.uint8)


class DataParallelTrainer:
    """Owns the engine and the device-resident dataset for one rank.

    data_mode:
      "owner"    rank 0 holds the dataset; it is uploaded once into rank 0's symmetric buffer and
                 every rank pulls its slice of each minibatch from there (reference semantics:
                 rank 0 reads MNIST and scatters, ref fpcode/main.cpp:170-210, :460-481)
      "sharded"  every rank was given the full host dataset and uploads it itself; slices are
                 read from local memory
    """

    def __init__(self, H, batch_size: int, info: DistInfo, precision: str = "bf16",
                 comm: str = "fused", data_mode: str = "owner"):
        from ..models.mlp_engine import MLPEngineBF16, MLPEngineFP64

        self.info = info
        self.H = list(H)
        self.batch_size = batch_size
        self.precision = precision
        self.data_mode = data_mode if info.world > 1 else "sharded"
        self.max_rows = max(1, batch_size // info.world)
        if precision == "bf16":
            self.engine = MLPEngineBF16(H, self.max_rows, info, comm=comm)
        elif precision == "fp64":
            self.engine = MLPEngineFP64(H, self.max_rows, info, comm=comm)
        else:
            raise ValueError("precision must be 'bf16' or 'fp64'")
        self.data = None
        self.labels = None
        self.N = 0
        self.graph = True          # replay whole steps as CUDA graphs (one per minibatch position)

    def set_data(self, X_u8: torch.Tensor | None, labels: torch.Tensor | None) -> None:
        """Upload the dataset once.  In "owner" mode only rank 0 needs to pass tensors."""
        info = self.info
        meta = torch.zeros(2, dtype=torch.int64, device=info.device)
        if X_u8 is not None:
            meta[0], meta[1] = X_u8.shape[0], X_u8.shape[1]
        if info.world > 1 and self.data_mode == "owner":
            dist.broadcast(meta, src=0)
        self.N, D = int(meta[0].item()), int(meta[1].item())
        if D != self.H[0]:
            raise ValueError(f"data has {D} features, the model expects {self.H[0]}")
        if self.data_mode == "owner":
            self.data = SymmetricTensor(self.N * D + 32, torch.uint8, info, zero=False)
            lab = torch.zeros(self.N, dtype=torch.int32, device=info.device)
            if info.rank == 0:
                self.data.tensor[: self.N * D].copy_(X_u8.reshape(-1).to(info.device, non_blocking=True))
                lab.copy_(labels.to(info.device, torch.int32))
            dist.broadcast(lab, src=0)     # class ids, once at set-up (ref scattered 10xN one-hot per step)
            self.labels = lab
            torch.cuda.synchronize()
            dist.barrier()
            self._data_base = self.data.peer_ptrs[0]
        else:
            t = X_u8.contiguous().to(info.device)
            self.data = t
            self.labels = labels.to(info.device, torch.int32)
            self._data_base = int(t.data_ptr())

    def train(self, nn, lr: float, reg: float, epochs: int, print_every: int = -1, debug_log=None,
              log=print, debug: bool = False) -> dict:
        info, eng = self.info, self.engine
        eng.load_params(nn)
        D = self.H[0]
        N, world, bs = self.N, info.world, self.batch_size
        num_batches = (N + bs - 1) // bs
        it = 0
        losses = []
        fp64_x = None
        for _epoch in range(epochs):
            for batch in range(num_batches):
                start = batch * bs
                in_proc = M.dp_in_proc(bs, N - start, world)
                want_loss = print_every > 0 and it % print_every == 0
                if in_proc > 0:
                    lo = start + info.rank * in_proc
                    lab = self.labels[lo: lo + in_proc]
                    if self.precision == "bf16":
                        eng.step(self._data_base + lo * D, lab, in_proc, lr, reg, in_proc * world,
                                 want_loss=want_loss, graph=self.graph)
                        if self.data_mode == "owner" and world > 1:
                            # pull the next minibatch's slice from the owner while this step computes
                            nb = batch + 1 if batch + 1 < num_batches else (0 if _epoch + 1 < epochs else None)
                            if nb is not None:
                                n_in = M.dp_in_proc(bs, N - nb * bs, world)
                                if n_in > 0:
                                    eng.prefetch(self._data_base + (nb * bs + info.rank * n_in) * D, n_in,
                                                 local=info.rank == 0)
                    else:
                        src = self.data if not isinstance(self.data, SymmetricTensor) else None
                        if src is None:
                            # fp64 parity path reads the owner's buffer through a local view
                            # (single node): peer-mapped address -> pull with the copy kernel
                            from ..ops import device as dev

                            if fp64_x is None:
                                fp64_x = torch.empty(self.max_rows * D + 48, dtype=torch.uint8,
                                                     device=info.device)
                            base = self._data_base + lo * D
                            first = base - (base % 16)
                            # whole 16-byte vectors covering the slice, never past the owner's buffer
                            # (the symmetric allocation is padded by 32 bytes in set_data)
                            nbytes = (base + in_proc * D - first + 15) // 16 * 16
                            dev.peer_copy(first, fp64_x, nbytes)
                            xs = fp64_x[base % 16: base % 16 + in_proc * D].view(in_proc, D)
                        else:
                            xs = src[lo: lo + in_proc]
                        if want_loss:
                            eng.loss.zero_()
                        eng.step(xs.to(torch.float64), lab, in_proc, lr, reg, in_proc * world,
                                 want_loss=want_loss)
                    if want_loss:
                        val = float(eng.loss.item())
                        if world > 1:
                            t = torch.tensor([val], dtype=torch.float64, device=info.device)
                            dist.all_reduce(t)
                            val = float(t.item())
                        losses.append((it, val / (in_proc * world)))
                        if info.is_main:
                            log(f"Loss at iteration {it} of epoch {_epoch}/{epochs} = {losses[-1][1]}")
                if debug and M.print_flag_for(batch, it, print_every):
                    eng.store_params(nn)        # collective: the fp32 master is sharded across ranks
                    if debug_log is not None and info.is_main:
                        debug_log.write(nn, it)
                it += 1
        eng.store_params(nn)             # raises if a cross-GPU wait timed out (bf16 engine)
        return {"iterations": it, "losses": losses}


def parallel_train(nn, X, labels, learning_rate: float = 0.001, reg: float = 0.0, epochs: int = 1,
                   batch_size: int = 800, grad_check: bool = False, print_every: int = -1,
                   debug: bool = False, info: DistInfo | None = None, precision: str = "bf16",
                   comm: str = "fused", data_mode: str = "owner", debug_log=None, log=print) -> dict:
    """Same call shape as the reference's parallel_train (ref fpcode/inc/neural_network.h:44-47).
    X: [N, D] uint8 images (a float tensor must hold exact integers 0..255, see as_pixels_u8; the
    CPU schedule takes any real-valued X), labels: [N] class ids; only rank 0 needs them in
    "owner" mode.  data_mode="stream" keeps the dataset in pinned host memory and feeds every step
    through the native runtime (stream_train).  On return `nn` holds the trained weights on
    every rank."""
    del grad_check  # the analytic/numeric check lives in the serial trainer and the tests
    info = info or DistInfo(device=torch.device("cuda" if torch.cuda.is_available() else "cpu"))
    if info.device.type == "cpu":
        return parallel_train_cpu(nn, X, labels, learning_rate, reg, epochs, batch_size, info,
                                  debug_log=debug_log if debug else None, print_every=print_every)
    t0 = time.perf_counter()
    if data_mode == "stream":
        if precision != "bf16" or X is None:
            raise ValueError('data_mode="stream" is the bf16 engine fed from host memory: every rank passes X')
        Xs = as_pixels_u8(X)
        out = stream_train(nn, Xs, labels, learning_rate, reg, epochs, batch_size, info, comm=comm,
                           print_every=print_every, log=log)
        out["seconds"] = time.perf_counter() - t0
        return out
    trainer = DataParallelTrainer(nn.H, batch_size, info, precision=precision, comm=comm,
                                  data_mode=data_mode)
    Xu8 = None
    if X is not None:
        Xu8 = as_pixels_u8(X)
    trainer.set_data(Xu8, labels)
    out = trainer.train(nn, learning_rate, reg, epochs, print_every=print_every,
                        debug_log=debug_log if debug else None, log=log, debug=bool(debug))
    torch.cuda.synchronize()
    out["seconds"] = time.perf_counter() - t0
    return out


class StreamingFeeder:
    """Per-step host -> device input path for the bf16 engine.

    Every step copies this rank's slice of the minibatch (uint8 pixels + int32 labels) from
    pinned host memory into one of two device staging buffers on a copy stream, so the copy of
    step i+1 overlaps the compute of step i, and reads the previous step's loss back into pinned
    host memory.  This is the "call a user makes" path that the end-to-end benchmark times."""

    def __init__(self, engine, D: int, graph: bool = True):
        self.eng = engine
        self.D = D
        self.graph = graph
        d = engine.info.device
        self.copy_stream = torch.cuda.Stream(device=d)
        self.xbuf = [torch.empty((engine.max_rows, D), dtype=torch.uint8, device=d) for _ in range(2)]
        self.lbuf = [torch.empty(engine.max_rows, dtype=torch.int32, device=d) for _ in range(2)]
        self.ready = [torch.cuda.Event() for _ in range(2)]
        self.consumed = [torch.cuda.Event() for _ in range(2)]
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def prefetch(self, slot: int, x_pinned: torch.Tensor, lab_pinned: torch.Tensor) -> None:
        rows = x_pinned.shape[0]
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.consumed[slot])
            self.xbuf[slot][:rows].copy_(x_pinned, non_blocking=True)
            self.lbuf[slot][:rows].copy_(lab_pinned, non_blocking=True)
            self.ready[slot].record(self.copy_stream)
        self.h2d_bytes += x_pinned.numel() + lab_pinned.numel() * 4

    def run(self, x_host: torch.Tensor, lab_host: torch.Tensor, batches, lr: float, reg: float,
            world: int, rank: int, loss_out: torch.Tensor) -> int:
        """batches: list of (start, in_proc).  x_host / lab_host are pinned.  loss_out: pinned
        float32 [len(batches)] that receives the summed cross-entropy of every step."""
        eng = self.eng
        cur = torch.cuda.current_stream()
        for ev in self.consumed:
            ev.record(cur)

        def host_slice(b):
            start, in_proc = b
            lo = start + rank * in_proc
            return x_host[lo: lo + in_proc], lab_host[lo: lo + in_proc]

        if batches:
            self.prefetch(0, *host_slice(batches[0]))
        for i, (start, in_proc) in enumerate(batches):
            slot = i & 1
            if i + 1 < len(batches):
                self.prefetch(slot ^ 1, *host_slice(batches[i + 1]))
            cur.wait_event(self.ready[slot])
            eng.step(self.xbuf[slot].data_ptr(), self.lbuf[slot][:in_proc], in_proc, lr, reg,
                     in_proc * world, want_loss=True, graph=self.graph)
            self.consumed[slot].record(cur)
            loss_out[i: i + 1].copy_(eng.loss, non_blocking=True)
            self.d2h_bytes += 4
        return len(batches)


class NativeFeeder:
    """Same job as StreamingFeeder with the loop in C++ (csrc/mlp_runtime.cu): one call queues
    every step of the run (H2D on a copy stream, one CUDA graph per step, loss read-back), so
    the Python interpreter is not on the per-step path at all.  The step inside is
    cme_mlp_step_ex, the same schedule MLPEngineBF16 issues (six launches, the gradient
    all-reduce + SGD inside the dW1 GEMM)."""

    def __init__(self, engine, D: int, graph: bool = True):
        from ..ops import device as dev

        if D != engine.D:
            raise ValueError(f"feeder built for {D} features, the engine has {engine.D}")
        if engine.comm == "nccl" or engine.tail != "tc":
            raise ValueError("the native runtime drives the fused tensor-core step only")
        self.eng = engine
        self.rt = dev.MlpRuntime(engine, graphs=graph)
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def run(self, x_host: torch.Tensor, lab_host: torch.Tensor, batches, lr: float, reg: float,
            world: int, rank: int, loss_out: torch.Tensor | None) -> int:
        """batches: list of (start, in_proc) as for StreamingFeeder.run."""
        batches = [b for b in batches if b[1] > 0]
        if not batches:
            return 0
        before = self.rt.stats()
        begin = [start + rank * in_proc for start, in_proc in batches]
        rows = [in_proc for _, in_proc in batches]
        glob = [in_proc * world for _, in_proc in batches]
        n = self.rt.run(x_host, lab_host, begin, rows, glob, lr, reg, loss_out)
        after = self.rt.stats()
        self.h2d_bytes += after["h2d_bytes"] - before["h2d_bytes"]
        self.d2h_bytes += after["d2h_bytes"] - before["d2h_bytes"]
        return n


def stream_train(nn, X_u8: torch.Tensor, labels: torch.Tensor, learning_rate: float, reg: float,
                 epochs: int, batch_size: int, info: DistInfo, comm: str = "fused", native: bool = True,
                 print_every: int = -1, log=print) -> dict:
    """parallel_train for a dataset that stays in host memory: every rank holds (or maps) the
    full uint8 dataset in pinned memory and streams its slice of each minibatch to the device
    (ref fpcode/neural_network.cpp:443-567 does this with blocking copies and host scatters).
    Returns {'iterations', 'losses'} like the resident trainer; losses are the per-sample
    cross-entropy of this rank's slice at the print_every iterations."""
    from ..models.mlp_engine import MLPEngineBF16

    world, rank = info.world, info.rank
    N, D = X_u8.shape
    eng = MLPEngineBF16(nn.H, max(1, batch_size // world), info, comm=comm)
    eng.load_params(nn)
    xp = X_u8 if X_u8.is_pinned() else X_u8.contiguous().pin_memory()
    lab32 = labels.to(torch.int32)
    lp = lab32 if lab32.is_pinned() else lab32.contiguous().pin_memory()
    feeder = (NativeFeeder if native else StreamingFeeder)(eng, D)
    num_batches = (N + batch_size - 1) // batch_size
    batches = []
    for _ in range(epochs):
        for b in range(num_batches):
            start = b * batch_size
            batches.append((start, M.dp_in_proc(batch_size, N - start, world)))
    live = [b for b in batches if b[1] > 0]
    loss_host = torch.zeros(max(1, len(live)), dtype=torch.float32).pin_memory()
    feeder.run(xp, lp, live, learning_rate, reg, world, rank, loss_host)
    torch.cuda.synchronize()
    losses = []
    if print_every > 0:
        k = 0
        for it, (start, in_proc) in enumerate(batches):
            if in_proc <= 0:
                continue
            if it % print_every == 0:
                losses.append((it, float(loss_host[k]) / in_proc))
                if info.is_main:
                    log(f"Loss at iteration {it} = {losses[-1][1]}")
            k += 1
    eng.store_params(nn)
    return {"iterations": len(batches), "losses": losses, "h2d_bytes": feeder.h2d_bytes}
