"""Device engines of the MLP: the bf16 tensor-core engine and the fp64 parity engine.

An engine owns every device buffer of one rank for a fixed maximum local batch (the
reference's `device_cache`, ref fpcode/inc/gpu_func.h:49-101, without its unused Z1/Z2 and
without re-uploading weights every step) and exposes `step()`: forward,
backward, gradient all-reduce and SGD for one minibatch slice, all on the device.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops import device as dev
from ..parallel.dist import DistInfo, SymmetricTensor


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


def flat_layout(D: int, H: int, C: int) -> dict:
    """Offsets of the flat fp32 parameter / gradient vector [W1aug (Hp x ldx) | W2 (C x H) | b2 (C)],
    every segment padded to 4 floats; the C side computes the same (cme_mlp_layout)."""
    K = D + 1
    ldx, Hp = _round_up(K, 64), _round_up(H, 64)
    n_w1 = Hp * ldx
    o_b2 = n_w1 + _round_up(C * H, 4)
    return {"K": K, "ldx": ldx, "Hp": Hp, "ldh": Hp, "n_w1": n_w1, "o_w2": n_w1, "o_b2": o_b2,
            "n_total": o_b2 + _round_up(C, 4)}


class MLPEngineBF16:
    """bf16 operands / fp32 accumulate on tcgen05, fp32 master weights.

    Layout of the flat fp32 parameter and gradient vectors:
        [ W1aug (Hp x ldx; columns 0..D-1 weights, column D = b1, rest 0) | W2 (C x H) | b2 (C) ]
    `w1c` is the bf16 compute copy of W1aug that the forward GEMM reads (K-major).  With more
    than one rank all three live in symmetric memory: the fused all-reduce + SGD kernel reads
    every rank's gradients and writes every rank's weights through peer pointers."""

    def __init__(self, H, max_rows: int, info: DistInfo | None = None, comm: str = "fused",
                 tail: str = "tc"):
        """tail: "tc" runs layer 2 / dW2 on the tensor cores, "simt" on the CUDA-core kernels."""
        self.info = info or DistInfo(device=torch.device("cuda", torch.cuda.current_device()))
        self.tail = tail
        self.fused_tail = tail == "tc"            # dZ1 + dW2 in one pass over A1
        self.D, self.Hh, self.C = H
        if self.C != 10:
            raise ValueError("the bf16 engine is built for 10 classes")
        lay = flat_layout(self.D, self.Hh, self.C)
        self.K = lay["K"]                         # pixels + the ones column that carries b1
        self.ldx, self.Hp, self.ldh = lay["ldx"], lay["Hp"], lay["ldh"]
        self.max_rows = max_rows
        self.comm = ("nccl" if comm == "nccl" else "fused") if self.info.world > 1 else "local"
        self.n_w1, self.o_w2, self.o_b2, self.n_total = lay["n_w1"], lay["o_w2"], lay["o_b2"], lay["n_total"]
        d = self.info.device
        self.params = SymmetricTensor(self.n_total, torch.float32, self.info)
        self.grads = SymmetricTensor(self.n_total, torch.float32, self.info)
        self.w1c = SymmetricTensor(self.n_w1, torch.bfloat16, self.info)
        self.pads = SymmetricTensor(1024, torch.int32, self.info)
        self.pads_dev = self.pads.ptr_table_dev()
        self.xb = torch.zeros((max_rows, self.ldx), dtype=torch.bfloat16, device=d)
        self.a1 = torch.zeros((max_rows, self.ldh), dtype=torch.bfloat16, device=d)
        self.dz1 = torch.zeros((max_rows, self.ldh), dtype=torch.bfloat16, device=d)
        self.dz2 = torch.zeros((max_rows, 16), dtype=torch.float32, device=d)
        self.dz2b = torch.zeros((max_rows, 64), dtype=torch.bfloat16, device=d)   # bf16 copy for dW2
        self.w2c = torch.zeros((16, self.ldh), dtype=torch.bfloat16, device=d)    # bf16 W2, rows >= C zero
        self.w2t = torch.zeros((self.Hp, 64), dtype=torch.bfloat16, device=d)     # bf16 W2^T, 16 live columns
        self.pred = torch.zeros(max_rows, dtype=torch.int32, device=d)
        self.loss = torch.zeros(1, dtype=torch.float32, device=d)
        # NVSwitch multicast (multimem.ld_reduce / multimem.st) wins from 4 ranks up (measured:
        # 0.401 vs 0.434 ms/step at 8 ranks, a few percent at 4, slightly slower at 2)
        want_mc = comm == "mc" or (comm == "fused" and self.info.world >= 4)
        self.use_mc = bool(want_mc and self.info.world > 1 and self.grads.multicast_ptr
                           and self.params.multicast_ptr and self.w1c.multicast_ptr)
        if comm == "mc" and self.info.world > 1 and not self.use_mc:
            raise RuntimeError("comm='mc': symmetric memory reports no multicast address on this box")
        self._graphs = {}                         # CUDA graphs of whole steps, keyed by their arguments
        self.side_stream = torch.cuda.Stream(device=d)
        self.ev = [torch.cuda.Event() for _ in range(4)]
        # overlap the all-reduce of the first half of W1 with the second half of the backward pass
        self.overlap = self.info.world > 1 and self.comm == "fused" and self.Hh >= 1024
        if self.info.world > 1:
            dist.barrier()                        # every rank's buffers are zeroed before use

    # ---- views into the flat vectors ----
    def _views(self, flat: torch.Tensor):
        w1 = flat[: self.n_w1].view(self.Hp, self.ldx)
        w2 = flat[self.o_w2: self.o_w2 + self.C * self.Hh].view(self.C, self.Hh)
        b2 = flat[self.o_b2: self.o_b2 + self.C]
        return w1, w2, b2

    def load_params(self, nn) -> None:
        """Host model (reference layout W[i] = [H_out, H_in] fp64) -> device buffers."""
        w1, w2, b2 = self._views(self.params.tensor)
        w1.zero_()
        w1[: self.Hh, : self.D] = nn.W[0].to(w1.device, torch.float32)
        w1[: self.Hh, self.D] = nn.b[0].to(w1.device, torch.float32)
        w2.copy_(nn.W[1].to(w2.device, torch.float32))
        b2.copy_(nn.b[1].to(b2.device, torch.float32))
        self.w1c.tensor.copy_(self.params.tensor[: self.n_w1].to(torch.bfloat16))
        self._recast_w2()
        torch.cuda.synchronize()
        if self.info.world > 1:
            dist.barrier()

    def gather_master(self) -> None:
        """The fp32 master of W1 is sharded across ranks (every update range is cut into one slice
        per rank); rebuild the full copy on every rank: keep what is mine, zero the rest, then a
        sum all-reduce (x + 0 is exact)."""
        if self.info.world == 1 or self.comm == "nccl":
            return
        w1 = self.params.tensor[: self.n_w1]
        mine = torch.zeros_like(w1)
        ranges, _ = self.buckets()
        for r0, r1 in ranges:
            quads = (r1 - r0) // 4
            b = r0 + quads * self.info.rank // self.info.world * 4
            e = r0 + quads * (self.info.rank + 1) // self.info.world * 4
            b, e = min(b, self.n_w1), min(e, self.n_w1)
            if e > b:
                mine[b:e] = w1[b:e]
        dist.all_reduce(mine)
        w1.copy_(mine)

    def store_params(self, nn) -> None:
        torch.cuda.synchronize()
        self.gather_master()
        torch.cuda.synchronize()
        w1, w2, b2 = self._views(self.params.tensor)
        nn.W[0] = w1[: self.Hh, : self.D].double().cpu().contiguous()
        nn.b[0] = w1[: self.Hh, self.D].double().cpu().contiguous()
        nn.W[1] = w2.double().cpu().contiguous()
        nn.b[1] = b2.double().cpu().contiguous()

    def _recast_w2(self) -> None:
        _, w2, _ = self._views(self.params.tensor)
        dev.mlp_recast_w2(w2, self.w2c, self.w2t, self.Hh, self.C)

    # ---- one minibatch slice ----
    def stage(self, x_u8_ptr: int, rows: int) -> None:
        """uint8 pixels at x_u8_ptr (local or peer-mapped) -> bf16 rows + ones column."""
        dev.u8_to_bf16_rows(x_u8_ptr, self.xb, rows, self.D, add_ones=True)

    def forward(self, rows: int, labels: torch.Tensor | None, scale: float, want_loss: bool = False,
                want_grad: bool = True, skip_fwd1: bool = False) -> None:
        _, w2, b2 = self._views(self.params.tensor)
        _, _, gb2 = self._views(self.grads.tensor)
        if not skip_fwd1:
            dev.mlp_fwd1(self.xb, self.w1c.tensor, self.a1, rows, self.K, self.Hh)
        if self.tail == "tc":
            dev.mlp_fwd2_tc(self.a1, self.w2c, b2, labels, self.dz2 if want_grad else None,
                            self.dz2b if want_grad else None, self.loss if want_loss else None,
                            self.pred, gb2 if want_grad else None, scale, rows, self.Hh, self.C)
        else:
            dev.mlp_fwd2_softmax(self.a1, w2, b2, labels, self.dz2 if want_grad else None,
                                 self.loss if want_loss else None, self.pred,
                                 gb2 if want_grad else None, scale, rows, self.Hh, self.C)

    def backward(self, rows: int, h0: int = 0, h1: int | None = None, with_dw2: bool = True) -> None:
        """Backward for hidden units [h0, h1): dZ1 and dW1 of that column block (pointer offsets
        into the same buffers), plus dW2 of the whole layer when with_dw2."""
        h1 = self.Hh if h1 is None else h1
        _, w2, _ = self._views(self.params.tensor)
        gw1, gw2, _ = self._views(self.grads.tensor)
        hb = h1 - h0
        if self.tail == "tc":
            a1_blk, dz1_blk = self.a1[:, h0:], self.dz1[:, h0:]
            if self.fused_tail:
                # dZ1 and this column block's share of dW2 in one pass over A1
                dev.mlp_bwd_tail_tc(a1_blk, self.dz2b, self.w2t[h0:], dz1_blk, gw2[:, h0:], self.Hh,
                                    rows, hb, ld=self.ldh, C=self.C)
            else:
                dev.mlp_bwd_dz1_tc(a1_blk, self.dz2b, self.w2t[h0:], dz1_blk, rows, hb, ld=self.ldh)
            dev.mlp_dw1(dz1_blk, self.xb, gw1[h0:], rows, self.K, hb, self.ldx, ldh=self.ldh)
            if with_dw2 and not self.fused_tail:
                dev.mlp_dw2_tc(self.a1, self.dz2b, gw2, rows, self.Hh, self.C)
        else:
            assert h0 == 0 and h1 == self.Hh
            dev.mlp_bwd_elem(self.a1, self.dz2, w2, self.dz1, gw2, rows, self.Hh, self.C)
            dev.mlp_dw1(self.dz1, self.xb, gw1, rows, self.K, self.Hh, self.ldx)

    def _reduce_range(self, lr: float, reg: float, rng: tuple[int, int], ch: int) -> None:
        """Barrier, fused all-reduce + L2 + SGD + re-cast of one range of the flat vector, barrier."""
        info = self.info
        w2_begin, w2_end = self.o_w2, self.o_w2 + self.C * self.Hh
        if info.world > 1:
            dev.peer_barrier(self.pads_dev, info.rank, info.world, ch, 0)      # gradients of the range complete
        if self.use_mc:
            dev.allreduce_sgd_mc(self.grads.multicast_ptr, self.params.tensor.data_ptr(),
                                 self.params.multicast_ptr, self.w1c.multicast_ptr, info.rank,
                                 info.world, self.n_total, self.n_w1, self.ldx, self.D, w2_begin,
                                 w2_end, lr, reg, rng)
        else:
            dev.allreduce_sgd(self.grads.peer_ptrs, self.params.peer_ptrs, self.w1c.peer_ptrs,
                              info.rank, self.n_total, self.n_w1, self.ldx, self.D, w2_begin, w2_end,
                              lr, reg, rng)
        if info.world > 1:
            dev.peer_barrier(self.pads_dev, info.rank, info.world, ch + 1, 0)  # new weights visible

    def update(self, lr: float, reg: float) -> None:
        """Gradient all-reduce + L2 + SGD + bf16 re-cast (ref fpcode/neural_network.cpp:501-541)."""
        if self.comm == "nccl":
            # baseline path: library all-reduce, then the same update kernel on local pointers
            w2_begin, w2_end = self.o_w2, self.o_w2 + self.C * self.Hh
            dist.all_reduce(self.grads.tensor)
            dev.allreduce_sgd([self.grads.tensor.data_ptr()], [self.params.tensor.data_ptr()],
                              [self.w1c.tensor.data_ptr()], 0, self.n_total, self.n_w1, self.ldx,
                              self.D, w2_begin, w2_end, lr, reg)
        else:
            self._reduce_range(lr, reg, (0, self.n_total), 0)
        if self.tail == "tc":
            self._recast_w2()

    def buckets(self):
        """Update ranges of the flat vector: with overlap, W1 rows [0, Hsplit) form the first
        bucket (reduced on the side stream while the rest of the backward pass runs)."""
        if not self.overlap:
            return [(0, self.n_total)], self.Hh
        hs = (self.Hh // 2 + 255) // 256 * 256
        if hs <= 0 or hs >= self.Hh:
            return [(0, self.n_total)], self.Hh
        return [(0, hs * self.ldx), (hs * self.ldx, self.n_total)], hs

    def _step_body(self, x_u8_ptr: int, labels: torch.Tensor, rows: int, lr: float, reg: float,
                   norm: float, want_loss: bool) -> None:
        main = torch.cuda.current_stream()
        side = self.side_stream
        # gradient memset on the side stream, hidden under staging + the forward GEMM
        self.ev[0].record(main)
        side.wait_event(self.ev[0])
        with torch.cuda.stream(side):
            self.grads.tensor.zero_()
            self.ev[1].record(side)
        self.stage(x_u8_ptr, rows)
        if want_loss:
            self.loss.zero_()
        dev.mlp_fwd1(self.xb, self.w1c.tensor, self.a1, rows, self.K, self.Hh)
        main.wait_event(self.ev[1])
        self.forward(rows, labels, 1.0 / norm, want_loss=want_loss, skip_fwd1=True)
        ranges, hs = self.buckets()
        if len(ranges) == 1 or self.comm == "nccl" or self.tail != "tc":
            self.backward(rows)
            self.update(lr, reg)
            return
        # bucket 0: hidden units [0, hs) -> its all-reduce + SGD runs on the side stream while the
        # main stream computes the second half of the backward pass
        self.backward(rows, 0, hs, with_dw2=False)
        self.ev[2].record(main)
        with torch.cuda.stream(side):
            side.wait_event(self.ev[2])
            self._reduce_range(lr, reg, ranges[0], 0)
            self.ev[3].record(side)
        self.backward(rows, hs, self.Hh, with_dw2=True)
        self._reduce_range(lr, reg, ranges[1], 2)
        main.wait_event(self.ev[3])
        self._recast_w2()

    def step(self, x_u8_ptr: int, labels: torch.Tensor, rows: int, lr: float, reg: float,
             norm: float, want_loss: bool = False, graph: bool = False) -> None:
        """One training step on `rows` local samples; `norm` = global sample count of the
        minibatch (in_proc * P, ref fpcode/neural_network.cpp:330-334).  With want_loss the summed
        cross-entropy of this step is left in `self.loss`.

        graph=True replays the step as one CUDA graph: the first call with a given argument set
        runs eagerly and records the launch sequence (staging, memset, the tensor-core kernels,
        the flag barriers with their device-side epochs and the fused all-reduce), later calls
        submit the whole step with a single launch.  Every pointer is part of the key, so callers
        should cycle through a small set of staging buffers."""
        if rows > self.max_rows:
            raise ValueError(f"step: {rows} rows exceed the engine capacity {self.max_rows}")
        if not graph or self.comm == "nccl":
            self._step_body(x_u8_ptr, labels, rows, lr, reg, norm, want_loss)
            return
        key = (x_u8_ptr, labels.data_ptr(), rows, lr, reg, norm, want_loss)
        entry = self._graphs.get(key)
        if entry is not None:
            entry[0].replay()
            dev._count(entry[1])
            return
        self._step_body(x_u8_ptr, labels, rows, lr, reg, norm, want_loss)     # the real step
        before = dev.launches()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._step_body(x_u8_ptr, labels, rows, lr, reg, norm, want_loss)  # recorded, not run
        n = dev.launches() - before
        dev._count(-n)
        self._graphs[key] = (g, n)

    def predict(self, x_u8: torch.Tensor) -> torch.Tensor:
        """Arg-max classes for [n, D] uint8 images on the device (chunks of max_rows)."""
        out = torch.empty(x_u8.shape[0], dtype=torch.int32, device=self.info.device)
        for lo in range(0, x_u8.shape[0], self.max_rows):
            hi = min(x_u8.shape[0], lo + self.max_rows)
            chunk = x_u8[lo:hi].contiguous().to(self.info.device)
            self.stage(chunk.data_ptr(), hi - lo)
            self.forward(hi - lo, None, 0.0, want_grad=False)
            out[lo:hi] = self.pred[: hi - lo]
        return out


class MLPEngineFP64:
    """fp64 parity engine on this repo's SIMT kernels (grade modes, 1e-7 vs the serial run).

    Buffers are column-major like the reference's device_cache; a column-major M x N matrix is
    held as a contiguous torch tensor of shape [N, M]."""

    def __init__(self, H, max_rows: int, info: DistInfo | None = None):
        self.info = info or DistInfo(device=torch.device("cuda", torch.cuda.current_device()))
        self.D, self.Hh, self.C = H
        self.max_rows = max_rows
        d, f = self.info.device, torch.float64
        D, Hh, C = self.D, self.Hh, self.C
        self.sizes = [Hh * D, Hh, C * Hh, C]
        self.offs = [0, Hh * D, Hh * D + Hh, Hh * D + Hh + C * Hh]
        n = sum(self.sizes)
        self.params = torch.zeros(n, dtype=f, device=d)
        self.grads = torch.zeros(n, dtype=f, device=d)
        self.x = torch.zeros((max_rows, D), dtype=f, device=d)
        self.a1 = torch.zeros((max_rows, Hh), dtype=f, device=d)
        self.da1 = torch.zeros((max_rows, Hh), dtype=f, device=d)
        self.dz1 = torch.zeros((max_rows, Hh), dtype=f, device=d)
        self.z2 = torch.zeros((max_rows, C), dtype=f, device=d)
        self.diff = torch.zeros((max_rows, C), dtype=f, device=d)
        self.loss = torch.zeros(1, dtype=f, device=d)

    def _seg(self, flat, i):
        return flat[self.offs[i]: self.offs[i] + self.sizes[i]]

    def load_params(self, nn) -> None:
        d = self.info.device
        # column-major H x D == contiguous [D, H] == W.T
        self._seg(self.params, 0).copy_(nn.W[0].T.contiguous().view(-1).to(d))
        self._seg(self.params, 1).copy_(nn.b[0].to(d))
        self._seg(self.params, 2).copy_(nn.W[1].T.contiguous().view(-1).to(d))
        self._seg(self.params, 3).copy_(nn.b[1].to(d))

    def store_params(self, nn) -> None:
        torch.cuda.synchronize()
        D, Hh, C = self.D, self.Hh, self.C
        nn.W[0] = self._seg(self.params, 0).view(D, Hh).T.contiguous().cpu()
        nn.b[0] = self._seg(self.params, 1).clone().cpu()
        nn.W[1] = self._seg(self.params, 2).view(Hh, C).T.contiguous().cpu()
        nn.b[1] = self._seg(self.params, 3).clone().cpu()

    def forward(self, rows: int, labels: torch.Tensor | None, scale: float, want_loss=False) -> None:
        D, Hh, C = self.D, self.Hh, self.C
        W1, b1, W2, b2 = (self._seg(self.params, i) for i in range(4))
        dev.dgemm(False, False, Hh, rows, D, 1.0, W1, Hh, self.x, D, 0.0, self.a1, Hh)
        dev.d_bias_sigmoid(self.a1, b1, Hh, rows)
        dev.dgemm(False, False, C, rows, Hh, 1.0, W2, C, self.a1, Hh, 0.0, self.z2, C)
        dev.d_bias_softmax_ce(self.z2, b2, labels, self.diff if labels is not None else None, scale,
                              self.loss if want_loss else None, C, rows)

    def backward(self, rows: int) -> None:
        D, Hh, C = self.D, self.Hh, self.C
        W2 = self._seg(self.params, 2)
        gW1, gb1, gW2, gb2 = (self._seg(self.grads, i) for i in range(4))
        dev.dgemm(False, True, C, Hh, rows, 1.0, self.diff, C, self.a1, Hh, 0.0, gW2, C)
        dev.d_row_sum(self.diff, gb2, C, rows)
        dev.dgemm(True, False, Hh, rows, C, 1.0, W2, C, self.diff, C, 0.0, self.da1, Hh)
        dev.d_dsigmoid(self.da1[:rows], self.a1[:rows], self.dz1)
        dev.dgemm(False, True, Hh, D, rows, 1.0, self.dz1, Hh, self.x, D, 0.0, gW1, Hh)
        dev.d_row_sum(self.dz1, gb1, Hh, rows)

    def update(self, lr: float, reg: float) -> None:
        if self.info.world > 1:
            dist.all_reduce(self.grads)             # parity path: library all-reduce
        for i, r in ((0, reg), (1, 0.0), (2, reg), (3, 0.0)):
            dev.d_sgd(self._seg(self.params, i), self._seg(self.grads, i), lr, r)

    def step(self, x: torch.Tensor, labels: torch.Tensor, rows: int, lr: float, reg: float,
             norm: float, want_loss: bool = False) -> None:
        self.x[:rows].copy_(x[:rows])
        self.forward(rows, labels, 1.0 / norm, want_loss=want_loss)
        self.backward(rows)
        self.update(lr, reg)

    def predict(self, x: torch.Tensor) -> torch.Tensor:
        out = torch.empty(x.shape[0], dtype=torch.int64, device=self.info.device)
        for lo in range(0, x.shape[0], self.max_rows):
            hi = min(x.shape[0], lo + self.max_rows)
            self.x[: hi - lo].copy_(x[lo:hi].to(self.info.device, torch.float64))
            self.forward(hi - lo, None, 0.0)
            out[lo:hi] = self.z2[: hi - lo].argmax(dim=1)
        return out
