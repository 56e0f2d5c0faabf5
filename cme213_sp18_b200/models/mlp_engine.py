"""Device engines of the MLP: the bf16 tensor-core engine and the fp64 parity engine.

An engine owns every device buffer of one rank for a fixed maximum local batch (the
reference's `device_cache`, ref fpcode/inc/gpu_func.h:49-101, without its unused Z1/Z2 and
without re-uploading weights every step) and exposes `step()`: forward,
backward, gradient all-reduce and SGD for one minibatch slice, all on the device.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from ..ops import device as dev
from ..ops._lib import NativeError, env_flag
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
    `w1c` is the bf16 compute copy of W1aug that the forward GEMM reads (K-major).  All three
    live in symmetric memory: the streaming update kernel reads every rank's gradients and
    writes every rank's weights through peer pointers or the NVSwitch multicast address.

    A step is issued by the native executor (csrc/mlp_step.cu, cme_mlp_step_ex); the Python-
    issued sequence is kept only for the NCCL baseline (`comm="nccl"`) and the CUDA-core tail
    (`tail="simt"`)."""

    def __init__(self, H, max_rows: int, info: DistInfo | None = None, comm: str = "fused",
                 tail: str = "tc", timeout_ms: int | None = None):
        """tail: "tc" runs layer 2 / dW2 on the tensor cores, "simt" on the CUDA-core kernels.
        timeout_ms bounds every cross-GPU wait of the update kernel (default 20 s, or
        CME_PEER_TIMEOUT_MS)."""
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
        self.pads = SymmetricTensor(dev.PAD_WORDS, torch.int32, self.info)
        self.pads_dev = self.pads.ptr_table_dev()
        self.sync_scratch = torch.zeros(dev.dw1_scratch_words(), dtype=torch.int32, device=d)
        self.step_args = torch.zeros(4, dtype=torch.int64, device=d)             # CmeStepArgs (32 bytes)
        # partial logits of the one-kernel forward pass (layer 2 rides on the forward GEMM's tiles)
        fused_fwd = tail == "tc" and self.Hp % 256 == 0 and max_rows >= 256 and not env_flag("CME_NO_FWD_FUSED")
        self.fwd_scratch = torch.zeros(dev.fwd_scratch_floats(max_rows), device=d) if fused_fwd else None
        # per-CTA slots of the dW2 product, added up in CTA order (bit-reproducible gradients)
        n_dw2 = int(dev.device_lib().cme_mlp_bwd_tail_scratch_floats(self.Hh)) if torch.cuda.is_available() else 0
        self.dw2_scratch = torch.zeros(n_dw2, device=d) if (tail == "tc" and n_dw2) else None
        self.xb = torch.zeros((max_rows, self.ldx), dtype=torch.bfloat16, device=d)
        self.a1 = torch.zeros((max_rows, self.ldh), dtype=torch.bfloat16, device=d)
        self.dz1 = torch.zeros((max_rows, self.ldh), dtype=torch.bfloat16, device=d)
        self.dz2 = torch.zeros((max_rows, 16), dtype=torch.float32, device=d)
        self.dz2b = torch.zeros((max_rows, 64), dtype=torch.bfloat16, device=d)   # bf16 copy for dW2
        self.w2c = torch.zeros((16, self.ldh), dtype=torch.bfloat16, device=d)    # bf16 W2, rows >= C zero
        self.w2t = torch.zeros((self.Hp, 64), dtype=torch.bfloat16, device=d)     # bf16 W2^T, 16 live columns
        self.pred = torch.zeros(max_rows, dtype=torch.int32, device=d)
        self.loss = torch.zeros(1, dtype=torch.float32, device=d)
        # NVSwitch multicast (multimem.ld_reduce / multimem.st) wins from 4 ranks up (measured in
        # round 1: 0.401 vs 0.434 ms/step at 8 ranks, a few percent at 4, slightly slower at 2)
        want_mc = comm == "mc" or (comm == "fused" and self.info.world >= 4)
        self.use_mc = bool(want_mc and self.info.world > 1 and self.grads.multicast_ptr
                           and self.params.multicast_ptr and self.w1c.multicast_ptr)
        if comm == "mc" and self.info.world > 1 and not self.use_mc:
            raise RuntimeError("comm='mc': symmetric memory reports no multicast address on this box")
        # rows of W1 per publication unit of the gradient; fixes the owner of every slice of the
        # sharded fp32 master, so it depends on the layer only (never on a step's row count)
        self.unit_rows = 256 if self.Hp >= 512 else 128
        self.nunits = (self.Hp + self.unit_rows - 1) // self.unit_rows
        self.timeout_ms = int(timeout_ms if timeout_ms is not None else os.environ.get("CME_PEER_TIMEOUT_MS", 20000))
        # the update kernel runs next to the dW1 GEMM; CME_NO_OVERLAP=1 puts it behind the GEMM
        # (for tools that serialise kernels: ncu, compute-sanitizer)
        self.overlap = not env_flag("CME_NO_OVERLAP")
        self._graphs = {}                         # CUDA graphs of whole steps, keyed by (rows, want_loss)
        self.side_stream = torch.cuda.Stream(device=d)
        self.ev = [torch.cuda.Event() for _ in range(4)]
        for e in self.ev:
            e.record()                            # torch creates the CUDA event on first use
        self.plan = dev.make_mlp_plan(self)
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

    def owned_slices(self, rank: int | None = None):
        """[(begin, end)] of the flat vector whose fp32 master `rank` keeps: its 1/world slice of
        every update unit (the same arithmetic as the update kernel, cme_mlp_unit_slice)."""
        rank = self.info.rank if rank is None else rank
        return [dev.mlp_unit_slice(self.plan, u, rank) for u in range(self.nunits + 1)]

    def gather_master(self) -> None:
        """The fp32 master of W1 is sharded across ranks (every update unit is cut into one slice
        per rank); rebuild the full copy on every rank: keep what is mine, zero the rest, then a
        sum all-reduce (x + 0 is exact)."""
        if self.info.world == 1 or self.comm == "nccl":
            return
        w1 = self.params.tensor[: self.n_w1]
        mine = torch.zeros_like(w1)
        for b, e in self.owned_slices():
            b, e = min(b, self.n_w1), min(e, self.n_w1)
            if e > b:
                mine[b:e] = w1[b:e]
        dist.all_reduce(mine)
        w1.copy_(mine)

    def check_health(self) -> None:
        """Raises if a cross-GPU wait of the update kernel ran into its timeout (a peer died or
        skipped a step).  Synchronises the device."""
        torch.cuda.synchronize(self.info.device)
        st = dev.mlp_status(self.plan)
        if st != 0:
            what = {1: "a gradient unit of some rank never arrived", 2: "a rank never finished its update",
                    3: "a k-chunk of the dW1 GEMM never got its turn (its grid was not fully resident)"}
            raise NativeError(f"rank {self.info.rank}: update kernel timed out after {self.timeout_ms} ms "
                              f"({what.get(st & 0xFF, 'unknown wait')}, detail {st >> 8}); "
                              "the replicas are no longer consistent")

    def store_params(self, nn) -> None:
        self.check_health()
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

    # ---- one minibatch slice, piece by piece (tests, profiling, the baseline paths) ----
    def stage(self, x_u8_ptr: int, rows: int) -> None:
        """uint8 pixels at x_u8_ptr (local or peer-mapped) -> bf16 rows + ones column; the same
        launch refreshes the bf16 copies of W2 (the step executor does likewise, so whatever
        runs a forward pass starts from the current W2)."""
        if self.tail == "tc":
            _, w2, _ = self._views(self.params.tensor)
            dev.mlp_stage_recast(x_u8_ptr, self.xb, rows, self.D, w2, self.w2c, self.w2t, self.Hh, self.C)
        else:
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

    def backward(self, rows: int, publish: bool = False) -> None:
        """dZ1, dW2 (accumulated: zero the gradient tail first) and dW1 (stored).  publish=True
        raises the per-unit flags the streaming update waits for."""
        _, w2, _ = self._views(self.params.tensor)
        gw1, gw2, _ = self._views(self.grads.tensor)
        if self.tail == "tc":
            if self.fused_tail:
                dev.mlp_bwd_tail_tc(self.a1, self.dz2b, self.w2t, self.dz1, gw2, self.Hh, rows, self.Hh,
                                    ld=self.ldh, C=self.C)
            else:
                dev.mlp_bwd_dz1_tc(self.a1, self.dz2b, self.w2t, self.dz1, rows, self.Hh, ld=self.ldh)
                dev.mlp_dw2_tc(self.a1, self.dz2b, gw2, rows, self.Hh, self.C)
        else:
            dev.mlp_bwd_elem(self.a1, self.dz2, w2, self.dz1, gw2, rows, self.Hh, self.C)
        dev.mlp_dw1_sync(self.dz1, self.xb, gw1, rows, self.K, self.Hh, self.ldx, self.sync_scratch,
                         self.pads_dev if publish else None, self.info.world, self.unit_rows, ldh=self.ldh,
                         status_ptr=self.pads.tensor.data_ptr() + 4 * dev.PAD_STATUS)

    def update(self, lr: float, reg: float) -> None:
        """Gradient all-reduce + L2 + SGD + bf16 re-cast (ref fpcode/neural_network.cpp:501-541).
        Follows backward(publish=True) on the same stream for the fused paths."""
        if self.comm == "nccl":
            # baseline path: library all-reduce, then the one-rank update kernel on local pointers
            w2_begin, w2_end = self.o_w2, self.o_w2 + self.C * self.Hh
            dist.all_reduce(self.grads.tensor)
            dev.allreduce_sgd([self.grads.tensor.data_ptr()], [self.params.tensor.data_ptr()],
                              [self.w1c.tensor.data_ptr()], 0, self.n_total, self.n_w1, self.ldx,
                              self.D, w2_begin, w2_end, lr, reg)
        else:
            dev.mlp_update_stream(self.plan, lr, reg)

    def _step_body(self, x_u8_ptr: int, labels: torch.Tensor, rows: int, lr: float, reg: float,
                   norm: float, want_loss: bool) -> None:
        """The step issued piece by piece from Python on one stream: the NCCL baseline and the
        CUDA-core tail; the product path is the native executor (see step())."""
        self.grads.tensor[self.o_w2:].zero_()
        self.stage(x_u8_ptr, rows)
        if want_loss:
            self.loss.zero_()
        self.forward(rows, labels, 1.0 / norm, want_loss=want_loss)
        self.backward(rows, publish=self.comm != "nccl")
        self.update(lr, reg)

    def step(self, x_u8_ptr: int, labels: torch.Tensor, rows: int, lr: float, reg: float,
             norm: float, want_loss: bool = False, graph: bool = False) -> None:
        """One training step on `rows` local samples; `norm` = global sample count of the
        minibatch (in_proc * P, ref fpcode/neural_network.cpp:330-334).  With want_loss the summed
        cross-entropy of this step is left in `self.loss`.

        graph=True replays the step as one CUDA graph.  The graph holds the step in its indirect
        form: inputs, labels, lr, reg and 1/norm are read on the device from `step_args`, which a
        one-thread kernel rewrites before every replay, so ONE graph per row count serves every
        minibatch position and every learning rate."""
        if rows > self.max_rows:
            raise ValueError(f"step: {rows} rows exceed the engine capacity {self.max_rows}")
        x_u8_ptr, pf_slot = self._prefetched(x_u8_ptr, rows)
        if self.comm == "nccl" or self.tail != "tc":
            self._step_body(x_u8_ptr, labels, rows, lr, reg, norm, want_loss)
            self._prefetch_consumed(pf_slot)
            return
        scale = 1.0 / norm
        if not graph:
            if want_loss:
                self.loss.zero_()
            dev.mlp_step_native(self.plan, x_u8_ptr, labels, rows, lr, reg, scale,
                                self.loss if want_loss else None, self.side_stream, self.ev)
            self._prefetch_consumed(pf_slot)
            return
        dev.mlp_set_step_args(self.plan, x_u8_ptr, labels, lr, reg, scale)
        key = (rows, want_loss)
        entry = self._graphs.get(key)
        if entry is None:
            self._indirect_step(rows, want_loss)                  # the real step (also warms every kernel up)
            before = dev.launches()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._indirect_step(rows, want_loss)              # recorded, not run
            n = dev.launches() - before
            dev._count(-n)
            self._graphs[key] = (g, n)
            self._prefetch_consumed(pf_slot)
            return
        entry[0].replay()
        dev._count(entry[1])
        self._prefetch_consumed(pf_slot)

    # ---- owner-mode prefetch: the next step's slice crosses NVLink while this step computes ----
    def prefetch(self, x_u8_ptr: int, rows: int, local: bool = False) -> None:
        """Start copying `rows` rows of uint8 pixels at the (peer-mapped) address x_u8_ptr into a
        local buffer on a copy stream.  The next step() called with this address stages from the
        local copy, so the pull over NVLink (the reference's MPI_Scatter of the minibatch,
        ref fpcode/neural_network.cpp:460-481) is off the step's critical path.  Call it right
        after issuing the step that precedes the one it feeds.  local=True (the address is in this
        GPU's own memory: the owner rank itself) does nothing: a copy would only cost bandwidth."""
        if local:
            return
        if not hasattr(self, "_pf"):
            d = self.info.device
            n = self.max_rows * self.D + 48
            self._pf = {"buf": [torch.empty(n, dtype=torch.uint8, device=d) for _ in range(2)],
                        "ready": [torch.cuda.Event() for _ in range(2)],
                        "read": [torch.cuda.Event() for _ in range(2)], "read_valid": [False, False],
                        "src": [None, None], "next": 0,
                        "stream": torch.cuda.Stream(device=d)}
        pf = self._pf
        slot = pf["next"]
        pf["next"] ^= 1
        nbytes = rows * self.D
        cs = pf["stream"]
        if pf["read_valid"][slot]:
            cs.wait_event(pf["read"][slot])               # the step that staged from this slot is done
        with torch.cuda.stream(cs):
            # on a copy engine: the step's persistent GEMM kernels hold every SM's registers, and a
            # copy KERNEL was only scheduled in the gaps between them (8 GPUs: 0.374 ms/step against
            # 0.382 without any prefetch and 0.330 with the data already local)
            dev.peer_memcpy(x_u8_ptr, pf["buf"][slot], nbytes)
            pf["ready"][slot].record(cs)
        pf["src"][slot] = (x_u8_ptr, rows, pf["buf"][slot].data_ptr())

    def _prefetched(self, x_u8_ptr: int, rows: int):
        """(local address, slot) if this slice was prefetched, else (x_u8_ptr, None); makes the
        current stream wait for the copy."""
        pf = getattr(self, "_pf", None)
        if pf is not None:
            for slot in (0, 1):
                src = pf["src"][slot]
                if src is not None and src[0] == x_u8_ptr and src[1] >= rows:
                    torch.cuda.current_stream().wait_event(pf["ready"][slot])
                    pf["src"][slot] = None
                    return src[2], slot
        return x_u8_ptr, None

    def _prefetch_consumed(self, slot) -> None:
        if slot is not None:
            self._pf["read"][slot].record()
            self._pf["read_valid"][slot] = True

    def _indirect_step(self, rows: int, want_loss: bool) -> None:
        if want_loss:
            self.loss.zero_()
        dev.mlp_step_native(self.plan, None, None, rows, 0.0, 0.0, 0.0, self.loss if want_loss else None,
                            self.side_stream, self.ev)

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

    def __init__(self, H, max_rows: int, info: DistInfo | None = None, comm: str = "fused",
                 timeout_ms: int | None = None):
        """comm: "fused" = one kernel per step does the gradient all-reduce, L2 and SGD over peer
        memory (cme_allreduce_sgd_f64); "nccl" = dist.all_reduce + local SGD kernels (baseline)."""
        self.info = info or DistInfo(device=torch.device("cuda", torch.cuda.current_device()))
        self.D, self.Hh, self.C = H
        self.max_rows = max_rows
        d, f = self.info.device, torch.float64
        D, Hh, C = self.D, self.Hh, self.C
        self.sizes = [Hh * D, Hh, C * Hh, C]
        self.offs = [0, Hh * D, Hh * D + Hh, Hh * D + Hh + C * Hh]
        n = sum(self.sizes)
        self.n_pad = n + (n & 1)                  # the update kernel moves pairs of doubles
        self.comm = ("nccl" if comm == "nccl" else "fused") if self.info.world > 1 else "local"
        self.timeout_ms = int(timeout_ms if timeout_ms is not None else os.environ.get("CME_PEER_TIMEOUT_MS", 20000))
        if self.comm == "nccl":
            self.params = torch.zeros(self.n_pad, dtype=f, device=d)
            self.grads = torch.zeros(self.n_pad, dtype=f, device=d)
        else:
            self._params = SymmetricTensor(self.n_pad, f, self.info)
            self._grads = SymmetricTensor(self.n_pad, f, self.info)
            self._pads = SymmetricTensor(dev.PAD_WORDS, torch.int32, self.info)
            self._pads_dev = self._pads.ptr_table_dev()
            self.params, self.grads = self._params.tensor, self._grads.tensor
            if self.info.world > 1:
                dist.barrier()                    # every rank's pad is zeroed before anyone signals
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
        torch.cuda.synchronize()
        if self.info.world > 1:
            dist.barrier()                        # no peer's update may land before this copy

    def store_params(self, nn) -> None:
        self.check_health()
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
        if self.comm == "nccl":
            dist.all_reduce(self.grads)             # baseline: library all-reduce + four SGD kernels
            for i, r in ((0, reg), (1, 0.0), (2, reg), (3, 0.0)):
                dev.d_sgd(self._seg(self.params, i), self._seg(self.grads, i), lr, r)
            return
        dev.allreduce_sgd_f64(self._grads.peer_ptrs, self._params.peer_ptrs, self._pads_dev, self.info.rank,
                              self.n_pad, self.offs[1], self.offs[2], self.offs[3], lr, reg, self.timeout_ms)

    def check_health(self) -> None:
        """Raises if a cross-GPU wait of the update kernel expired.  Synchronises the device."""
        if self.comm == "nccl":
            return
        torch.cuda.synchronize(self.info.device)
        st = int(self._pads.tensor[dev.F64_STATUS_WORD].item())
        if st:
            raise NativeError(f"rank {self.info.rank}: fp64 update timed out after {self.timeout_ms} ms "
                              f"waiting for rank {st - 1}; the replicas are no longer consistent")

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
