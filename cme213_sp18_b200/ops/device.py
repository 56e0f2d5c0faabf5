"""torch-facing wrappers over the hand-written sm_100a kernels in libcme213b.so.

Every function here launches this repo's own CUDA kernels through the C ABI declared in
include/cme213b.h; none of them falls back to a PyTorch op.  A module-level launch counter
(`launches()`) counts kernel launches issued through this module so benches can report how
many of the repo's kernels ran inside a timed region.
"""
from __future__ import annotations

import torch

from ._lib import check, device_lib, ptr, require_cuda, stream_ptr

_launches = 0


def launches() -> int:
    return _launches


def _count(n: int = 1) -> None:
    global _launches
    _launches += n


def device_info():
    import ctypes as C

    sm, maj, mnr = C.c_int(0), C.c_int(0), C.c_int(0)
    check(device_lib().cme_device_info(C.addressof(sm), C.addressof(maj), C.addressof(mnr)),
          "cme_device_info")
    return {"sm_count": sm.value, "cc": (maj.value, mnr.value)}


def l2_fetch_granularity(nbytes: int = 0) -> int:
    """Set the device's L2 fetch granularity (32 / 64 / 128 bytes; 0 = only query); returns the
    previous value.  Gather-heavy kernels (PageRank on graphs larger than L2) gain from 32."""
    import ctypes as C

    old = C.c_int(0)
    check(device_lib().cme_l2_fetch_granularity(nbytes, C.addressof(old)), "cme_l2_fetch_granularity")
    return old.value


# ------------------------------------------------------------------ hw2 q1
SHIFT_VARIANTS = {"char": 0, "uint": 1, "uint2": 2, "uint4": 3}


def shift_bytes(src: torch.Tensor, shift: int, variant: str = "uint4",
                out: torch.Tensor | None = None) -> torch.Tensor:
    """out[i] = src[i] + shift mod 256 (ref hw2code/shift.cu:13-99)."""
    require_cuda(src)
    assert src.dtype == torch.uint8 and src.is_contiguous()
    if out is None:
        out = torch.empty_like(src)
    check(device_lib().cme_shift_bytes(ptr(src), ptr(out), src.numel(), shift & 0xFF,
                                       SHIFT_VARIANTS[variant], stream_ptr()), "cme_shift_bytes")
    _count(1 if src.numel() % 16 == 0 or variant == "char" else 2)
    return out


# ------------------------------------------------------------------ hw2 q2
def pagerank_prescale(rank: torch.Tensor, inv_deg: torch.Tensor, contrib: torch.Tensor) -> None:
    check(device_lib().cme_pagerank_prescale(ptr(rank), ptr(inv_deg), ptr(contrib), rank.numel(),
                                             stream_ptr()), "cme_pagerank_prescale")
    _count()


def pagerank_step(row_ptr, edges_ptr: int, contrib_in, inv_deg, rank_out, contrib_out,
                  row_begin: int, row_end: int, num_nodes: int, peer_contrib: torch.Tensor | None = None,
                  nranks: int = 1) -> None:
    """One propagation step (ref hw2code/pagerank.cu:9-28). edges_ptr is an address so a
    row-partitioned caller can pass (local_edges - first_local_edge)."""
    check(device_lib().cme_pagerank_step(ptr(row_ptr), edges_ptr, ptr(contrib_in), ptr(inv_deg),
                                         ptr(rank_out), ptr(contrib_out), ptr(peer_contrib), nranks,
                                         row_begin, row_end, num_nodes, stream_ptr()),
          "cme_pagerank_step")
    _count()


def pagerank_iterate(row_ptr, edges, rank0, inv_deg, contrib_a, contrib_b, rank_out, iterations: int) -> None:
    """All iterations in one cooperative launch with grid barriers (ref hw2code/pagerank.cu:59-136)."""
    check(device_lib().cme_pagerank_iterate(ptr(row_ptr), ptr(edges), ptr(rank0), ptr(inv_deg),
                                            ptr(contrib_a), ptr(contrib_b), ptr(rank_out), iterations,
                                            rank0.numel(), stream_ptr()), "cme_pagerank_iterate")
    _count(1)            # one cooperative launch: pre-scale + every iteration


def pagerank_block_step(blk_ptr_addr: int, edges_b, contrib_in, partial, inv_deg, rank_out, contrib_out,
                        row_begin: int, nrows: int, num_nodes: int, first: bool, last: bool,
                        peer_contrib: torch.Tensor | None = None, nranks: int = 1) -> None:
    """One pass of the column-blocked propagation; blk_ptr_addr is the address of this block's
    nrows + 1 int32 offsets into edges_b."""
    check(device_lib().cme_pagerank_block_step(blk_ptr_addr, ptr(edges_b), ptr(contrib_in), ptr(partial), ptr(inv_deg),
                                               ptr(rank_out), ptr(contrib_out), ptr(peer_contrib), nranks, row_begin,
                                               nrows, num_nodes, int(first), int(last), stream_ptr()),
          "cme_pagerank_block_step")
    _count()


def pagerank_gen_edges(edges: torch.Tensor, edge_begin: int, num_nodes: int, seed: int) -> None:
    check(device_lib().cme_pagerank_gen_edges(ptr(edges), edge_begin, edges.numel(), num_nodes,
                                              seed, stream_ptr()), "cme_pagerank_gen_edges")
    _count()


# ------------------------------------------------------------------ hw3
HEAT_VARIANTS = {"global": 0, "loop": 1, "shared": 2}


def heat_step(nxt: torch.Tensor, cur: torch.Tensor, order: int, xcfl: float, ycfl: float,
              variant: str = "shared", bc_scale: float | None = None) -> None:
    """One time step of the 2-D heat stencil (ref hw3code/gpuStencil.cu:62-74, 163-179).
    If bc_scale is given the border decay rides in the same launch (ref hw3code/BC.h:7-38)."""
    require_cuda(nxt, cur)
    gy, gx = cur.shape
    check(device_lib().cme_heat_step(ptr(nxt), ptr(cur), gx, gy, order, xcfl, ycfl,
                                     HEAT_VARIANTS[variant], 0 if bc_scale is None else 1,
                                     0.0 if bc_scale is None else bc_scale, stream_ptr()),
          "cme_heat_step")
    _count()


def heat_bc(nxt: torch.Tensor, cur: torch.Tensor, order: int, bc_scale: float) -> None:
    gy, gx = cur.shape
    check(device_lib().cme_heat_bc(ptr(nxt), ptr(cur), gx, gy, order, bc_scale, stream_ptr()),
          "cme_heat_bc")
    _count()


def heat_slab_step(nxt: torch.Tensor, cur: torch.Tensor, up_ghost_ptr: int | None,
                   down_ghost_ptr: int | None, rows_local: int, order: int, xcfl: float,
                   ycfl: float, bc_scale: float, is_top: bool, is_bottom: bool) -> None:
    gx = cur.shape[1]
    check(device_lib().cme_heat_slab_step(ptr(nxt), ptr(cur), up_ghost_ptr, down_ghost_ptr, gx,
                                          rows_local, order, xcfl, ycfl, bc_scale, int(is_top),
                                          int(is_bottom), stream_ptr()), "cme_heat_slab_step")
    _count()


# ------------------------------------------------------------------ hw4
def sanitize_compact(raw: torch.Tensor):
    """Lower-case and keep letters only (ref hw4code/create_cipher.cu:184-195).
    Returns (buffer, count tensor); buffer[:count] is the clean text."""
    require_cuda(raw)
    n = raw.numel()
    out = torch.empty(max(n, 1), dtype=torch.uint8, device=raw.device)
    count = torch.zeros(1, dtype=torch.int64, device=raw.device)
    scratch = torch.empty(4 * (n // 2048 + 2) + 16, dtype=torch.uint8, device=raw.device)
    check(device_lib().cme_sanitize_compact(ptr(raw), n, ptr(out), ptr(count), ptr(scratch),
                                            scratch.numel(), stream_ptr()), "cme_sanitize_compact")
    _count(3)
    return out, count


def byte_histogram(text: torch.Tensor, stride: int = 1, offset: int = 0) -> torch.Tensor:
    require_cuda(text)
    counts = torch.empty(256, dtype=torch.int64, device=text.device)
    check(device_lib().cme_byte_histogram(ptr(text), text.numel(), stride, offset, ptr(counts),
                                          stream_ptr()), "cme_byte_histogram")
    _count()
    return counts


def strided_histograms(text: torch.Tensor, period: int) -> torch.Tensor:
    require_cuda(text)
    counts = torch.empty((period, 256), dtype=torch.int64, device=text.device)
    check(device_lib().cme_strided_histograms(ptr(text), text.numel(), period, ptr(counts),
                                              stream_ptr()), "cme_strided_histograms")
    _count()
    return counts


def vigenere_shift(text: torch.Tensor, shifts: torch.Tensor, wrap26: bool = False) -> torch.Tensor:
    require_cuda(text, shifts)
    assert shifts.dtype == torch.int32
    out = torch.empty_like(text)
    check(device_lib().cme_vigenere_shift(ptr(text), ptr(out), text.numel(), ptr(shifts),
                                          shifts.numel(), int(wrap26), stream_ptr()),
          "cme_vigenere_shift")
    _count()
    return out


def lag_match_counts(text: torch.Tensor, lag0: int, num_lags: int) -> torch.Tensor:
    require_cuda(text)
    out = torch.empty(num_lags, dtype=torch.int64, device=text.device)
    check(device_lib().cme_lag_match_counts(ptr(text), text.numel(), lag0, num_lags, ptr(out),
                                            stream_ptr()), "cme_lag_match_counts")
    _count()
    return out


# ------------------------------------------------------------------ hw1 on the GPU
def parity_sum(values: torch.Tensor) -> torch.Tensor:
    require_cuda(values)
    assert values.dtype in (torch.uint32, torch.int32)
    out = torch.empty(2, dtype=torch.int64, device=values.device)
    check(device_lib().cme_parity_sum(ptr(values), values.numel(), ptr(out), stream_ptr()),
          "cme_parity_sum")
    _count()
    return out


def radix_sort(keys: torch.Tensor) -> torch.Tensor:
    """Stable LSD radix sort of 32-bit keys, treated as unsigned. Returns a sorted copy."""
    require_cuda(keys)
    assert keys.dtype in (torch.uint32, torch.int32) and keys.is_contiguous()
    k = keys.clone()
    tmp = torch.empty_like(k)
    nbytes = int(device_lib().cme_radix_scratch_bytes(k.numel()))
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=k.device)
    check(device_lib().cme_radix_sort(ptr(k), ptr(tmp), k.numel(), ptr(scratch), nbytes,
                                      stream_ptr()), "cme_radix_sort")
    _count(12)
    return k


# ------------------------------------------------------------------ fp64 parity path
def dgemm(transA: bool, transB: bool, M: int, N: int, K: int, alpha: float, A: torch.Tensor, lda: int,
          B: torch.Tensor, ldb: int, beta: float, Cm: torch.Tensor, ldc: int, simt: bool = False) -> None:
    """Column-major C = alpha*op(A)*op(B) + beta*C (ref fpcode/gpu_func.cu:259-297) on the fp64
    tensor path (DMMA); simt=True runs the round-1 CUDA-core kernel (A/B timing only)."""
    require_cuda(A, B, Cm)
    assert A.dtype == B.dtype == Cm.dtype == torch.float64
    fn = device_lib().cme_dgemm_simt if simt else device_lib().cme_dgemm
    check(fn(int(transA), int(transB), M, N, K, alpha, ptr(A), lda, ptr(B), ldb, beta, ptr(Cm), ldc, stream_ptr()),
          "cme_dgemm")
    _count()


def d_bias_sigmoid(z, bias, rows, cols):
    check(device_lib().cme_d_bias_sigmoid(ptr(z), ptr(bias), rows, cols, stream_ptr()),
          "cme_d_bias_sigmoid")
    _count()


def d_bias_softmax_ce(z, bias, labels, diff, scale, loss_sum, rows, cols):
    check(device_lib().cme_d_bias_softmax_ce(ptr(z), ptr(bias), ptr(labels), ptr(diff), scale,
                                             ptr(loss_sum), rows, cols, stream_ptr()),
          "cme_d_bias_softmax_ce")
    _count()


def d_dsigmoid(da, a, dz):
    check(device_lib().cme_d_dsigmoid(ptr(da), ptr(a), ptr(dz), a.numel(), stream_ptr()),
          "cme_d_dsigmoid")
    _count()


def d_row_sum(m, out, rows, cols):
    check(device_lib().cme_d_row_sum(ptr(m), ptr(out), rows, cols, stream_ptr()), "cme_d_row_sum")
    _count()


def d_sgd(p, g, lr, reg):
    check(device_lib().cme_d_sgd(ptr(p), ptr(g), p.numel(), lr, reg, stream_ptr()), "cme_d_sgd")
    _count()


def d_axpby(c, a, b, alpha, beta):
    check(device_lib().cme_d_axpby(ptr(c), ptr(a), ptr(b), alpha, beta, c.numel(), stream_ptr()),
          "cme_d_axpby")
    _count()


# ------------------------------------------------------------------ bf16 tensor-core path
def u8_to_bf16_rows(src_ptr: int, dst: torch.Tensor, rows: int, cols: int, add_ones: bool = True) -> None:
    """uint8 rows -> bf16 rows of dst (leading dim = dst.shape[1]); src_ptr may be a peer address
    (the minibatch scatter, ref fpcode/neural_network.cpp:460-481)."""
    ld = dst.shape[1]
    check(device_lib().cme_u8_to_bf16_rows(src_ptr, ptr(dst), rows, cols, ld, int(add_ones),
                                           stream_ptr()), "cme_u8_to_bf16_rows")
    _count()


def mlp_stage_recast(src_ptr: int, dst: torch.Tensor, rows: int, cols: int, W2: torch.Tensor, w2c: torch.Tensor,
                     w2t: torch.Tensor, H: int, C: int = 10) -> None:
    """The staging launch of a step: uint8 rows -> bf16 rows + ones column, and in the same launch
    the bf16 copies of W2 that layer 2 and the backward tail read."""
    check(device_lib().cme_mlp_stage_recast(src_ptr, None, ptr(dst), rows, cols, dst.shape[1], ptr(W2), ptr(w2c),
                                            w2c.shape[1], ptr(w2t), H, C, stream_ptr()), "cme_mlp_stage_recast")
    _count()


def mlp_fwd1(X: torch.Tensor, W1c: torch.Tensor, A1: torch.Tensor, rows: int, K: int, H: int) -> None:
    """A1 = sigmoid(X @ W1c.T) on tcgen05 (ref fpcode/neural_network.cpp:289-300)."""
    check(device_lib().cme_mlp_fwd1(ptr(X), ptr(W1c), None, ptr(A1), rows, K, H, X.shape[1],
                                    A1.shape[1], stream_ptr()), "cme_mlp_fwd1")
    _count()


def gemm_bf16(A: torch.Tensor, B: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """C[M,N] = A[M,K] @ B[N,K].T in bf16 with fp32 accumulation on tcgen05."""
    require_cuda(A, B)
    assert A.dtype == B.dtype == torch.bfloat16 and A.stride(1) == 1 and B.stride(1) == 1
    M, K = A.shape
    N = B.shape[0]
    if out is None:
        out = torch.empty((M, N), dtype=torch.bfloat16, device=A.device)
    check(device_lib().cme_gemm_bf16(ptr(A), ptr(B), ptr(out), M, N, K, A.stride(0), B.stride(0),
                                     out.stride(0), stream_ptr()), "cme_gemm_bf16")
    _count()
    return out


def mlp_fwd2_softmax(A1, W2, b2, labels, dz2, loss_sum, pred, gb2, scale: float, rows: int, H: int,
                     C: int = 10) -> None:
    check(device_lib().cme_mlp_fwd2_softmax(ptr(A1), ptr(W2), ptr(b2), ptr(labels), ptr(dz2),
                                            ptr(loss_sum), ptr(pred), ptr(gb2), scale, rows, H,
                                            A1.shape[1], C, stream_ptr()), "cme_mlp_fwd2_softmax")
    _count()


def mlp_bwd_elem(A1, dz2, W2, dZ1, gW2, rows: int, H: int, C: int = 10) -> None:
    check(device_lib().cme_mlp_bwd_elem(ptr(A1), ptr(dz2), ptr(W2), ptr(dZ1), ptr(gW2), rows, H,
                                        A1.shape[1], C, stream_ptr()), "cme_mlp_bwd_elem")
    _count()


def mlp_fwd2_tc(A1, W2c, b2, labels, dz2, dz2b, loss_sum, pred, gb2, scale: float, rows: int, H: int,
                C: int = 10) -> None:
    """Layer 2 + softmax + CE gradient on tcgen05 with everything after the GEMM in the epilogue."""
    check(device_lib().cme_mlp_fwd2_tc(ptr(A1), ptr(W2c), ptr(b2), ptr(labels), ptr(dz2), ptr(dz2b),
                                       ptr(loss_sum), ptr(pred), ptr(gb2), scale, rows, H,
                                       A1.shape[1], C, stream_ptr()), "cme_mlp_fwd2_tc")
    _count()


def mlp_bwd_dz1_tc(A1, dz2b, W2T, dZ1, rows: int, H: int, ld: int | None = None) -> None:
    """dZ1 = (dz2 @ W2) * a * (1 - a) on tcgen05, streaming A1 through a TMA staging ring.
    A1 / dZ1 may be column-block views of wider buffers: pass their row pitch as ld."""
    check(device_lib().cme_mlp_bwd_dz1_tc(ptr(A1), ptr(dz2b), ptr(W2T), ptr(dZ1), rows, H,
                                          ld if ld is not None else A1.shape[1], stream_ptr()),
          "cme_mlp_bwd_dz1_tc")
    _count()


def mlp_bwd_tail_tc(A1, dz2b, W2T, dZ1, gW2, ldg2: int, rows: int, H: int, ld: int | None = None,
                    C: int = 10) -> None:
    """dZ1 and dW2 in one pass over A1 (fused backward tail on tcgen05)."""
    check(device_lib().cme_mlp_bwd_tail_tc(ptr(A1), ptr(dz2b), ptr(W2T), ptr(dZ1), ptr(gW2), ldg2, rows, H,
                                           ld if ld is not None else A1.shape[1], C, stream_ptr()),
          "cme_mlp_bwd_tail_tc")
    _count()


def fwd_fused_ok(rows: int, H: int) -> bool:
    return bool(device_lib().cme_mlp_fwd_fused_ok(rows, H))


def fwd_scratch_floats(max_rows: int) -> int:
    return int(device_lib().cme_mlp_fwd_scratch_floats(max_rows))


def mlp_fwd_fused(X, W1c, w2c, b2, labels, A1, dz2, dz2b, loss_sum, pred, gb2, scratch, max_rows: int,
                  scale: float, rows: int, K: int, H: int, C: int = 10) -> None:
    """Forward pass in one tensor-core kernel (layer-2 logits ride on the staged A1 tiles) plus the
    per-row softmax tail; see cme_mlp_fwd_fused."""
    check(device_lib().cme_mlp_fwd_fused(ptr(X), ptr(W1c), ptr(w2c), ptr(b2), ptr(labels), None, ptr(A1),
                                         ptr(dz2), ptr(dz2b), ptr(loss_sum), ptr(pred), ptr(gb2), ptr(scratch),
                                         max_rows, scale, rows, K, H, X.shape[1], A1.shape[1], C, stream_ptr()),
          "cme_mlp_fwd_fused")
    _count(2)


def mlp_recast_w2(W2, w2c, w2t, H: int, C: int = 10) -> None:
    check(device_lib().cme_mlp_recast_w2(ptr(W2), ptr(w2c), w2c.shape[1], ptr(w2t), H, C, stream_ptr()),
          "cme_mlp_recast_w2")
    _count()


def mlp_dw2_tc(A1, dz2b, gW2, rows: int, H: int, C: int = 10) -> None:
    check(device_lib().cme_mlp_dw2_tc(ptr(A1), ptr(dz2b), ptr(gW2), rows, H, A1.shape[1], C,
                                      stream_ptr()), "cme_mlp_dw2_tc")
    _count()


def f32_to_bf16_2d(src, ld_src: int, dst, ld_dst: int, rows: int, cols: int) -> None:
    check(device_lib().cme_f32_to_bf16_2d(ptr(src), ld_src, ptr(dst), ld_dst, rows, cols,
                                          stream_ptr()), "cme_f32_to_bf16_2d")
    _count()


def mlp_dw1(dZ1, X, gW1, rows: int, K: int, H: int, ldg: int, ldh: int | None = None) -> None:
    check(device_lib().cme_mlp_dw1(ptr(dZ1), ptr(X), ptr(gW1), rows, K, H, X.shape[1],
                                   ldh if ldh is not None else dZ1.shape[1], ldg, stream_ptr()),
          "cme_mlp_dw1")
    _count()


def _ptr_table(ptrs):
    import ctypes as C

    arr = (C.c_void_p * len(ptrs))(*[C.c_void_p(int(p)) for p in ptrs])
    return arr


def allreduce_sgd(grad_ptrs, param_ptrs, w1c_ptrs, rank: int, n_total: int, n_w1: int, ldx: int,
                  kreg: int, w2_begin: int, w2_end: int, lr: float, reg: float,
                  rng: tuple[int, int] | None = None) -> None:
    """Fused gradient all-reduce + L2 + SGD + bf16 re-cast over peer pointers
    (ref fpcode/neural_network.cpp:501-541)."""
    n = len(grad_ptrs)
    check(device_lib().cme_allreduce_sgd(_ptr_table(grad_ptrs), _ptr_table(param_ptrs),
                                         _ptr_table(w1c_ptrs), rank, n, n_total, n_w1, ldx, kreg,
                                         w2_begin, w2_end, lr, reg, *(rng or (0, n_total)), stream_ptr()),
          "cme_allreduce_sgd")
    _count()


def allreduce_sgd_mc(mc_grads: int, local_params: int, mc_params: int, mc_w1c: int, rank: int,
                     nranks: int, n_total: int, n_w1: int, ldx: int, kreg: int, w2_begin: int,
                     w2_end: int, lr: float, reg: float, rng: tuple[int, int] | None = None) -> None:
    check(device_lib().cme_allreduce_sgd_mc(mc_grads, local_params, mc_params, mc_w1c, rank, nranks,
                                            n_total, n_w1, ldx, kreg, w2_begin, w2_end, lr, reg,
                                            *(rng or (0, n_total)), stream_ptr()), "cme_allreduce_sgd_mc")
    _count()


F64_STATUS_WORD = 6 * 64 + 61     # pad word the fp64 update leaves its timeout code in


def allreduce_sgd_f64(grad_ptrs, param_ptrs, pads_dev: torch.Tensor, rank: int, n: int, w1_end: int,
                      w2_begin: int, w2_end: int, lr: float, reg: float, timeout_ms: int = 0) -> None:
    """fp64 gradient all-reduce + L2 + SGD with its own peer handshake, one launch
    (ref fpcode/neural_network.cpp:496-541 at the reference's precision)."""
    check(device_lib().cme_allreduce_sgd_f64(_ptr_table(grad_ptrs), _ptr_table(param_ptrs), ptr(pads_dev),
                                             rank, len(grad_ptrs), n, w1_end, w2_begin, w2_end, lr, reg,
                                             timeout_ms, stream_ptr()), "cme_allreduce_sgd_f64")
    _count()


def peer_memcpy(src_ptr: int, dst: torch.Tensor, nbytes: int) -> None:
    """Copy from a (peer-mapped) address on a copy engine (no SMs involved)."""
    check(device_lib().cme_peer_memcpy(src_ptr, ptr(dst), nbytes, stream_ptr()), "cme_peer_memcpy")


def peer_barrier(pads_dev: torch.Tensor, rank: int, nranks: int, channel: int, epoch: int,
                 signal: bool = True, wait: bool = True) -> None:
    check(device_lib().cme_peer_barrier(ptr(pads_dev), rank, nranks, channel, epoch & 0xFFFFFFFF,
                                        int(signal), int(wait), stream_ptr()), "cme_peer_barrier")
    _count()


def peer_copy(src_ptr: int, dst: torch.Tensor, nbytes: int, max_blocks: int = 0) -> None:
    """Copy from a (peer-mapped) address with this repo's copy kernel; max_blocks caps its grid."""
    check(device_lib().cme_peer_copy_bg(src_ptr, ptr(dst), nbytes, max_blocks, stream_ptr()), "cme_peer_copy")
    _count()


# ------------------------------------------------------------------ native step executor
PAD_WORDS = 4096          # csrc/common.cuh kPadWords
PAD_STATUS = 3106         # kPadStatus


def dw1_scratch_words() -> int:
    return int(device_lib().cme_mlp_dw1_scratch_bytes()) // 4


def _plan_type():
    import ctypes as C

    class CmeMlpPlan(C.Structure):
        _fields_ = [("D", C.c_int), ("H", C.c_int), ("C", C.c_int), ("Hp", C.c_int), ("ldx", C.c_int),
                    ("ldh", C.c_int), ("rank", C.c_int), ("nranks", C.c_int), ("max_rows", C.c_int64),
                    ("params", C.c_void_p * 16), ("grads", C.c_void_p * 16), ("w1c", C.c_void_p * 16),
                    ("xb", C.c_void_p), ("a1", C.c_void_p), ("dz1", C.c_void_p), ("dz2b", C.c_void_p),
                    ("w2c", C.c_void_p), ("w2t", C.c_void_p), ("dz2", C.c_void_p), ("pred", C.c_void_p),
                    ("pads_dev", C.c_void_p), ("overlap", C.c_int), ("use_mc", C.c_int),
                    ("grads_mc", C.c_void_p), ("params_mc", C.c_void_p), ("w1c_mc", C.c_void_p),
                    ("unit_rows", C.c_int), ("timeout_ms", C.c_int), ("pad_local", C.c_void_p),
                    ("sync_scratch", C.c_void_p), ("step_args", C.c_void_p), ("fwd_scratch", C.c_void_p),
                    ("dw2_scratch", C.c_void_p)]

    return CmeMlpPlan


def make_mlp_plan(eng):
    """ctypes mirror of CmeMlpPlan (include/cme213b.h) for an MLPEngineBF16."""
    p = _plan_type()()
    p.D, p.H, p.C, p.Hp, p.ldx, p.ldh = eng.D, eng.Hh, eng.C, eng.Hp, eng.ldx, eng.ldh
    p.rank, p.nranks, p.max_rows = eng.info.rank, eng.info.world, eng.max_rows
    for r in range(eng.info.world):
        p.params[r] = eng.params.peer_ptrs[r]
        p.grads[r] = eng.grads.peer_ptrs[r]
        p.w1c[r] = eng.w1c.peer_ptrs[r]
    p.xb, p.a1, p.dz1, p.dz2b = ptr(eng.xb), ptr(eng.a1), ptr(eng.dz1), ptr(eng.dz2b)
    p.w2c, p.w2t, p.dz2, p.pred = ptr(eng.w2c), ptr(eng.w2t), ptr(eng.dz2), ptr(eng.pred)
    p.pads_dev = ptr(eng.pads_dev)
    p.pad_local = ptr(eng.pads.tensor)
    p.sync_scratch = ptr(eng.sync_scratch)
    p.step_args = ptr(eng.step_args)
    p.fwd_scratch = ptr(eng.fwd_scratch)
    p.dw2_scratch = ptr(eng.dw2_scratch)
    p.unit_rows, p.timeout_ms = eng.unit_rows, eng.timeout_ms
    p.overlap, p.use_mc = int(eng.overlap), int(eng.use_mc)
    if eng.use_mc:
        p.grads_mc, p.params_mc, p.w1c_mc = eng.grads.multicast_ptr, eng.params.multicast_ptr, eng.w1c.multicast_ptr
    return p


def mlp_step_native(plan, x_u8_ptr: int | None, labels, rows: int, lr: float, reg: float,
                    scale: float, loss_sum: torch.Tensor | None = None, side_stream=None, events=None) -> None:
    """The whole step issued by the native executor (one C call).  x_u8_ptr None = the indirect
    form (inputs, labels, lr, reg, scale from the plan's device-side step arguments).  With a side
    stream (and 4 events) the update kernel runs next to the dW1 GEMM."""
    import ctypes as C

    lib = device_lib()
    if side_stream is not None:
        evs = (C.c_void_p * 4)(*[C.c_void_p(int(e.cuda_event)) for e in events])
        check(lib.cme_mlp_step_ex(C.addressof(plan), x_u8_ptr, ptr(labels), rows, lr, reg, scale, ptr(loss_sum),
                                  stream_ptr(), int(side_stream.cuda_stream), evs), "cme_mlp_step_ex")
    else:
        check(lib.cme_mlp_step(C.addressof(plan), x_u8_ptr, ptr(labels), rows, lr, reg, scale, ptr(loss_sum),
                               stream_ptr()), "cme_mlp_step")
    _count(lib.cme_mlp_step_launches(C.addressof(plan)))


def mlp_set_step_args(plan, x_u8_ptr: int, labels, lr: float, reg: float, scale: float) -> None:
    import ctypes as C

    check(device_lib().cme_mlp_set_step_args(C.addressof(plan), x_u8_ptr, ptr(labels), lr, reg, scale,
                                             stream_ptr()), "cme_mlp_set_step_args")
    _count()


def mlp_update_stream(plan, lr: float, reg: float) -> None:
    import ctypes as C

    check(device_lib().cme_mlp_update_stream(C.addressof(plan), lr, reg, stream_ptr()), "cme_mlp_update_stream")
    _count()


def mlp_status(plan) -> int:
    import ctypes as C

    out = C.c_uint32(0)
    check(device_lib().cme_mlp_status(C.addressof(plan), C.addressof(out)), "cme_mlp_status")
    return int(out.value)


def mlp_unit_slice(plan, unit: int, rank: int) -> tuple[int, int]:
    import ctypes as C

    out = (C.c_int64 * 2)()
    check(device_lib().cme_mlp_unit_slice(C.addressof(plan), unit, rank, out), "cme_mlp_unit_slice")
    return int(out[0]), int(out[1])


def mlp_dw1_sync(dZ1, X, gW1, rows: int, K: int, H: int, ldg: int, scratch, pads_dev, nranks: int,
                 unit_rows: int, ldh: int | None = None, status_ptr: int | None = None) -> None:
    check(device_lib().cme_mlp_dw1_sync(ptr(dZ1), ptr(X), ptr(gW1), rows, K, H, X.shape[1],
                                        ldh if ldh is not None else dZ1.shape[1], ldg, ptr(scratch),
                                        ptr(pads_dev), nranks, unit_rows, status_ptr, stream_ptr()),
          "cme_mlp_dw1_sync")
    _count()


def warmup() -> int:
    """Context + module warm-up with one trivial kernel (ref fpcode/gpu_func.cu:11-36 device_add_one);
    returns 1 when the kernel ran."""
    import ctypes as C

    lib = device_lib()
    lib.cme_warmup.argtypes = [C.POINTER(C.c_int)]
    lib.cme_warmup.restype = C.c_int
    v = C.c_int(0)
    check(lib.cme_warmup(C.byref(v)), "cme_warmup")
    _count(1)
    return v.value


class MlpRuntime:
    """Handle on the native streaming trainer (csrc/mlp_runtime.cu, cme_mlp_runtime_*): the whole
    minibatch loop of parallel_train (ref fpcode/neural_network.cpp:443-567) runs in C++, fed
    from pinned host memory, with one CUDA graph per step."""

    def __init__(self, engine, graphs: bool = True):
        import ctypes as C

        lib = device_lib()
        lib.cme_mlp_runtime_create.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
        lib.cme_mlp_runtime_create.restype = C.c_int
        lib.cme_mlp_runtime_run.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                            C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        lib.cme_mlp_runtime_run.restype = C.c_int
        lib.cme_mlp_runtime_stats.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
        lib.cme_mlp_runtime_stats.restype = C.c_int
        lib.cme_mlp_runtime_destroy.argtypes = [C.c_void_p]
        lib.cme_mlp_runtime_destroy.restype = C.c_int
        self._lib = lib
        self.engine = engine               # keeps the buffers named by the plan alive
        self.plan = make_mlp_plan(engine)
        h = C.c_void_p()
        with torch.cuda.device(engine.info.device):
            check(lib.cme_mlp_runtime_create(C.addressof(self.plan), int(graphs), C.byref(h)),
                  "cme_mlp_runtime_create")
        self._h = h
        self._counted = 0

    def run(self, x_host: torch.Tensor, lab_host: torch.Tensor, row_begin, rows, global_rows,
            lr: float, reg: float, loss_host: torch.Tensor | None = None) -> int:
        """x_host [N, D] uint8 and lab_host [N] int32 must be pinned; row_begin / rows /
        global_rows are per-step sequences (this rank's first row, its row count, the global
        sample count of the minibatch).  Asynchronous: synchronise the current stream (or read
        an event) before touching loss_host."""
        import ctypes as C

        n = len(rows)
        if not (x_host.is_pinned() and lab_host.is_pinned()):
            raise ValueError("MlpRuntime.run needs pinned host tensors")
        if x_host.dtype != torch.uint8 or lab_host.dtype != torch.int32 or not x_host.is_contiguous():
            raise ValueError("MlpRuntime.run needs contiguous uint8 pixels and int32 labels")
        if x_host.shape[1] != self.plan.D:
            raise ValueError(f"rows have {x_host.shape[1]} features, the plan expects {self.plan.D}")
        hi = max((b + r for b, r in zip(row_begin, rows)), default=0)
        if hi > x_host.shape[0] or hi > lab_host.shape[0]:
            raise ValueError("a step reads past the end of the host arrays")
        if loss_host is not None and (not loss_host.is_pinned() or loss_host.numel() < n
                                      or loss_host.dtype != torch.float32):
            raise ValueError("loss_host must be pinned float32 with one element per step")
        arr = C.c_int64 * n
        rb, rr, gr = arr(*row_begin), arr(*rows), arr(*global_rows)
        with torch.cuda.device(self.engine.info.device):
            check(self._lib.cme_mlp_runtime_run(self._h, x_host.data_ptr(), lab_host.data_ptr(), rb, rr, gr, n,
                                                lr, reg, loss_host.data_ptr() if loss_host is not None else None,
                                                stream_ptr()), "cme_mlp_runtime_run")
        st = self.stats()
        _count(st["launches"] - self._counted)
        self._counted = st["launches"]
        return n

    def stats(self) -> dict:
        import ctypes as C

        out = (C.c_int64 * 5)()
        check(self._lib.cme_mlp_runtime_stats(self._h, out), "cme_mlp_runtime_stats")
        return dict(zip(("h2d_bytes", "d2h_bytes", "launches", "steps", "graphs"), [int(v) for v in out]))

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.cme_mlp_runtime_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass
