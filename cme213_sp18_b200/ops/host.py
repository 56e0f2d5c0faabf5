"""NumPy-facing wrappers over libcme213b_host.so (the native C++/OpenMP host workloads).

These run on any box, GPU or not.  They are the hw1 programs of the reference
(ref hw1code/main_q1.cpp, hw1code/main_q2.cpp) and the CPU oracles for hw2-hw4.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import host_lib, np_ptr


def _u32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.uint32)


def num_threads() -> int:
    return int(host_lib().cmeh_num_threads())


# ---------------------------------------------------------------- hw1 q1
def parity_sum(values, parallel: bool = True, threads: int = 0):
    """(even_sum, odd_sum) of a uint32 vector as exact Python ints, plus the mod-2^32 pair the
    reference prints (ref hw1code/main_q1.cpp:16-45)."""
    v = _u32(values)
    out32 = np.zeros(2, dtype=np.uint32)
    out64 = np.zeros(2, dtype=np.uint64)
    lib = host_lib()
    if parallel:
        rc = lib.cmeh_parity_sum_omp(np_ptr(v), v.size, threads, np_ptr(out32), np_ptr(out64))
    else:
        rc = lib.cmeh_parity_sum_serial(np_ptr(v), v.size, np_ptr(out32), np_ptr(out64))
    if rc:
        raise ValueError("parity_sum: bad argument")
    return (int(out64[0]), int(out64[1])), (int(out32[0]), int(out32[1]))


# ---------------------------------------------------------------- hw1 q2
def radix_block_histograms(keys, num_blocks, num_bits, start_bit, block_size):
    k = _u32(keys)
    out = np.zeros(num_blocks << num_bits, dtype=np.uint32)
    rc = host_lib().cmeh_radix_block_histograms(np_ptr(k), k.size, num_blocks, num_bits, start_bit,
                                                block_size, np_ptr(out))
    if rc:
        raise ValueError("radix_block_histograms: bad argument")
    return out


def radix_reduce_histograms(block_histo, num_blocks, num_buckets):
    b = _u32(block_histo)
    out = np.zeros(num_buckets, dtype=np.uint32)
    host_lib().cmeh_radix_reduce_histograms(np_ptr(b), num_blocks, num_buckets, np_ptr(out))
    return out


def radix_scan_global(global_histo):
    g = _u32(global_histo)
    out = np.zeros_like(g)
    host_lib().cmeh_radix_scan_global(np_ptr(g), g.size, np_ptr(out))
    return out


def radix_block_exscan(global_exscan, block_histo, num_blocks, num_buckets):
    g, b = _u32(global_exscan), _u32(block_histo)
    out = np.zeros(num_blocks * num_buckets, dtype=np.uint32)
    host_lib().cmeh_radix_block_exscan(np_ptr(g), np_ptr(b), num_blocks, num_buckets, np_ptr(out))
    return out


def radix_populate(block_exscan, num_blocks, num_bits, start_bit, block_size, keys):
    bx, k = _u32(block_exscan), _u32(keys)
    out = np.zeros_like(k)
    rc = host_lib().cmeh_radix_populate(np_ptr(bx), num_blocks, num_bits, start_bit, block_size,
                                        np_ptr(k), k.size, np_ptr(out))
    if rc:
        raise ValueError("radix_populate: bad argument")
    return out


def radix_sort(keys, num_bits: int = 16, num_blocks: int = 8, parallel: bool = True,
               threads: int = 0) -> np.ndarray:
    """LSD radix sort of uint32 keys (ref hw1code/main_q2.cpp:151-218). Returns a sorted copy."""
    k = _u32(keys).copy()
    tmp = np.empty_like(k)
    lib = host_lib()
    if parallel:
        rc = lib.cmeh_radix_sort_parallel(np_ptr(k), np_ptr(tmp), k.size, num_bits, num_blocks, threads)
    else:
        rc = lib.cmeh_radix_sort_serial(np_ptr(k), np_ptr(tmp), k.size, num_bits)
    if rc:
        raise ValueError("radix_sort: num_bits must divide 16 and num_blocks must be > 0")
    return k


def std_sort(keys) -> np.ndarray:
    k = _u32(keys).copy()
    host_lib().cmeh_std_sort(np_ptr(k), k.size)
    return k


# ---------------------------------------------------------------- hw2
def shift_bytes(data: np.ndarray, shift: int) -> np.ndarray:
    d = np.ascontiguousarray(data, dtype=np.uint8)
    out = np.empty_like(d)
    host_lib().cmeh_shift_bytes(np_ptr(d), np_ptr(out), d.size, shift & 0xFF)
    return out


def pagerank_graph(num_nodes: int, avg_edges: int, seed: int = 1, with_edges: bool = True):
    """Reference-shaped synthetic graph (ref hw2code/main_q2.cu:87-120): returns
    (row_ptr int64[N+1], edges uint32[M] or None, inv_deg float32[N])."""
    row_ptr = np.zeros(num_nodes + 1, dtype=np.int64)
    inv_deg = np.zeros(num_nodes, dtype=np.float32)
    lib = host_lib()
    lib.cmeh_pagerank_graph(num_nodes, avg_edges, seed, np_ptr(row_ptr), None, np_ptr(inv_deg))
    edges = None
    if with_edges:
        edges = np.zeros(int(row_ptr[-1]), dtype=np.uint32)
        lib.cmeh_pagerank_graph(num_nodes, avg_edges, seed, np_ptr(row_ptr), np_ptr(edges),
                                np_ptr(inv_deg))
    return row_ptr, edges, inv_deg


def pagerank(row_ptr, edges, rank0, inv_deg, iterations: int) -> np.ndarray:
    rp = np.ascontiguousarray(row_ptr, dtype=np.int64)
    e = _u32(edges)
    r0 = np.ascontiguousarray(rank0, dtype=np.float32)
    inv = np.ascontiguousarray(inv_deg, dtype=np.float32)
    out = np.empty_like(r0)
    rc = host_lib().cmeh_pagerank(np_ptr(rp), np_ptr(e), np_ptr(r0), np_ptr(out), np_ptr(inv),
                                  iterations, r0.size)
    if rc:
        raise ValueError("pagerank: bad argument")
    return out


# ---------------------------------------------------------------- hw3
def heat_init(gx: int, gy: int, dx: float, dy: float) -> np.ndarray:
    g = np.empty((gy, gx), dtype=np.float32)
    host_lib().cmeh_heat_init(np_ptr(g), gx, gy, dx, dy)
    return g


def heat_cpu(grid: np.ndarray, order: int, iters: int, xcfl: float, ycfl: float, bc_scale: float,
             semantics: str = "gpu", threads: int = 0) -> np.ndarray:
    """CPU heat-diffusion oracle. semantics 'gpu' follows the reference's device path
    (next.border = curr.border*s, ref hw3code/gpuStencil.cu:115), 'cpu' its host path
    (ref hw3code/main.cu:184); see SURVEY quirk Q7."""
    g = np.ascontiguousarray(grid, dtype=np.float32).copy()
    gy, gx = g.shape
    scratch = np.empty_like(g)
    rc = host_lib().cmeh_heat_cpu(np_ptr(g), np_ptr(scratch), gx, gy, order, iters, xcfl, ycfl,
                                  bc_scale, 1 if semantics == "cpu" else 0, threads)
    if rc:
        raise ValueError("heat_cpu: bad argument")
    return g


# ---------------------------------------------------------------- ULPs
def max_ulp(a: np.ndarray, b: np.ndarray):
    """(max ULP distance, flat index) between two float32 or two float64 arrays."""
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b, dtype=a.dtype)
    if a.shape != b.shape:
        raise ValueError("max_ulp: shape mismatch")
    where = C.c_size_t(0)
    lib = host_lib()
    if a.dtype == np.float32:
        d = lib.cmeh_max_ulp_f32(np_ptr(a), np_ptr(b), a.size, C.addressof(where))
    elif a.dtype == np.float64:
        d = lib.cmeh_max_ulp_f64(np_ptr(a), np_ptr(b), a.size, C.addressof(where))
    else:
        raise TypeError("max_ulp: float32 or float64 only")
    return int(d), int(where.value)


# ---------------------------------------------------------------- hw4
def letter_histogram(text: np.ndarray, stride: int = 1, offset: int = 0) -> np.ndarray:
    t = np.ascontiguousarray(text, dtype=np.uint8)
    out = np.zeros(256, dtype=np.uint64)
    host_lib().cmeh_letter_histogram(np_ptr(t), t.size, stride, offset, np_ptr(out))
    return out


# ---------------------------------------------------------------- fpcode: IDX files
def idx_info(path) -> tuple[int, tuple[int, int, int]]:
    """(magic, (count, rows, cols)) of an IDX file (ref fpcode/utils/mnist.cpp:9-35)."""
    magic = C.c_int32(0)
    dims = (C.c_int64 * 3)()
    off = host_lib().cmeh_idx_info(str(path).encode(), C.addressof(magic), dims)
    if off < 0:
        raise ValueError(f"{path}: not a readable IDX file (code {off})")
    return int(magic.value), (int(dims[0]), int(dims[1]), int(dims[2]))


def idx_read(path, first_item: int, num_items: int, item_bytes: int) -> np.ndarray:
    out = np.empty(num_items * item_bytes, dtype=np.uint8)
    got = host_lib().cmeh_idx_read(str(path).encode(), first_item, num_items, np_ptr(out))
    if got != out.size:
        raise ValueError(f"{path}: short or bad IDX read (code {got})")
    return out


# ------------------------------------------------------------------ fpcode: native serial trainer
def mlp_train(X: np.ndarray, labels: np.ndarray, W1: np.ndarray, b1: np.ndarray, W2: np.ndarray, b2: np.ndarray,
              lr: float, reg: float, epochs: int, batch: int, loss_every: int = 0):
    """The serial fp64 trainer in C++/OpenMP (ref fpcode/neural_network.cpp:219-279).  X [N, D]
    float64, labels [N]; W1 [H, D], b1 [H], W2 [C, H], b2 [C] are float64 C-contiguous arrays
    updated IN PLACE.  Returns (iterations, losses) with one loss per loss_every iterations."""
    X = np.ascontiguousarray(X, dtype=np.float64)
    labels = np.ascontiguousarray(labels, dtype=np.int32)
    for a in (W1, b1, W2, b2):
        if a.dtype != np.float64 or not a.flags["C_CONTIGUOUS"]:
            raise ValueError("mlp_train: the weights must be float64 C-contiguous arrays (updated in place)")
    N, D = X.shape
    H, Cn = W1.shape[0], W2.shape[0]
    if W1.shape != (H, D) or b1.shape != (H,) or W2.shape != (Cn, H) or b2.shape != (Cn,) or labels.shape != (N,):
        raise ValueError("mlp_train: shapes do not fit together")
    nb = (N + batch - 1) // batch
    cap = (epochs * nb + loss_every - 1) // loss_every if loss_every > 0 else 0
    losses = np.zeros(max(cap, 1), dtype=np.float64)
    it = host_lib().cmeh_mlp_train(np_ptr(X), np_ptr(labels), N, D, H, Cn, np_ptr(W1), np_ptr(b1), np_ptr(W2), np_ptr(b2),
                                   lr, reg, epochs, batch, loss_every, np_ptr(losses), cap)
    if it < 0:
        raise ValueError("cmeh_mlp_train: bad argument" if it == -1 else "cmeh_mlp_train: a label is outside [0, C)")
    return int(it), losses[:cap]


def mlp_forward(X: np.ndarray, W1, b1, W2, b2, want_probs: bool = True):
    """Forward pass of the native CPU model: (probs [N, C] or None, pred [N] int32)."""
    X = np.ascontiguousarray(X, dtype=np.float64)
    N, D = X.shape
    H, Cn = W1.shape[0], W2.shape[0]
    probs = np.empty((N, Cn), dtype=np.float64) if want_probs else None
    pred = np.empty(N, dtype=np.int32)
    w = [np.ascontiguousarray(a, dtype=np.float64) for a in (W1, b1, W2, b2)]
    rc = host_lib().cmeh_mlp_forward(np_ptr(X), N, D, H, Cn, *(np_ptr(a) for a in w),
                                     np_ptr(probs) if want_probs else None, np_ptr(pred))
    if rc != 0:
        raise ValueError("cmeh_mlp_forward: bad argument")
    return probs, pred
