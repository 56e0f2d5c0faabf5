"""PageRank workload (ref hw2code/pagerank.cu:59-136 device_graph_iterate, main_q2.cu:151-233):
the reference's byte model, the one-launch iteration on one GPU, and the graph set-up of the
column-blocked propagation used once the contribution vector no longer fits in L2."""
from __future__ import annotations

import torch

from ..ops import device as dev


def pagerank_bytes(num_nodes: int, avg_edges: int, iterations: int) -> int:
    """The reference's byte model (ref hw2code/main_q2.cu:224-226), kept for comparable GB/s."""
    n, e, it = num_nodes, avg_edges, iterations
    return 2 * n * it * (1 + e) * 4 + n * it * (1 + 2 * e) * 4


def pagerank_device(row_ptr: torch.Tensor, edges: torch.Tensor, rank0: torch.Tensor,
                    inv_deg: torch.Tensor, iterations: int) -> torch.Tensor:
    """`iterations` propagation steps on one GPU, ping-ponging two contrib buffers; returns the
    final rank vector.  No host synchronisation inside."""
    contrib_a, contrib_b = torch.empty_like(rank0), torch.empty_like(rank0)
    rank = torch.empty_like(rank0)
    dev.pagerank_iterate(row_ptr, edges, rank0, inv_deg, contrib_a, contrib_b, rank, iterations)
    return rank


def column_blocked_csr(row_ptr_local: torch.Tensor, edges: torch.Tensor, num_nodes: int, block_nodes: int):
    """Regroup a CSR slice by L2-sized blocks of SOURCE nodes: returns (blk_ptr, edges_b, nblocks) with
    blk_ptr int32 [nblocks * nrows + 1] and the edges of (block b, row i) at
    edges_b[blk_ptr[b * nrows + i] : blk_ptr[b * nrows + i + 1]], in their original relative order.
    Done once per graph on the device, outside every timed loop (torch sort / scan: set-up work)."""
    nrows = row_ptr_local.numel() - 1
    nb = (num_nodes + block_nodes - 1) // block_nodes
    E = edges.numel()
    if E >= 2**31:
        raise ValueError("column_blocked_csr: a slice holds at most 2^31 - 1 edges")
    deg = (row_ptr_local[1:] - row_ptr_local[:-1]).to(torch.int64)
    rows = torch.repeat_interleave(torch.arange(nrows, device=edges.device, dtype=torch.int64), deg)
    # uint32 node ids held in an int32 tensor: widen before dividing
    src = edges.to(torch.int64) & 0xFFFFFFFF
    key = (src // block_nodes) * nrows + rows
    del src, rows, deg
    order = torch.argsort(key, stable=True)
    edges_b = edges[order].contiguous()
    del order
    counts = torch.bincount(key, minlength=nb * nrows)
    del key
    blk_ptr = torch.zeros(nb * nrows + 1, dtype=torch.int32, device=edges.device)
    blk_ptr[1:] = torch.cumsum(counts, 0).to(torch.int32)
    return blk_ptr, edges_b, nb
