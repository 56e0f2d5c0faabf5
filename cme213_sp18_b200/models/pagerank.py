"""PageRank workload (ref hw2code/pagerank.cu:59-136 device_graph_iterate, main_q2.cu:151-233)."""
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
