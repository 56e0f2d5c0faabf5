"""Command-line drivers: python -m cme213_sp18_b200.cli.<mlp|heat|pagerank|shift|cipher|radix|parity|gemm_bench>."""
