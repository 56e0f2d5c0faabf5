"""fp64 GEMM timing on a few shapes (CUDA events, 20 calls back to back)."""
import sys
from pathlib import Path
import torch
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from cme213_sp18_b200.ops import device as dev
from cme213_sp18_b200.utils.timing import time_cuda_ms

def bench(M, N, K, ta=False, tb=False):
    A = torch.randn((K, M) if not ta else (M, K), dtype=torch.float64, device="cuda")   # column-major buffers
    B = torch.randn((N, K) if not tb else (K, N), dtype=torch.float64, device="cuda")
    C = torch.zeros((N, M), dtype=torch.float64, device="cuda")
    lda = K if ta else M
    ldb = N if tb else K
    def call():
        for _ in range(20):
            dev.dgemm(ta, tb, M, N, K, 1.0, A, lda, B, ldb, 0.0, C, M)
    ms = time_cuda_ms(call, warmup=1, iters=5) / 20
    a2, b2 = torch.randn(M, K, dtype=torch.float64, device="cuda"), torch.randn(K, N, dtype=torch.float64, device="cuda")
    def lib():
        for _ in range(20):
            torch.matmul(a2, b2)
    ms_lib = time_cuda_ms(lib, warmup=1, iters=5) / 20
    print(f"({M},{N},{K}) ta={int(ta)} tb={int(tb)}: {ms*1e3:8.1f} us = {2.0*M*N*K/ms/1e9:6.1f} TFLOP/s | cuBLAS {ms_lib*1e3:8.1f} us = {2.0*M*N*K/ms_lib/1e9:6.1f}")

if len(sys.argv) > 1 and sys.argv[1] == "one":
    bench(1000, 800, 784)
else:
    for s in ((1000, 800, 784), (1000, 784, 800), (4096, 4096, 4096), (8192, 8192, 784), (2048, 2048, 2048), (1000, 100, 784), (512, 512, 512)):
        bench(*s)
    bench(1000, 784, 800, False, True)
    bench(1000, 800, 10, True, False)
