cd $GRAFT_REPO_ROOT
timeout 120 python -m pytest tests/test_gpu_kernels.py -q -m gpu --timeout 60 -x -k "gemm_bf16_kmajor or fwd1" 2>&1 | tail -15
timeout 120 python -c "
import torch, sys
sys.path.insert(0,'.')
from cme213_sp18_b200.ops import device as dev
from cme213_sp18_b200.utils.timing import time_cuda_ms
for (M,N,K) in [(65536,8192,785),(8192,8192,785),(8192,8192,8192)]:
    ld=(K+63)//64*64
    a=torch.randn((M,ld),device='cuda').bfloat16(); b=torch.randn((N,ld),device='cuda').bfloat16()
    out=torch.empty((M,N),dtype=torch.bfloat16,device='cuda')
    ms=time_cuda_ms(lambda: dev.gemm_bf16(a[:,:K],b[:,:K],out),3,10)
    ref=time_cuda_ms(lambda: torch.matmul(a[:,:K],b[:,:K].T),3,10)
    want=(a[:2048,:K].float()@b[:2048,:K].float().T)
    err=float((out[:2048,:2048].float()-want).abs().max()/want.abs().max())
    print(M,N,K,'pair-kernel ms',round(ms,4),'TFLOPs',round(2*M*N*K/ms/1e9),'cublas ms',round(ref,4),'err',err)
"
