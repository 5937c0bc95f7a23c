import torch
for (M,N,K) in ((8192,152064,3584),(8192,3584,152064),(152064,3584,8192)):
    A=torch.randn(M,K,device='cuda').bfloat16(); B=torch.randn(N,K,device='cuda').bfloat16()
    out=torch.empty(M,N,device='cuda',dtype=torch.bfloat16)
    torch.matmul(A,B.t(),out=out); torch.cuda.synchronize()
