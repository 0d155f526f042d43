import torch, sys
m=n=k=8192
torch.manual_seed(0)
a=torch.randn(m,k,device="cuda").bfloat16(); b=torch.randn(n,k,device="cuda").bfloat16(); c=torch.empty(m,n,device="cuda",dtype=torch.bfloat16)
for _ in range(int(sys.argv[1])): torch.matmul(a,b.t(),out=c)
torch.cuda.synchronize(); print("ok")
