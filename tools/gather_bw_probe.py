import sys, os
sys.path.insert(0, os.getcwd())
import torch, dgi_b200
from dgi_b200 import ops
dev="cuda:0"
for B,S in ((4,4096),(8,4096),(4,32768),(16,1024)):
    Hkv,D,bs=8,128,16
    nb=(S+bs-1)//bs
    mb=B*nb+8
    k=torch.randn((mb,Hkv,bs,D),device=dev).bfloat16(); v=torch.randn_like(k)
    t=torch.randperm(mb,device=dev)[:B*nb].to(torch.int32).view(B,nb)
    for _ in range(3): ops.kv_gather_dense(k,v,t,S)
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): ops.kv_gather_dense(k,v,t,S)
    e1.record(); torch.cuda.synchronize()
    ms=e0.elapsed_time(e1)/20
    byt=2*2*B*S*Hkv*D*2
    print(B,S,f"{ms*1e3:.1f} us {byt/ms/1e6:.0f} GB/s (read+write; includes the two torch.empty allocations)")
