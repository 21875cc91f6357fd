"""Read-only streaming bandwidth on this GPU (torch reductions over 2 GiB) next to the copy figure: is the copy peak the
right denominator for kernels that only read?  Output of the round-2 run: profiles/r02_read_bw_probe.txt."""
import torch
dev="cuda:0"
for dt,name in ((torch.float32,"f32"),(torch.int32,"i32"),(torch.bfloat16,"bf16")):
    n=(2<<30)//torch.empty((),dtype=dt).element_size()
    x=torch.ones(n,dtype=dt,device=dev)
    for fn,fname in ((lambda t:t.sum(),"sum"),(lambda t:t.max(),"max")):
        for _ in range(3): fn(x)
        torch.cuda.synchronize()
        e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20): fn(x)
        e1.record(); torch.cuda.synchronize()
        ms=e0.elapsed_time(e1)/20
        print(name,fname,f"{ms*1e3:.1f} us  {2*(1<<30)/ms/1e6:.0f} GB/s")
    del x
a=torch.empty(1<<30,dtype=torch.uint8,device=dev); b=torch.empty_like(a)
for _ in range(3): b.copy_(a)
torch.cuda.synchronize()
e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): b.copy_(a)
e1.record(); torch.cuda.synchronize()
ms=e0.elapsed_time(e1)/20
print("copy 1GiB", f"{2*(1<<30)/ms/1e6:.0f} GB/s (read+write)")
