import torch
a=torch.empty(56_400_000//2,dtype=torch.bfloat16,device='cuda'); b=torch.empty_like(a)
big=torch.empty(300_000_000,dtype=torch.uint8,device='cuda')
ts=[]
for i in range(20):
    big.zero_()   # flush L2
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record(); b.copy_(a); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1)*1e3)
ts.sort(); print("copy 56.4MB->56.4MB (112.8 MB traffic): median %.1f us min %.1f us => %.0f GB/s"%(ts[10],ts[0],112.8e6/ts[10]/1e3))
ts=[]
for i in range(20):
    big.zero_()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record(); a.mul_(1.0001); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1)*1e3)
ts.sort(); print("in-place scale of 56.4MB (112.8 MB traffic): median %.1f us min %.1f us => %.0f GB/s"%(ts[10],ts[0],112.8e6/ts[10]/1e3))
