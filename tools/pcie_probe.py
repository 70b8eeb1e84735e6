import torch, time
n=256<<20
h=torch.empty(n,dtype=torch.uint8).pin_memory(); d=torch.empty(n,dtype=torch.uint8,device='cuda')
for _ in range(2): d.copy_(h,non_blocking=True)
torch.cuda.synchronize()
e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
e0.record(); 
for _ in range(5): d.copy_(h,non_blocking=True)
e1.record(); torch.cuda.synchronize(); print('H2D GB/s', 5*n/e0.elapsed_time(e1)/1e6)
e0.record(); 
for _ in range(5): h.copy_(d,non_blocking=True)
e1.record(); torch.cuda.synchronize(); print('D2H GB/s', 5*n/e0.elapsed_time(e1)/1e6)
s2=torch.cuda.Stream(); h2=torch.empty(n,dtype=torch.uint8).pin_memory(); d2=torch.empty(n,dtype=torch.uint8,device='cuda')
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(5):
    d.copy_(h,non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2,non_blocking=True)
torch.cuda.synchronize(); dt=time.perf_counter()-t; print('duplex each GB/s', 5*n/dt/1e9)
