import torch, time
B=512; n=B*752*480
p=torch.empty(n,dtype=torch.uint8,pin_memory=True); d=torch.empty(n,dtype=torch.uint8,device='cuda')
o=torch.empty(37750784,dtype=torch.uint8,device='cuda'); op=torch.empty(37750784,dtype=torch.uint8,pin_memory=True)
s2=torch.cuda.Stream()
for both in (0,1):
    for _ in range(3): d.copy_(p,non_blocking=True)
    torch.cuda.synchronize(); t=time.perf_counter()
    for _ in range(10):
        d.copy_(p,non_blocking=True)
        if both:
            with torch.cuda.stream(s2): op.copy_(o,non_blocking=True)
        torch.cuda.synchronize()
    dt=(time.perf_counter()-t)/10
    print("both" if both else "h2d only", round(dt*1e3,3),"ms", round(n/dt/1e9,1),"GB/s")
for chunk in (16,64):
    c=chunk*752*480
    torch.cuda.synchronize(); t=time.perf_counter()
    for _ in range(10):
        for i in range(0,n,c): d[i:i+c].copy_(p[i:i+c],non_blocking=True)
        torch.cuda.synchronize()
    dt=(time.perf_counter()-t)/10
    print("chunk",chunk, round(dt*1e3,3),"ms", round(n/dt/1e9,1),"GB/s")
