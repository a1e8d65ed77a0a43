import torch, time
for mb in (0.2, 0.65, 1.3, 1.8, 2.65):
    n=int(mb*1e6)
    h=torch.empty(n,dtype=torch.uint8).pin_memory(); d=torch.empty(n,dtype=torch.uint8,device='cuda')
    for _ in range(5): d.copy_(h,non_blocking=True)
    torch.cuda.synchronize(); t=time.perf_counter()
    for _ in range(50): d.copy_(h,non_blocking=True); torch.cuda.synchronize()
    a=(time.perf_counter()-t)/50
    t=time.perf_counter()
    for _ in range(50): h.copy_(d,non_blocking=True); torch.cuda.synchronize()
    b=(time.perf_counter()-t)/50
    print(f"{mb} MB: H2D {a*1e6:.0f} us ({n/a/1e9:.1f} GB/s)  D2H {b*1e6:.0f} us ({n/b/1e9:.1f} GB/s)")
