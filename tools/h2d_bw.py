import torch,time
for mb in (1,8,34,128):
    n=mb*1024*1024//4
    h=torch.empty(n).pin_memory(); d=torch.empty(n,device='cuda')
    for _ in range(3): d.copy_(h,non_blocking=True)
    torch.cuda.synchronize()
    e0=torch.cuda.Event(enable_timing=True); e1=torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): d.copy_(h,non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print(mb,'MB H2D GB/s', 10*n*4/ (e0.elapsed_time(e1)*1e-3)/1e9)
