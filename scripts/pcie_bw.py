import torch, time
dev=torch.device('cuda:0')
for mb in (1, 4, 64):
    n=mb<<20
    h=torch.empty(n,dtype=torch.uint8).pin_memory(); d=torch.empty(n,dtype=torch.uint8,device=dev)
    for _ in range(3): d.copy_(h,non_blocking=True)
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    K=max(4, 512//mb)
    e0.record()
    for _ in range(K): d.copy_(h,non_blocking=True)
    e1.record(); torch.cuda.synchronize()
    print("H2D %d MB: %.1f GB/s"%(mb, n*K/e0.elapsed_time(e1)/1e6))
