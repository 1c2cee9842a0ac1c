import torch,time
x=torch.empty(1<<30,dtype=torch.uint8).pin_memory(); d=torch.empty(1<<30,dtype=torch.uint8,device='cuda')
for name,(a,b) in {'h2d':(d,x),'d2h':(x,d)}.items():
    a.copy_(b,non_blocking=True); torch.cuda.synchronize(); t=time.perf_counter()
    for _ in range(3): a.copy_(b,non_blocking=True)
    torch.cuda.synchronize(); print(name, 3*(1<<30)/(time.perf_counter()-t)/1e9,'GB/s')
s1=torch.cuda.Stream(); s2=torch.cuda.Stream(); y=torch.empty(1<<30,dtype=torch.uint8).pin_memory(); e=torch.empty(1<<30,dtype=torch.uint8,device='cuda')
torch.cuda.synchronize(); t=time.perf_counter()
for _ in range(3):
    with torch.cuda.stream(s1): d.copy_(x,non_blocking=True)
    with torch.cuda.stream(s2): y.copy_(e,non_blocking=True)
torch.cuda.synchronize(); print('bidir each', 3*(1<<30)/(time.perf_counter()-t)/1e9,'GB/s')
