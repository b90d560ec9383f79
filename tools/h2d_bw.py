import torch,time
x=torch.empty(610*1024*1024//4,dtype=torch.float32).pin_memory(); d=torch.empty_like(x,device='cuda')
for n in (1,7,35):
    ch=x.numel()//n
    torch.cuda.synchronize(); t=time.time()
    for r in range(5):
        for i in range(n): d[i*ch:(i+1)*ch].copy_(x[i*ch:(i+1)*ch],non_blocking=True)
    torch.cuda.synchronize(); dt=(time.time()-t)/5
    print(n,'chunks H2D GB/s',x.numel()*4/dt/1e9)
y=torch.empty(64*1024*1024//4,dtype=torch.float32).pin_memory()
s2=torch.cuda.Stream()
torch.cuda.synchronize(); t=time.time()
for r in range(5):
    d.copy_(x,non_blocking=True)
    with torch.cuda.stream(s2): y.copy_(d[:y.numel()],non_blocking=True)
torch.cuda.synchronize(); dt=(time.time()-t)/5
print('H2D with concurrent D2H GB/s',x.numel()*4/dt/1e9)
