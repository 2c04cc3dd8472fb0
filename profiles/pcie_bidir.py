import torch, time
n=16*1024*1024
h1=torch.empty(n,dtype=torch.float64).pin_memory(); h2=torch.empty(n,dtype=torch.float64).pin_memory()
d1=torch.empty(n,dtype=torch.float64,device='cuda'); d2=torch.empty(n,dtype=torch.float64,device='cuda')
s1,s2=torch.cuda.Stream(),torch.cuda.Stream()
def t(f,reps=5):
    f(); torch.cuda.synchronize(); t0=time.perf_counter()
    for _ in range(reps): f()
    torch.cuda.synchronize(); return (time.perf_counter()-t0)/reps*1e3
def h2d():
    with torch.cuda.stream(s1): d1.copy_(h1,non_blocking=True)
def d2h():
    with torch.cuda.stream(s2): h2.copy_(d2,non_blocking=True)
def both(): h2d(); d2h()
def chunked(K=64):
    c=n//K
    for i in range(K):
        with torch.cuda.stream(s1): d1[i*c:(i+1)*c].copy_(h1[i*c:(i+1)*c],non_blocking=True)
        with torch.cuda.stream(s2): h2[i*c:(i+1)*c].copy_(d2[i*c:(i+1)*c],non_blocking=True)
print('h2d ms',t(h2d),'d2h ms',t(d2h),'both ms',t(both),'chunked64 both ms',t(chunked))
