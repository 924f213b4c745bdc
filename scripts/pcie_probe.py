import torch, time
n=512<<20
h_in=torch.empty(n,dtype=torch.uint8).pin_memory(); h_out=torch.empty(320<<20,dtype=torch.uint8).pin_memory()
d_in=torch.empty(n,dtype=torch.uint8,device='cuda'); d_out=torch.empty(320<<20,dtype=torch.uint8,device='cuda')
s1,s2=torch.cuda.Stream(),torch.cuda.Stream()
def t(f,reps=5):
    f(); torch.cuda.synchronize(); t0=time.perf_counter()
    for _ in range(reps): f()
    torch.cuda.synchronize(); return (time.perf_counter()-t0)/reps
def h2d():
    with torch.cuda.stream(s1): d_in.copy_(h_in,non_blocking=True)
def d2h():
    with torch.cuda.stream(s2): h_out.copy_(d_out,non_blocking=True)
def both(): h2d(); d2h()
a=t(h2d); b=t(d2h); c=t(both)
print(f"H2D alone {n/a/1e9:.1f} GB/s, D2H alone {(320<<20)/b/1e9:.1f} GB/s, both: {c*1e3:.2f} ms -> H2D {n/c/1e9:.1f} GB/s + D2H {(320<<20)/c/1e9:.1f} GB/s; rays/s if PCIe-bound {16.777216e6/c/1e6:.0f} M")
