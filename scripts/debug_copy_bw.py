import torch
M=32768
zx=torch.randn(M,3352,device='cuda').bfloat16()
out=torch.empty(M,1792,device='cuda',dtype=torch.bfloat16)
src_c=torch.randn(M,1792,device='cuda').bfloat16()
def t(f,n=20):
    for _ in range(3): f()
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)/n*1e3
big=torch.empty(256*1024*1024,device='cuda',dtype=torch.uint8)
def flush(): big.zero_()
a=t(lambda: out.copy_(zx[:,1536:1536+1792]))
b=t(lambda: out.copy_(src_c))
print(f"strided slice -> contiguous: {a:.1f} us = {2*M*1792*2/a/1e6:.2f} TB/s ; contiguous copy {b:.1f} us = {2*M*1792*2/b/1e6:.2f} TB/s")
