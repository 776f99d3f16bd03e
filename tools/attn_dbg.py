import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import ops as O
from ptv3_b200 import ops, _lib
lib = _lib.load()
def rel(a,b):
    a,b=a.float().cpu(),b.float().cpu(); return ((a-b).norm()/b.norm()).item()
for counts,K,H in [([128],128,1),([64],64,1),([1024],1024,2),([700,2300,64,1100],1024,4)]:
    torch.manual_seed(0)
    n,C=sum(counts),H*16
    qkv=(torch.randn(n,3*C)*1.5).bfloat16()
    offset=np.cumsum(counts)
    order=torch.cat([torch.randperm(c)+s for c,s in zip(counts,np.concatenate([[0],offset[:-1]]))])
    inverse=torch.empty_like(order); inverse[order]=torch.arange(n)
    ref=O.window_attention(qkv.float(),order,inverse,offset,H,K)
    w,nw=ops.build_windows(torch.from_numpy(offset).cuda(),K,n)
    for variant in (0,1):
        lib.ptv3_debug_set(1,variant)
        got=ops.attn_window(qkv.cuda(),order.cuda(),w,nw,H,K,algo=3)
        torch.cuda.synchronize()
        print(counts,K,H,"variant",variant,"rel",rel(got,ref), "nan", bool(torch.isnan(got.float()).any()))
    lib.ptv3_debug_set(1,0)
