import os, sys, torch
sys.path.insert(0, "/root/repo")
import pint_b200
from pint_b200 import ops
def timed(fn, reps=10):
    for _ in range(3): fn()
    ts=[]
    for _ in range(reps):
        s,e=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); e.synchronize(); ts.append(s.elapsed_time(e))
    ts.sort(); return ts[len(ts)//2]
n=1<<30
L=pint_b200.Layout(4,8,8,8,8)
a=torch.empty(2*n,dtype=torch.uint32,device="cuda"); ops.fill_splitmix(a,1)
for cps in (8, 4, 16):
    ops.set_option("ctas_per_sm", cps)
    for rep in range(2):
        for mode in (1,2):
            ops.set_option("vec256",mode)
            ms=timed(lambda: ops.pairs_reduce("add_signed_saturate",L,a))
            print("ctas_per_sm(opt)",cps,"vec256",mode, round(8*n/ms/1e6),"GB/s", flush=True)
ops.set_option("vec256",1)
