import os, sys, json, torch
sys.path.insert(0, "/root/repo")
import pint_b200
from pint_b200 import dense, ops
def timed(fn, reps=10):
    for _ in range(3): fn()
    ts=[]
    for _ in range(reps):
        s,e=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); e.synchronize(); ts.append(s.elapsed_time(e))
    ts.sort(); return ts[len(ts)//2]
n=1<<28
for eb,bits in ((8,(21,21,21)),(8,(11,11,11)),(8,(13,13,13)),(8,(15,15,15,15))):
    L=pint_b200.Layout(eb,*bits); S=sum(bits)
    x=torch.empty(n,dtype=torch.uint64,device="cuda"); ops.fill_splitmix(x,1); da=dense.pack(L,x); ops.fill_splitmix(x,2); db=dense.pack(L,x); do=torch.empty_like(da)
    for threads,tile in ((0,0),(64,0),(32,0),(64,8192),(128,0),(256,0)):
        ops.set_option("dense_threads",threads); ops.set_option("dense_tile_bytes",tile)
        row=dict(layout=str(L),S=S,threads=threads,tile=tile)
        for op in ("add_wrap","add_unsigned_saturate","add_signed_saturate"):
            row[op]=round(3*n*S/8/timed(lambda: dense.binary(op,L,da,db,n,out=do))/1e6)
        print(json.dumps(row),flush=True)
    ops.set_option("dense_threads",0); ops.set_option("dense_tile_bytes",0)
    del x,da,db,do; torch.cuda.empty_cache()
