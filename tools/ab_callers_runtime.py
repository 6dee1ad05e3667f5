import json, os, sys, torch
sys.path.insert(0, "/root/repo")
import pint_b200
from pint_b200 import ops
def timed(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); e1.synchronize()
    return e0.elapsed_time(e1) / reps
n = 1 << 29
for eb, bits, dt in ((4, (7, 9, 3), torch.uint32), (4, (5, 6, 5, 3), torch.uint32), (8, (10, 10, 10, 10), torch.uint64)):
    L = pint_b200.Layout(eb, *bits)
    x = torch.empty(n, dtype=dt, device="cuda"); ops.fill_splitmix(x, 3)
    o = torch.empty_like(x)
    row = {"layout": repr(L)}
    for op in ("add_signed_saturate", "min_signed"):
        ms = timed(lambda: ops.binary_scalar(op, L, x, 0x01234567, out=o)); row["binary_scalar " + op] = round(2 * eb * n / ms / 1e6, 1)
        ms = timed(lambda: ops.binary_pairs(op, L, x, out=o[: n // 2])); row["binary_pairs " + op] = round(3 * eb * (n // 2) / ms / 1e6, 1)
    print(json.dumps(row), flush=True)
