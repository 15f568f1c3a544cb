"""Write-only and copy HBM bandwidth probe (context for the materialisation kernel, whose reads come from L2)."""
import json, torch
n = 16 << 30
x = torch.empty(n, dtype=torch.uint8, device="cuda")
y = torch.empty(n, dtype=torch.uint8, device="cuda")
res = {}
def timeit(fn, reps=5):
    best = 1e9
    for _ in range(reps):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best
fn = lambda: x.fill_(7)
fn(); res["fill_GBps"] = n / timeit(fn) / 1e6
fn = lambda: x.zero_()
fn(); res["memset_GBps"] = n / timeit(fn) / 1e6
fn = lambda: y.copy_(x)
fn(); res["copy_rw_GBps"] = 2 * n / timeit(fn) / 1e6
x32 = x.view(torch.int32)
fn = lambda: x32.sum()
fn(); res["read_GBps"] = n / timeit(fn) / 1e6
print(json.dumps(res))
