"""HBM read-only / write-only / copy bandwidth with torch kernels (CUDA events, best of 10): what a write-dominated
epilogue (FF1, conv3) can expect.  usage: python profiles/microbench/hbm_rw.py"""
import json, torch
n = 1 << 30                     # 1 Gi bf16 elements = 2 GiB
a = torch.empty(n, dtype=torch.bfloat16, device="cuda").normal_()
b = torch.empty_like(a)
def best(fn, reps=10):
    t = []
    for _ in range(reps):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        t.append(s.elapsed_time(e))
    return min(t)
res = {}
res["write_only_fill_GBs"] = 2 * n / best(lambda: b.fill_(1.0)) / 1e6
res["read_only_sum_GBs"] = 2 * n / best(lambda: a.view(torch.int16).max()) / 1e6
res["copy_read_plus_write_GBs"] = 4 * n / best(lambda: b.copy_(a)) / 1e6
print(json.dumps(res))
