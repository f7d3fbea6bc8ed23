"""Experiment: two contexts in flight (encode of batch i+1 under the decode loop of batch i)."""
import os, sys, json, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from retr_b200 import synthetic

B = int(os.environ.get("PB", "256"))
STEPS = int(os.environ.get("PSTEPS", "8"))
model, cfg = bench.build_model("bf16")
model.max_batch = B
model = model.to("cuda")
dev = torch.device("cuda", 0)
data = synthetic.make_batch(B, shard=0)
host = [data["t_img"], data["t_mask"].to(torch.uint8), data["g_img"], data["g_mask"].to(torch.uint8), data["loc"]]
host = [t.contiguous().pin_memory() for t in host]
devt = [t.to(dev) for t in host]
out_host = torch.empty(B, 128, dtype=torch.int64).pin_memory()
engs = model.pipeline_engines(B, 2)
cur_stream = torch.cuda.current_stream()

def timed(fn, steps, prime, drain):
    torch.cuda.synchronize()
    prime()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for e in engs: e.stream.wait_stream(cur_stream)
    for i in range(steps):
        fn(i)
    for e in engs: cur_stream.wait_stream(e.stream)
    e1.record()
    torch.cuda.synchronize()
    drain()
    return e0.elapsed_time(e1) / steps

# reference ids (serial, one context)
ids_ref, _ = engs[0].generate_host(*host, 101, 102, 128)
ids_ref = ids_ref.clone()

def serial_dev(i):
    engs[0].encode(*devt); engs[0].decode_greedy(101, 102, 128)
def serial_host(i):
    engs[0].generate_host(*host, 101, 102, 128, out=out_host)
def pipe_dev(i):
    engs[(i + 1) % 2].encode(*devt)
    engs[i % 2].decode_greedy(101, 102, 128)
def pipe_host(i):
    engs[(i + 1) % 2].encode_host(*host)
    engs[i % 2].decode_host(101, 102, 128, out=out_host)
nop = lambda: None
res = {}
for _ in range(2): serial_dev(0)
res["serial_dev_ms"] = timed(serial_dev, STEPS, nop, nop)
res["serial_host_ms"] = timed(serial_host, STEPS, nop, nop)
for cap in [int(c) for c in os.environ.get("PCAPS", "0,132,116,100,84").split(",")]:
    for e in engs: e.set_encode_cta_cap(cap)
    # warm both contexts (graph capture of context 1)
    engs[0].encode(*devt); pipe_dev(0); pipe_dev(1); engs[0].decode_greedy(101, 102, 128)
    ms = timed(pipe_dev, STEPS, lambda: engs[0].encode(*devt), lambda: engs[STEPS % 2].decode_greedy(101, 102, 128))
    msh = timed(pipe_host, STEPS, lambda: engs[0].encode_host(*host), lambda: engs[STEPS % 2].decode_host(101, 102, 128, out=out_host))
    same = bool(torch.equal(out_host, ids_ref))
    res[f"cap{cap}"] = {"dev_ms": ms, "host_ms": msh, "dev_expr_s": B / ms * 1e3, "host_expr_s": B / msh * 1e3, "ids_equal": same,
                        "phases0": engs[0].phase_ms(), "phases1": engs[1].phase_ms()}
print(json.dumps(res))
