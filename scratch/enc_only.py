"""Profiling target: encode() only (backbone over both image sets in one pass + encoder), B=256."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from retr_b200 import synthetic
B = 256
model, cfg = bench.build_model("bf16"); model.max_batch = B; model = model.to("cuda")
data = synthetic.make_batch(B, shard=0)
dev = [data["t_img"].cuda(), data["t_mask"].to(torch.uint8).cuda(), data["g_img"].cuda(), data["g_mask"].to(torch.uint8).cuda(), data["loc"].cuda()]
eng = model.engine(B)
for _ in range(int(os.environ.get("N_ENC", "2"))):
    eng.encode(*dev)
torch.cuda.synchronize()
print("phases", eng.phase_ms())
