"""One B-sample greedy decode through the one-kernel decode step (RETR_DECODE_STACK=1).  Built with
RETR_EXTRA_NVCC_FLAGS=-DDS_PROFILE the kernel prints thread 0 / CTA 0's accumulated SM cycles per phase at step 60;
RETR_VERBOSE=1 prints the co-resident cluster count.  Usage: python profiles/decode_stack_profile.py [B]"""
import os, sys
os.environ.setdefault("RETR_DECODE_STACK", "1")
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from retr_b200 import synthetic
model, cfg = bench.build_model("bf16")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
model.max_batch = B
model = model.cuda()
s = tuple(x.to("cuda") for x in synthetic.pack(synthetic.make_batch(B, shard=0)))
with torch.no_grad():
    for _ in range(3):
        ids, n = model.greedy_native(s, 128, 101, 102)
torch.cuda.synchronize()
print("steps", n, "phase ms", model.engine(B).phase_ms())
