"""Throughput of retr_preprocess (one launch per batch of raw COCO-sized images) next to the reference's PIL /
torchvision pipeline on one host core.  Prints one JSON line."""
import json, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import preprocess_oracle as P
from retr_b200 import native
from retr_b200.data_utils import preprocess_batch
B = 256
scenes = [P.synthetic_scene(1000 + i) for i in range(B)]
images, boxes = [s[0] for s in scenes], [s[1] for s in scenes]
for _ in range(2):
    out = preprocess_batch(images, boxes)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(3):
    out = preprocess_batch(images, boxes)
torch.cuda.synchronize()
e2e = (time.perf_counter() - t0) / 3
# kernel alone: inputs resident
lib = native.lib()
raw = torch.from_numpy(np.concatenate([i.reshape(-1) for i in images])).cuda()
offs = torch.from_numpy(np.cumsum([0] + [i.size for i in images[:-1]]).astype(np.int64)).cuda()
hw = torch.tensor([i.shape[:2] for i in images], dtype=torch.int32).cuda()
bb = torch.tensor(boxes, dtype=torch.float64).cuda()
o = [torch.empty(B, 3, 224, 224, device="cuda"), torch.empty(B, 224, 224, dtype=torch.uint8, device="cuda"),
     torch.empty(B, 3, 224, 224, device="cuda"), torch.empty(B, 224, 224, dtype=torch.uint8, device="cuda"),
     torch.empty(B, 5, device="cuda")]
def launch():
    native.check(lib.retr_preprocess(native.ptr(raw), native.ptr(offs), native.ptr(hw), native.ptr(bb), B, 224,
                                     *[native.ptr(t) for t in o], native.stream_ptr(torch.device("cuda"))))
launch(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    launch()
e1.record(); torch.cuda.synchronize()
k_ms = e0.elapsed_time(e1) / 10
torch.set_num_threads(1)
t0 = time.perf_counter()
for img, b in scenes[:32]:
    P.preprocess_sample(img, b)
cpu = (time.perf_counter() - t0) / 32
in_bytes = raw.numel()
out_bytes = sum(t.numel() * t.element_size() for t in o)
print(json.dumps({"batch": B, "kernel_ms": k_ms, "samples_per_s_kernel": B / (k_ms / 1e3),
                  "bytes_in_out_per_launch": [in_bytes, out_bytes],
                  "algorithmic_GBps": (in_bytes + out_bytes) / (k_ms / 1e3) / 1e9,
                  "samples_per_s_end_to_end_from_numpy": B / e2e,
                  "reference_pipeline_samples_per_s_one_core": 1 / cpu}))
