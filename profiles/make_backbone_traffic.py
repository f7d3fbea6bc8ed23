"""profiles/backbone_dram_traffic.json from the ncu CSV of r2_profile_pass_final.sh step (3): DRAM read + write bytes
summed over the launches of one 512-image backbone pass, keyed by a hash of the kernel sources so that bench.py drops
the figure when the kernels change.  usage: make_backbone_traffic.py ncu.csv [launches_per_pass=108] [images=512]"""
import csv, hashlib, json, os, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path = sys.argv[1]
n_launch = int(sys.argv[2]) if len(sys.argv) > 2 else 105      # fused stem + 104 convolutions
images = int(sys.argv[3]) if len(sys.argv) > 3 else 512
rows = list(csv.DictReader(l for l in open(path) if not l.startswith("==")))
per = {}
for r in rows:
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1.0, "us": 1e3, "ms": 1e6}.get(unit, 1.0)
    per.setdefault(int(r["ID"]), {})[r["Metric Name"]] = v * scale
ids = sorted(per)[:n_launch]
rd = sum(per[i]["dram__bytes_read.sum"] for i in ids)
wr = sum(per[i]["dram__bytes_write.sum"] for i in ids)
ns = sum(per[i]["gpu__time_duration.sum"] for i in ids)
sources = ["retr_b200/csrc/gemm_sm100.cu", "retr_b200/csrc/stem_fused.cu"]
h = hashlib.sha256()
for f in sources:
    h.update(open(os.path.join(ROOT, f), "rb").read())
out = {"dram_bytes_per_image": (rd + wr) / images, "images_in_pass": images, "dram_read_gb": rd / 1e9,
       "dram_write_gb": wr / 1e9, "launches": len(ids), "sum_launch_ms_under_ncu": ns / 1e6,
       "from": "profiles/r2_profile_pass_final.sh step (3): ncu dram__bytes_read.sum + dram__bytes_write.sum over the "
               "%d launches of one %d-image backbone pass (bf16)" % (len(ids), images),
       "sources": sources, "sources_sha16": h.hexdigest()[:16]}
json.dump(out, open(os.path.join(ROOT, "profiles", "backbone_dram_traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
