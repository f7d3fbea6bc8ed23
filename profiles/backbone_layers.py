"""Per-convolution table of one backbone pass from a compact launch list (profiles/*_launches.csv[.gz] as written by
summarize_launches.py): time, algorithmic TFLOP/s and GB/s, and the ratio to max(tensor time, HBM time) at the measured
peaks.  usage: backbone_layers.py launches.csv[.gz] n_images"""
import gzip, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from analyze_backbone import convs


def main():
    path, nimg = sys.argv[1], int(sys.argv[2])
    peaks = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")))
    tf_peak = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1367.0)))
    bw_peak = float(peaks.get("hbm_gbs", 6538.0))
    rows = []
    for l in (gzip.open(path, "rt") if path.endswith(".gz") else open(path)):
        if l.startswith("#") or l.startswith("id,"):
            continue
        a = l.rstrip("\n").split(",")
        rows.append((",".join(a[1:-3]), float(a[-1])))
    first = next(i for i, r in enumerate(rows) if r[0].startswith(("stem_im2col", "stem_fused")))
    seq = [r for r in rows[first:] if r[0].startswith(("stem_im2col", "stem_fused", "gemm_tcgen05", "maxpool"))]
    n_im2col = sum(1 for r in seq[:3] if r[0].startswith("stem_im2col"))
    n_fused = sum(1 for r in seq[:3] if r[0].startswith("stem_fused"))
    if n_fused:      # fused stem (conv + BN + ReLU + max-pool in one kernel, one launch per image set): one table row
        seq = [("-", 0.0), ("stem_fused x%d" % n_fused, sum(r[1] for r in seq[:n_fused])), ("-", 0.0)] + seq[n_fused:]
    else:
        seq = [("stem_im2col x%d" % n_im2col, sum(r[1] for r in seq[:n_im2col]))] + seq[n_im2col:]
    tot = floor = 0.0
    print("%-12s %-28s %9s %9s %8s %9s %6s" % ("conv", "kernel", "us", "TFLOP/s", "GB/s", "floor us", "ratio"))
    for (name, shape), (kern, ns) in zip(convs(), seq):
        us = ns / 1000
        tot += us
        if shape:
            cin, cout, k, oh, ow = shape
            M = nimg * oh * ow
            fl = 2.0 * M * cout * cin * k * k
            by = 2.0 * (M * cin + M * cout * (2 if name.endswith("c3") else 1))
            fl_us = max(fl / (tf_peak * 1e12), by / (bw_peak * 1e9)) * 1e6
            floor += fl_us
            print("%-12s %-28s %9.1f %9.1f %8.0f %9.1f %6.2f" % (name, kern[:28], us, fl / us / 1e6, by / us / 1e3, fl_us, us / fl_us))
        else:
            print("%-12s %-28s %9.1f" % (name, kern[:28], us))
    print("pass total %.2f ms; sum of per-layer floors %.2f ms (peaks: %.0f TFLOP/s, %.0f GB/s)" % (tot / 1e3, floor / 1e3, tf_peak, bw_peak))


if __name__ == "__main__":
    main()
