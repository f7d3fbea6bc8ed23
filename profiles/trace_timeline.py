"""Timeline of one decode step from a -DRETR_TRACE build (see csrc/common.cuh).

    RETR_EXTRA_NVCC_FLAGS=-DRETR_TRACE python -m retr_b200.build --force
    RETR_TRACE_FILE=gpurun_out/trace.bin python bench.py --steps 1 --warmup 3 --no-cpu-baseline
    python profiles/trace_timeline.py gpurun_out/trace.bin [step]

Records: first/last CTA of small_gemm (1), self-attention (2), cross-attention (3), greedy_update (6), ffn_fused (8).
Prints, for the chosen step, every traced launch: group (from the output pointer), kernel, grid, start of its first
CTA, end of its last CTA (us from the step's start), and a per-kernel-class busy summary.
"""
import sys
from collections import defaultdict

import numpy as np

NAMES = {1: "small_gemm", 2: "self_attn", 3: "cross_attn", 6: "greedy_update", 8: "ffn_fused", 11: "out_ln_proj"}


def main():
    raw = np.fromfile(sys.argv[1], dtype=np.uint64)
    n = int(raw[0])
    rec = raw[8:8 + 8 * n].reshape(n, 8)
    step = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    upd = rec[rec[:, 0] == 6]
    upd = upd[np.argsort(upd[:, 3])]
    t_lo, t_hi = int(upd[step - 1, 4]), int(upd[step, 4])
    sel = rec[(rec[:, 3] >= t_lo) & (rec[:, 3] < t_hi) & np.isin(rec[:, 0], [1, 2, 3, 8, 11])]
    # group launches: key = (id, grid, tag) within the step, ordered by time; first+last CTA records pair up
    launches = defaultdict(list)
    for r in sel:
        launches[(int(r[0]), int(r[1]), int(r[6]))].append(r)
    rows = []
    for (kid, grid, tag), rs in launches.items():
        rs = sorted(rs, key=lambda r: int(r[3]))
        # several launches may share (id, grid, tag) (e.g. att buffer reused by self/cross has different id) -> split
        # by pairing consecutive first-CTA records
        firsts = [r for r in rs if int(r[2]) == 0]
        lasts = [r for r in rs if int(r[2]) == grid - 1]
        for i, f in enumerate(firsts):
            l = lasts[i] if i < len(lasts) else f
            rows.append((int(f[3]) - t_lo, max(int(l[4]), int(f[4])) - t_lo, kid, grid, tag, int(f[4]) - int(f[3])))
    rows.sort()
    tags = sorted({r[4] for r in rows})
    print(f"step {step}: {(t_hi - t_lo) / 1000:.1f} us, {len(rows)} traced launches")
    busy = defaultdict(float)
    for t0, t1, kid, grid, tag, cta0 in rows:
        busy[(NAMES[kid], grid)] += (t1 - t0) / 1000
        print(f"  {t0 / 1000:8.1f} -> {t1 / 1000:8.1f} us  ({(t1 - t0) / 1000:6.1f})  cta0 {cta0 / 1000:5.1f}  "
              f"{NAMES[kid]:<12} grid {grid:<5} tag {tag:#x}")
    print("sum of launch spans per class (us):")
    for k, v in sorted(busy.items(), key=lambda kv: -kv[1]):
        print(f"  {k[0]:<12} grid {k[1]:<5} {v:8.1f}")


if __name__ == "__main__":
    main()
