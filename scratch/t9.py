import sys, numpy as np
raw = np.fromfile(sys.argv[1], dtype=np.uint64); n = int(raw[0]); rec = raw[8:8+8*n].reshape(n, 8)
idx = np.where(rec[:, 0] == 9)[0]
idx = idx[len(idx)//2: len(idx)//2 + 12]
for i in idx:
    a, b = rec[i], rec[i+1]
    assert int(b[0]) == 10
    t0 = int(a[3]); ts = [int(a[4]), int(a[5]), int(a[6]), int(a[7]), int(b[1]), int(b[2]), int(b[3]), int(b[4]), int(b[5]), int(b[6])]
    d = [(ts[0]-t0)/1000] + [(ts[k]-ts[k-1])/1000 for k in range(1, 10)]
    print(f"cta {int(a[2]):3d}: pdl_wait {d[0]:5.2f} | panels " + " ".join(f"{x:4.2f}" for x in d[1:9]) + f" | reduce {d[9]:4.2f} | after-wait total {(ts[9]-ts[0])/1000:5.2f}")
