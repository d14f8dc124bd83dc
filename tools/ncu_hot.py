"""Hot SASS regions of an .ncu-rep by executed-instruction share (run-length groups of equal execution count)."""
import csv
import subprocess
import sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
ie, src, ss = hdr.index("Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
body = rows[2:]
tot = sum(int(r[ie]) for r in body)
tsamp = sum(int(r[ss]) for r in body)
print("total warp instructions", tot, "SASS lines", len(body), "samples", tsamp)
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01
grp, prev, start = [], None, 0
for i, r in enumerate(body):
    n = int(r[ie])
    if prev is None or abs(n - prev) > 0.02 * max(prev, 1):
        if prev is not None:
            grp.append((start, i - 1, prev))
        start, prev = i, n
grp.append((start, len(body) - 1, prev))
for a, b, n in grp:
    samples = sum(int(body[k][ss]) for k in range(a, b + 1))
    if n * (b - a + 1) > thr * tot or samples > thr * tsamp:
        ops = []
        for k in range(a, b + 1):
            t = body[k][src].split()
            ops.append(t[1] if t[0].startswith("@") else t[0])
        print(f"[{a:4d}-{b:4d}] len={b - a + 1:3d} exec={n:9d} inst%={n * (b - a + 1) / tot * 100:5.1f} samp%={samples / max(tsamp, 1) * 100:5.1f}", " ".join(ops)[:200])
