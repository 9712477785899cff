"""Per-region instruction / stall-sample breakdown of one kernel from an .ncu-rep captured with --import-source on.
usage: ncu_source_segments.py file.ncu-rep [dump_from dump_to]"""
import csv, io, subprocess, sys
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]; col = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
I = lambda r, k: int(r[col[k]] or 0)
tot = sum(I(r, "Instructions Executed") for r in data); stot = sum(I(r, "# Samples") for r in data)
print("total warp instructions", tot, "samples", stot)
prev = None; start = 0; seg = []
for k, r in enumerate(data):
    n = I(r, "Instructions Executed")
    if prev is None or abs(n - prev) > 0.15 * max(n, prev, 1):
        if prev is not None: seg.append((start, k - 1))
        start = k
    prev = n
seg.append((start, len(data) - 1))
for a, b in seg:
    n = sum(I(r, "Instructions Executed") for r in data[a:b + 1]); sm = sum(I(r, "# Samples") for r in data[a:b + 1])
    if n > 0.004 * tot or sm > 0.01 * stot:
        print(f"{a:5d}-{b:5d} len={b-a+1:4d} inst={n/tot*100:5.1f}% samples={sm/stot*100:5.1f}% per-instr={I(data[a],'Instructions Executed')} thr={data[a][col['Avg. Threads Executed']]}  {data[a][col['Source']][:50]}")
if len(sys.argv) > 3:
    for k in range(int(sys.argv[2]), int(sys.argv[3]) + 1):
        r = data[k]
        print(k, r[col["Source"]][:72].ljust(72), r[col["# Samples"]], r[col["Avg. Threads Executed"]])
