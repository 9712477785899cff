"""Per-launch table from an `ncu --csv` log (any --metrics list): kernel, then one column per metric.
usage: python tools/ncu_kernels.py gpurun_out/x.csv"""
import csv, sys
from collections import OrderedDict
lines = [ln for ln in open(sys.argv[1]) if ln.startswith('"')]
rows = OrderedDict()
for r in csv.DictReader(lines):
    k = (r["ID"], r["Kernel Name"].split("(")[0].split("::")[-1])
    rows.setdefault(k, OrderedDict())[r["Metric Name"]] = (r["Metric Value"], r["Metric Unit"])
for (i, name), m in rows.items():
    print(i, name, " ".join(f"{k.split('.')[0].replace('sm__','').replace('smsp__','')}={v[0]}{v[1] if v[1] not in ('%','') else ''}" for k, v in m.items()))
