"""Dev script (GPU box): kernel timeline of one frame graph via torch.profiler (CUPTI activity records).

Prints, for the last profiled frame, every kernel's start offset and duration in microseconds, so that the
overlap of the graph branches can be read off. CUPTI adds per-kernel overhead: use for the *shape* of the
timeline, not for absolute numbers. usage: python tools/timeline.py [cols rows]"""
import json, os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
from another_cuda_sift_b200 import CudaSiftConfig, Detector
from another_cuda_sift_b200.synth import synth_frame

cols, rows = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1920, 1200)
det = Detector(CudaSiftConfig(col_width=cols, row_width=rows))
det.gpuWarmUpAndAllocate()
img = synth_frame(rows, cols, 0)
for _ in range(5):
    det.detectAndCompute(img)
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        det.detectAndCompute(img)
    torch.cuda.synchronize()
path = os.path.join(tempfile.gettempdir(), "trace.json")
prof.export_chrome_trace(path)
ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
ev.sort(key=lambda e: e["ts"])
if not ev:
    print("no GPU activity records captured")
    sys.exit(0)
# split into frames at gaps > 200 us
frames, cur = [], [ev[0]]
for e in ev[1:]:
    if e["ts"] - (cur[-1]["ts"] + cur[-1]["dur"]) > 200:
        frames.append(cur); cur = []
    cur.append(e)
frames.append(cur)
f = frames[-1]
t0 = f[0]["ts"]
print(f"{len(frames)} frames; last frame: {len(f)} records, span {f[-1]['ts'] + f[-1]['dur'] - t0:.1f} us")
for e in f:
    name = e["name"].split("(")[0].split("::")[-1][:34]
    a = e.get("args", {})
    print(f"{e['ts'] - t0:8.1f} +{e['dur']:7.1f}  strm {a.get('stream', '?'):>3}  grid {str(a.get('grid', '')):>16}  {name}")
