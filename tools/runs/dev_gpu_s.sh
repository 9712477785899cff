set +e
mkdir -p gpurun_out
for sl in 16 32; do for fr in 64 128; do
  B200SIFT_BENCH_SLOTS=$sl python bench.py --no-extras --no-cpu-baseline --steps 5 --frames $fr 2>/dev/null | tail -1 > gpurun_out/r2s_slots${sl}_f${fr}.json
done; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2s_*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'e2e %.0f'%j['e2e']['value'], 'u8 %.0f'%j['e2e_u8']['value'], 'lat %.4f'%j['frame_graph_ms'], 'pipe %.4f'%j['frame_pipelined_ms'])
    except Exception as e:
        print(f, 'ERR', e)
PY
python - <<'PY'
import time, torch
from another_cuda_sift_b200 import CudaSiftConfig, Detector
for sl in (6, 16, 32):
    t=time.time(); d=Detector(CudaSiftConfig(col_width=1920,row_width=1200), frames_in_flight=sl); d.gpuWarmUpAndAllocate(); torch.cuda.synchronize(); print('create', sl, round(time.time()-t,3),'s', torch.cuda.memory_allocated(), torch.cuda.mem_get_info())
    del d
PY
