set +e
mkdir -p gpurun_out
python bench.py --no-extras --no-cpu-baseline --steps 5 --frames 16 2>/dev/null | tail -1 > gpurun_out/r2u_f16.json
B200SIFT_BENCH_SLOTS=6 python bench.py --no-extras --no-cpu-baseline --steps 5 --frames 16 2>/dev/null | tail -1 > gpurun_out/r2u_f16_slots6.json
python bench.py --no-extras --no-cpu-baseline --steps 5 --frames 64 2>/dev/null | tail -1 > gpurun_out/r2u_f64.json
B200SIFT_BENCH_SLOTS=6 python bench.py --no-extras --no-cpu-baseline --steps 5 --frames 64 2>/dev/null | tail -1 > gpurun_out/r2u_f64_slots6.json
B200SIFT_BENCH_SLOTS=32 python bench.py --no-extras --no-cpu-baseline --steps 5 --frames 64 2>/dev/null | tail -1 > gpurun_out/r2u_f64_slots32.json
python bench.py --no-extras --no-cpu-baseline --steps 5 --frames 128 2>/dev/null | tail -1 > gpurun_out/r2u_f128.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2u_*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'e2e %.0f'%j['e2e']['value'], 'u8 %.0f'%j['e2e_u8']['value'], 'kp', j['config']['keypoints_per_frame'], 'lat %.4f'%j['frame_graph_ms'], 'pipe %.4f'%j['frame_pipelined_ms'])
    except Exception as e:
        print(f, 'ERR', e)
PY
