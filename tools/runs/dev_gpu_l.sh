set +e
mkdir -p gpurun_out
for sl in 6 16 24 32; do
  B200SIFT_BENCH_SLOTS=$sl python bench.py --quick --steps 5 --frames 32 2>/dev/null | tail -1 > gpurun_out/r2l_ab_f32_slots${sl}.json
done
for sl in 16 32; do
  B200SIFT_BENCH_SLOTS=$sl python bench.py --quick --steps 5 --frames 32 --size 752x480 2>/dev/null | tail -1 > gpurun_out/r2l_ab752_f32_slots${sl}.json
done
B200SIFT_BENCH_SLOTS=16 python bench.py --no-extras --no-cpu-baseline --steps 5 2>/dev/null | tail -1 > gpurun_out/r2l_full_slots16.json
B200SIFT_BENCH_SLOTS=6 python bench.py --no-extras --no-cpu-baseline --steps 5 2>/dev/null | tail -1 > gpurun_out/r2l_full_slots6.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2l_*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'lat %.4f'%j['frame_graph_ms'], 'pipe %.4f'%j['frame_pipelined_ms'], 'e2e', j.get('e2e',{}).get('value'), 'u8', (j.get('e2e_u8') or {}).get('value'))
    except Exception as e:
        print(f, 'ERR', e)
PY
