set +e
mkdir -p gpurun_out
for rep in 1 2; do for c in 8 32; do
  CUDA_DEVICE_MAX_CONNECTIONS=$c python bench.py --no-extras --no-cpu-baseline --steps 5 2>/dev/null | tail -1 > gpurun_out/r2ii_conn${c}_$rep.json
done; done
for c in 8 32; do
  CUDA_DEVICE_MAX_CONNECTIONS=$c python bench.py --no-extras --no-cpu-baseline --steps 5 --size 752x480 2>/dev/null | tail -1 > gpurun_out/r2ii_752_conn${c}.json
  CUDA_DEVICE_MAX_CONNECTIONS=$c python bench.py --impl reference --steps 3 2>/dev/null | tail -1 > gpurun_out/r2ii_ref_conn${c}.json
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2ii_*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'e2e %.0f'%j['e2e']['value'], 'u8', (j.get('e2e_u8') or {}).get('value'), j.get('frame_graph_ms'))
    except Exception as e:
        print(f, 'ERR', e)
PY
