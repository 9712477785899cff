set +e
mkdir -p gpurun_out
for rep in 1 2; do for v in 1 2 4; do
  B200SIFT_COPY_STREAMS=$v python bench.py --no-extras --no-cpu-baseline --steps 5 2>/dev/null | tail -1 > gpurun_out/r2hh_copy${v}_$rep.json
done; done
B200SIFT_COPY_STREAMS=2 python bench.py --no-extras --no-cpu-baseline --steps 5 --size 752x480 2>/dev/null | tail -1 > gpurun_out/r2hh_752_copy2.json
B200SIFT_COPY_STREAMS=1 python bench.py --no-extras --no-cpu-baseline --steps 5 --size 752x480 2>/dev/null | tail -1 > gpurun_out/r2hh_752_copy1.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2hh_*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'e2e %.0f'%j['e2e']['value'], 'u8 %.0f'%j['e2e_u8']['value'])
    except Exception as e:
        print(f, 'ERR', e)
PY
