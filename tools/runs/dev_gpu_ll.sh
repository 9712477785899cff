set +e
mkdir -p gpurun_out
for rep in 1 2; do for c in 8 7 6 5; do
  B200SIFT_DESC_CTAS=$c python bench.py --no-extras --no-cpu-baseline --steps 5 2>/dev/null | tail -1 > gpurun_out/r2ll_ctas${c}_$rep.json
done; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2ll_*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'e2e %.0f'%j['e2e']['value'], 'u8 %.0f'%j['e2e_u8']['value'], 'lat %.4f'%j['frame_graph_ms'])
    except Exception as e:
        print(f, 'ERR', e)
PY
