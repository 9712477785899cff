set +e
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2j_pytest.log 2>&1; tail -3 gpurun_out/r2j_pytest.log
for rep in 1 2; do
  for c in 8 7 6; do
    B200SIFT_DESC_CTAS=$c python bench.py --quick --steps 5 2>/dev/null | tail -1 > gpurun_out/r2j_ab_ctas${c}_$rep.json
  done
done
for sl in 4 8; do
  B200SIFT_BENCH_SLOTS=$sl python bench.py --quick --steps 5 2>/dev/null | tail -1 > gpurun_out/r2j_ab_slots${sl}.json
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2j_ab*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'lat %.4f'%j['frame_graph_ms'], 'pipe %.4f'%j['frame_pipelined_ms'], 'pyr %.4f'%j['graph_cumulative_ms']['pyramid'])
    except Exception as e:
        print(f, 'ERR', e)
PY
