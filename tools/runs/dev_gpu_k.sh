set +e
mkdir -p gpurun_out
python -m pytest tests/test_gpu_options.py -m gpu -x -q > gpurun_out/r2k_pytest.log 2>&1; tail -3 gpurun_out/r2k_pytest.log
for sl in 6 8 10 12 16; do
  B200SIFT_BENCH_SLOTS=$sl python bench.py --quick --steps 5 2>/dev/null | tail -1 > gpurun_out/r2k_ab_slots${sl}.json
  B200SIFT_BENCH_SLOTS=$sl B200SIFT_DESC_CTAS=7 python bench.py --quick --steps 5 2>/dev/null | tail -1 > gpurun_out/r2k_ab_slots${sl}_c7.json
done
for sl in 6 10 16; do
  B200SIFT_BENCH_SLOTS=$sl python bench.py --quick --steps 5 --size 752x480 2>/dev/null | tail -1 > gpurun_out/r2k_ab752_slots${sl}.json
  B200SIFT_BENCH_SLOTS=$sl python bench.py --quick --steps 5 --size 1600x900 2>/dev/null | tail -1 > gpurun_out/r2k_ab1600_slots${sl}.json
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2k_ab*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'lat %.4f'%j['frame_graph_ms'], 'pipe %.4f'%j['frame_pipelined_ms'], 'pyr %.4f'%j['graph_cumulative_ms']['pyramid'])
    except Exception as e:
        print(f, 'ERR', e)
PY
