set +e
mkdir -p gpurun_out
for c in 8 7 6 5; do
  B200SIFT_DESC_CTAS=$c python bench.py --quick --steps 5 2>/dev/null | tail -1 > gpurun_out/r2p_ab_ctas${c}.json
done
for sl in 16 24 32; do
  B200SIFT_BENCH_SLOTS=$sl python bench.py --quick --steps 5 2>/dev/null | tail -1 > gpurun_out/r2p_ab_slots${sl}.json
done
B200SIFT_FUSE=0 python bench.py --quick --steps 5 2>/dev/null | tail -1 > gpurun_out/r2p_ab_fuse0.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2p_ab*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'lat %.4f'%j['frame_graph_ms'], 'pipe %.4f'%j['frame_pipelined_ms'], 'pyr %.4f'%j['graph_cumulative_ms']['pyramid'])
    except Exception as e:
        print(f, 'ERR', e)
PY
