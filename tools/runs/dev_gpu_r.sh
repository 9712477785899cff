set +e
mkdir -p gpurun_out
for rep in 1 2; do for v in 0 1; do
  B200SIFT_CHAIN_SHAPES=$v python bench.py --quick --steps 5 2>/dev/null | tail -1 > gpurun_out/r2r_ab_shapes${v}_$rep.json
done; done
B200SIFT_CHAIN_SHAPES=1 python bench.py --quick --steps 5 --size 752x480 2>/dev/null | tail -1 > gpurun_out/r2r_ab752_shapes1.json
B200SIFT_CHAIN_SHAPES=0 python bench.py --quick --steps 5 --size 752x480 2>/dev/null | tail -1 > gpurun_out/r2r_ab752_shapes0.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2r_ab*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'lat %.4f'%j['frame_graph_ms'], 'pipe %.4f'%j['frame_pipelined_ms'], 'pyr %.4f'%j['graph_cumulative_ms']['pyramid'])
    except Exception as e:
        print(f, 'ERR', e)
PY
