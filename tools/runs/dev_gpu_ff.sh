set +e
mkdir -p gpurun_out
for rep in 1 2; do for v in 1 2 4; do
  B200SIFT_VH_NB=$v python bench.py --quick --steps 5 2>/dev/null | tail -1 > gpurun_out/r2ff_ab_nb${v}_$rep.json
done; done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2ff_ab*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'lat %.4f'%j['frame_graph_ms'], 'pipe %.4f'%j['frame_pipelined_ms'], 'pyr %.4f'%j['graph_cumulative_ms']['pyramid'], j['kernel_ms_per_frame'].get('blur_oct0'))
    except Exception as e:
        print(f, 'ERR', e)
PY
