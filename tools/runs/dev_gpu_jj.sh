set +e
mkdir -p gpurun_out
python bench.py > gpurun_out/r2_bench_n1_final.json 2> gpurun_out/r2_bench_n1_final.err
python bench.py --impl reference > gpurun_out/r2_bench_ref_final.json 2>/dev/null
python bench.py --impl reference > gpurun_out/r2_bench_ref_final_b.json 2>/dev/null
python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -2
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_n1_final.json','gpurun_out/r2_bench_ref_final.json','gpurun_out/r2_bench_ref_final_b.json'):
    try:
        j=json.loads([l for l in open(f) if l.startswith('{')][-1])
        print(f, 'value', round(j['value'],1), 'e2e', round(j['e2e']['value'],1), 'u8', (j.get('e2e_u8') or {}).get('value'), j['config'].get('cuda_device_max_connections'), j.get('frame_graph_ms'), j.get('frame_pipelined_ms'))
        for c,v in (j.get('configs') or {}).items(): print('   ',c, round(v['value']), round(v['e2e']), round(v['e2e_u8']), v['keypoints_per_frame'], v.get('frame_graph_ms'))
        if j.get('match_100k'): print('   m100k', j['match_100k']['ms'], j['match_2k2k']['ms_median'], j['roofline']['frac'], j['graph_cumulative_ms'])
    except Exception as e: print(f,'ERR',e)
PY
