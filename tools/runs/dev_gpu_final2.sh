set +e
mkdir -p gpurun_out
python bench.py > gpurun_out/r2_bench_n1_final.json 2> gpurun_out/r2_bench_n1_final.err
python bench.py --impl reference > gpurun_out/r2_bench_ref_final.json 2> gpurun_out/r2_bench_ref_final.err
python bench.py --impl reference > gpurun_out/r2_bench_ref_final_b.json 2>/dev/null
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_n1_final.json','gpurun_out/r2_bench_ref_final.json','gpurun_out/r2_bench_ref_final_b.json'):
    try:
        j=json.loads([l for l in open(f) if l.startswith('{')][-1])
        print(f, 'value', round(j['value'],1), 'e2e', round(j['e2e']['value'],1), 'u8', (j.get('e2e_u8') or {}).get('value'), 'kp', j['config'].get('keypoints_per_frame'), j.get('breakdown_ms_per_frame'))
        for c,v in (j.get('configs') or {}).items(): print('   ',c, round(v['value']), round(v['e2e']), round(v['e2e_u8']), v['keypoints_per_frame'], v.get('frame_graph_ms'))
    except Exception as e: print(f,'ERR',e)
PY
