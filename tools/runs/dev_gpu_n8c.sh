set +e
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 400 $TR --nproc-per-node 8 --master-port 29631 bench.py --gpus 8 --no-cpu-baseline > gpurun_out/r2_bench_n8_final.json 2> gpurun_out/r2_bench_n8_final.err
timeout 200 $TR --nproc-per-node 4 --master-port 29632 bench.py --gpus 4 --no-cpu-baseline --no-extras > gpurun_out/r2_bench_n4_final.json 2> gpurun_out/r2_bench_n4_final.err
timeout 200 $TR --nproc-per-node 2 --master-port 29633 bench.py --gpus 2 --no-cpu-baseline --no-extras > gpurun_out/r2_bench_n2_final.json 2> gpurun_out/r2_bench_n2_final.err
timeout 200 python bench.py --gpus 1 --no-extras --no-cpu-baseline > gpurun_out/r2_bench_n1_samebox.json 2>/dev/null
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_n8_final.json','gpurun_out/r2_bench_n4_final.json','gpurun_out/r2_bench_n2_final.json','gpurun_out/r2_bench_n1_samebox.json'):
    try:
        j=json.loads([l for l in open(f) if l.startswith('{')][-1])
        print(f, 'value', round(j['value']), 'e2e', round(j['e2e']['value']), 'u8', round((j.get('e2e_u8') or {}).get('value') or 0), j['config'].get('keypoints_per_frame'))
        for c,v in (j.get('configs') or {}).items(): print('   ',c, round(v['value']), round(v['e2e']), round(v['e2e_u8']), v['frames_per_step_per_gpu'])
        m=j.get('match_100k'); 
        if m: print('  match_100k', m['ms'], m['stages_ms_max_over_ranks'], m['oracle_check']['mismatches'])
    except Exception as e: print(f,'ERR',e)
PY
