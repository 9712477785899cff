set +e
mkdir -p gpurun_out
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 200 $TR --nproc-per-node 8 --master-port 29611 tools/h2d_ceiling.py --affinity 2>/dev/null | grep '^{' > gpurun_out/r2_h2d_ceiling_8gpu.json
timeout 100 python tools/h2d_ceiling.py 2>/dev/null | grep '^{' > gpurun_out/r2_h2d_ceiling_1gpu.json
timeout 400 $TR --nproc-per-node 8 --master-port 29612 bench.py --gpus 8 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r2_pytest_multi_n8.log 2>&1; tail -3 gpurun_out/r2_pytest_multi_n8.log
timeout 200 python bench.py --gpus 1 --no-extras --no-cpu-baseline > gpurun_out/r2_bench_n1_samebox.json 2>/dev/null
timeout 300 $TR --nproc-per-node 4 --master-port 29613 bench.py --gpus 4 --no-cpu-baseline > gpurun_out/r2_bench_n4.json 2> gpurun_out/r2_bench_n4.err
cat gpurun_out/r2_h2d_ceiling_8gpu.json gpurun_out/r2_h2d_ceiling_1gpu.json
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_n8.json','gpurun_out/r2_bench_n4.json','gpurun_out/r2_bench_n1_samebox.json'):
    try:
        j=json.loads([l for l in open(f) if l.startswith('{')][-1])
        print(f, 'value', round(j['value']), 'e2e', round(j['e2e']['value']), 'u8', round((j.get('e2e_u8') or {}).get('value') or 0))
        for k,c in (j.get('configs') or {}).items(): print('  cfg', k, {x:c.get(x) for x in ('value','e2e','e2e_u8','frames_per_step_per_gpu')})
        m=j.get('match_100k'); 
        if m: print('  match_100k', m['ms'], m['stages_ms_max_over_ranks'], m['oracle_check']['mismatches'], m['roofline'])
    except Exception as e: print(f,'ERR',e)
PY
tail -c 400 gpurun_out/r2_bench_n8.err
