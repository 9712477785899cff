set +e
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2m_pytest.log 2>&1; tail -3 gpurun_out/r2m_pytest.log
python bench.py > gpurun_out/r2m_bench.json 2> gpurun_out/r2m_bench.err; tail -c 600 gpurun_out/r2m_bench.err
python bench.py --impl reference > gpurun_out/r2m_bench_ref.json 2> gpurun_out/r2m_bench_ref.err
python - <<'PY'
import json
for f in ('gpurun_out/r2m_bench.json','gpurun_out/r2m_bench_ref.json'):
    try:
        j=json.loads([l for l in open(f) if l.startswith('{')][-1])
        print(f, 'value', j['value'], 'e2e', j['e2e']['value'], 'ms/step', j['ms_per_step'], {k:j.get(k) for k in ('frame_graph_ms','frame_pipelined_ms')})
        print('  roofline', j.get('roofline')); print('  e2e_u8', (j.get('e2e_u8') or {}).get('value'))
        for c in j.get('configs',[]) or []: print('  cfg', {k:c.get(k) for k in ('workload','value','e2e','latency_ms','e2e_u8')})
        print('  match_100k', j.get('match_100k')); print('  match_2k2k', j.get('match_2k2k'))
    except Exception as e: print(f,'ERR',e)
PY
