set +e
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r2_pytest_gpu_final.log 2>&1; tail -3 gpurun_out/r2_pytest_gpu_final.log
python bench.py --impl reference > gpurun_out/r2_bench_ref_final.json 2> gpurun_out/r2_bench_ref_final.err
python bench.py > gpurun_out/r2_bench_n1_final.json 2> gpurun_out/r2_bench_n1_final.err
ncu --profile-from-start off --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --csv --log-file gpurun_out/r2_launches_final.csv python tools/prof_frame.py > /tmp/ncu_l.log 2>&1
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --quick --steps 2 --warmup 1 --frames 4 > /tmp/ncu_lb.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -f -o /tmp/full_final python tools/prof_frame.py > /tmp/ncu_full.log 2>&1
ncu -i /tmp/full_final.ncu-rep --page raw --csv > gpurun_out/r2_full_final_raw.csv 2>/dev/null
ncu -i /tmp/full_final.ncu-rep --page source --csv -k regex:descriptor > gpurun_out/r2_desc_source_final.csv 2>/dev/null
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:match -f -o /tmp/full_match python tools/prof_frame.py --match 100000 > /tmp/ncu_match.log 2>&1
ncu -i /tmp/full_match.ncu-rep --page raw --csv > gpurun_out/r2_full_match100k_raw.csv 2>/dev/null
tail -2 /tmp/ncu_full.log /tmp/ncu_match.log
python - <<'PY'
import json
for f in ('gpurun_out/r2_bench_n1_final.json','gpurun_out/r2_bench_ref_final.json'):
    try:
        j=json.loads([l for l in open(f) if l.startswith('{')][-1])
        print(f, 'value', j['value'], 'e2e', j['e2e']['value'], 'ms/step', j['ms_per_step'])
    except Exception as e: print(f,'ERR',e)
PY
du -sh gpurun_out
