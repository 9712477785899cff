set +e
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2h_pytest.log 2>&1; tail -3 gpurun_out/r2h_pytest.log
for f in 1 0; do
  B200SIFT_FUSE=$f python bench.py --quick --steps 5 2>/dev/null | tail -1 > gpurun_out/r2h_ab_fuse$f.json
done
python bench.py --quick --steps 5 --size 752x480 2>/dev/null | tail -1 > gpurun_out/r2h_ab752.json
python bench.py --quick --steps 5 --size 1600x900 2>/dev/null | tail -1 > gpurun_out/r2h_ab1600.json
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2h_launches.csv python tools/prof_frame.py > /tmp/ncu_l.log 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:descriptor -f -o /tmp/desc python tools/prof_frame.py > /tmp/ncu_desc.log 2>&1
ncu -i /tmp/desc.ncu-rep --page source --csv > gpurun_out/r2h_desc_source.csv 2>/dev/null
ncu -i /tmp/desc.ncu-rep --page raw --csv > gpurun_out/r2h_desc_raw.csv 2>/dev/null
tail -2 /tmp/ncu_desc.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2h_ab*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'lat %.4f'%j['frame_graph_ms'], 'pipe %.4f'%j['frame_pipelined_ms'], 'pyr %.4f'%j['graph_cumulative_ms']['pyramid'])
    except Exception as e:
        print(f, 'ERR', e)
PY
du -sh gpurun_out
