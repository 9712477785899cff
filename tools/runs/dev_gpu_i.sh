set +e
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2i_pytest.log 2>&1; tail -3 gpurun_out/r2i_pytest.log
L=another_cuda_sift_b200/libb200sift.so
cp $L /tmp/new.so
for rep in 1 2; do
  for v in old new; do
    if [ $v = old ]; then cp build_ab/libb200sift_olddesc.so $L; else cp /tmp/new.so $L; fi
    python bench.py --quick --steps 5 2>/dev/null | tail -1 > gpurun_out/r2i_ab_${v}_$rep.json
  done
done
cp /tmp/new.so $L
ncu --profile-from-start off --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --csv --log-file gpurun_out/r2i_launches.csv python tools/prof_frame.py > /tmp/ncu_l.log 2>&1
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2i_ab*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'lat %.4f'%j['frame_graph_ms'], 'pipe %.4f'%j['frame_pipelined_ms'], 'pyr %.4f'%j['graph_cumulative_ms']['pyramid'])
    except Exception as e:
        print(f, 'ERR', e)
PY
du -sh gpurun_out
