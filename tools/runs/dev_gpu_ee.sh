set +e
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_options.py -m gpu -x -q > gpurun_out/r2ee_pytest.log 2>&1; tail -3 gpurun_out/r2ee_pytest.log
for rep in 1 2; do for v in 0 1; do
  B200SIFT_FFMA2=$v python bench.py --quick --steps 5 2>/dev/null | tail -1 > gpurun_out/r2ee_ab_ffma2_${v}_$rep.json
done; done
for v in 0 1; do
  B200SIFT_FFMA2=$v python bench.py --quick --steps 5 --size 752x480 2>/dev/null | tail -1 > gpurun_out/r2ee_ab752_ffma2_${v}.json
done
ncu --profile-from-start off --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:blur --csv --log-file gpurun_out/r2ee_launches.csv python tools/prof_frame.py > /tmp/ncu_l.log 2>&1
python - <<'PY'
import json,glob,csv,re,collections
for f in sorted(glob.glob('gpurun_out/r2ee_ab*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'lat %.4f'%j['frame_graph_ms'], 'pipe %.4f'%j['frame_pipelined_ms'], 'pyr %.4f'%j['graph_cumulative_ms']['pyramid'])
    except Exception as e:
        print(f, 'ERR', e)
rows=list(csv.reader(open('gpurun_out/r2ee_launches.csv')))
h=[i for i,r in enumerate(rows) if r and r[0]=='ID'][0]
hdr=rows[h]; data=rows[h+1:]
iK=hdr.index('Kernel Name'); iM=hdr.index('Metric Name'); iV=hdr.index('Metric Value')
t=collections.OrderedDict()
for r in data:
    if len(r)<=iV: continue
    name=re.sub(r"\(.*","",r[iK]).split("::")[-1]
    t.setdefault((r[0],name),{})[r[iM]]=float(r[iV].replace(',',''))
for (i,k),m in t.items(): print(i, k[:50], round(m.get('gpu__time_duration.sum',0)/1000,2),'us', round(m.get('smsp__inst_executed.sum',0)/1e6,3),'M')
PY
