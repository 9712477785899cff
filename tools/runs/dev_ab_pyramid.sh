set +e
mkdir -p gpurun_out
for m in auto tile stream vh; do for f in 1 0; do
  B200SIFT_BLUR=$m B200SIFT_FUSE=$f python bench.py --quick --steps 5 2>/dev/null | tail -1 > gpurun_out/r2g_ab_${m}_fuse$f.json
done; done
for m in auto vh; do for f in 1 0; do
  B200SIFT_BLUR=$m B200SIFT_FUSE=$f python bench.py --quick --steps 5 --size 752x480 2>/dev/null | tail -1 > gpurun_out/r2g_ab752_${m}_fuse$f.json
done; done
for cfg in auto:1 tile:1 vh:0 stream:0; do m=${cfg%%:*}; f=${cfg##*:}
  B200SIFT_BLUR=$m B200SIFT_FUSE=$f ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2g_launches_${m}_fuse$f.csv python tools/prof_frame.py > /tmp/ncu_$m.log 2>&1
done
ncu --profile-from-start off --set full --clock-control none --import-source on -f -o /tmp/full_auto python tools/prof_frame.py > /tmp/ncu_full.log 2>&1
ncu -i /tmp/full_auto.ncu-rep --page raw --csv > gpurun_out/r2g_full_auto_raw.csv 2>/dev/null
B200SIFT_BLUR=vh B200SIFT_FUSE=0 ncu --profile-from-start off --set full --clock-control none -k regex:blur -f -o /tmp/full_vh python tools/prof_frame.py > /tmp/ncu_full_vh.log 2>&1
ncu -i /tmp/full_vh.ncu-rep --page raw --csv > gpurun_out/r2g_full_vh_raw.csv 2>/dev/null
tail -2 /tmp/ncu_full.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2g_ab*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'lat %.4f'%j['frame_graph_ms'], 'pipe %.4f'%j['frame_pipelined_ms'], 'pyr %.4f'%j['graph_cumulative_ms']['pyramid'])
    except Exception as e:
        print(f, 'ERR', e)
PY
du -sh gpurun_out
