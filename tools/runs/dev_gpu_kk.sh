set +e
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" python bench.py --no-extras --no-cpu-baseline --steps 5 2>/dev/null | tail -1 > gpurun_out/r2kk_$tag.json; }
run base X=0
run slots32 B200SIFT_BENCH_SLOTS=32
run slots24 B200SIFT_BENCH_SLOTS=24
run waves4 B200SIFT_DESC_WAVES=4
run prioa2 B200SIFT_PRIO_A=2
run shapes1 B200SIFT_CHAIN_SHAPES=1
run ctas6 B200SIFT_DESC_CTAS=6
run base2 X=0
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2kk_*.json')):
    try:
        j=json.loads(open(f).read())
        print(f.split('/')[-1], 'value %.0f'%j['value'], 'e2e %.0f'%j['e2e']['value'], 'u8 %.0f'%j['e2e_u8']['value'], 'lat %.4f'%j['frame_graph_ms'])
    except Exception as e:
        print(f, 'ERR', e)
PY
