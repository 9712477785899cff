run() { echo "== $*"; env "$@" python tools/dev_timeline.py 1920x1200 128 2>/dev/null | tail -1; env "$@" python bench.py --quick --steps 5 2>/dev/null | tail -1 | python -c "import json,sys; j=json.loads(sys.stdin.read()); print('   value %.0f lat %.4f pipe %.4f'%(j['value'], j['frame_graph_ms'], j['frame_pipelined_ms']))"; }
run X=0
run B200SIFT_PRIO_A=1
run B200SIFT_PRIO_A=3
run B200SIFT_DESC_WAVES=4
run B200SIFT_DESC_WAVES=4 B200SIFT_PRIO_A=2
run B200SIFT_DESC_WAVES=2 B200SIFT_PRIO_A=2
run B200SIFT_DESC_WAVES=4 B200SIFT_PRIO_A=2 CUDA_DEVICE_MAX_CONNECTIONS=32
