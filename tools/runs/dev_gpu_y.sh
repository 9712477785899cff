for c in 8 16 32; do
  echo "== CUDA_DEVICE_MAX_CONNECTIONS=$c"
  CUDA_DEVICE_MAX_CONNECTIONS=$c python tools/dev_timeline.py 1920x1200 48 2>/dev/null | tail -6
  CUDA_DEVICE_MAX_CONNECTIONS=$c python bench.py --quick --steps 5 2>/dev/null | tail -1 | python -c "import json,sys; j=json.loads(sys.stdin.read()); print('value %.0f lat %.4f pipe %.4f'%(j['value'], j['frame_graph_ms'], j['frame_pipelined_ms']))"
done
