for c in 8 6 4; do
  echo "== B200SIFT_DESC_CTAS=$c"
  B200SIFT_DESC_CTAS=$c python tools/dev_timeline.py 1920x1200 128 2>/dev/null | tail -1
done
echo "== 32 connections"; CUDA_DEVICE_MAX_CONNECTIONS=32 python tools/dev_timeline.py 1920x1200 128 2>/dev/null | tail -1
echo "== 32 slots"; python tools/dev_timeline.py 1920x1200 128 32 2>/dev/null | tail -1
echo "== 6 slots"; python tools/dev_timeline.py 1920x1200 128 6 2>/dev/null | tail -1
echo "== 752"; python tools/dev_timeline.py 752x480 128 2>/dev/null | tail -1
