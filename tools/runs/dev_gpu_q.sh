set +e
mkdir -p gpurun_out
python -m pytest tests/test_gpu_match.py -m gpu -x -q > gpurun_out/r2q_pytest.log 2>&1; tail -3 gpurun_out/r2q_pytest.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --quick --steps 2 --warmup 3 --frames 4 > /tmp/ncu_lb.log 2>&1
tail -3 /tmp/ncu_lb.log; wc -l gpurun_out/r2_launches_bench.csv
