set +e
mkdir -p gpurun_out
python -m pytest tests/test_gpu_match.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r2o_pytest.log 2>&1; tail -3 gpurun_out/r2o_pytest.log
python tools/bench_match.py --large 100000 2>/dev/null | tail -3 > gpurun_out/r2o_match.json; cat gpurun_out/r2o_match.json | cut -c1-1500
