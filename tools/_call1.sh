timeout 900 python -m pytest tests/test_gpu_assign.py tests/test_gpu_float.py -m gpu -x -q > gpurun_out/r03a_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/r03a_tests.log
timeout 300 python tools/perf_float.py > gpurun_out/r03a_perf_float.log 2>&1; echo "float rc=$?"; head -5 gpurun_out/r03a_perf_float.log | cut -c1-260
