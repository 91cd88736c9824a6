timeout 900 python -m pytest tests/test_gpu_assign.py tests/test_gpu_float.py -m gpu -x -q > gpurun_out/r02v_tests.log 2>&1; echo "tests rc=$?"; tail -12 gpurun_out/r02v_tests.log
timeout 300 python tools/perf_float.py > gpurun_out/r02v_perf_float.log 2>&1; echo "float rc=$?"; cut -c1-420 gpurun_out/r02v_perf_float.log
