timeout 900 python -m pytest tests/test_gpu_real_sift.py tests/test_gpu_parity.py tests/test_gpu_configs.py -m gpu -x -q > gpurun_out/r02m_tests.log 2>&1; echo "rc=$?" >> gpurun_out/r02m_tests.log; tail -15 gpurun_out/r02m_tests.log
timeout 200 python tools/perf_knn2.py > gpurun_out/r02m_perf_knn2.log 2>&1; cat gpurun_out/r02m_perf_knn2.log
