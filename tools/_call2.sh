timeout 600 python -m pytest tests/test_gpu_multirank.py tests/test_gpu_host_api.py -m gpu -x -q > gpurun_out/r03l_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/r03l_tests.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-secondary > gpurun_out/r03l_bench_n2.log 2>&1; echo "bench rc=$?"
