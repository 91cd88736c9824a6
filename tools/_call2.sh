timeout 300 python tools/perf_data.py > gpurun_out/r02s_perf_data.log 2>&1; echo "perf_data rc=$?"; cat gpurun_out/r02s_perf_data.log | cut -c1-700
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 4 > gpurun_out/r02s_bench_n4.log 2>&1; echo "bench rc=$?"; tail -c 400 gpurun_out/r02s_bench_n4.log
