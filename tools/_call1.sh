timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/r02r_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r02r_tests.log; tail -5 gpurun_out/r02r_tests.log
timeout 600 python bench.py > gpurun_out/r02r_bench_n1.log 2>&1; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02r_bench_ref.log 2>&1; echo "ref rc=$?"; tail -c 700 gpurun_out/r02r_bench_ref.log
timeout 300 python tools/perf_float.py > gpurun_out/r02r_perf_float.log 2>&1; echo "float rc=$?"
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02r_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r02r_smoke.log
