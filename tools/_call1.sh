timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r03f_tests.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/r03f_tests.log; tail -3 gpurun_out/r03f_tests.log
timeout 600 python bench.py > gpurun_out/r03f_bench_n1.log 2>&1; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r03f_bench_ref.log 2>&1; echo "ref rc=$?"
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r03f_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r03f_smoke.log
