K='regex:^(sweep_kernel|chunk_sweep.*|refine_kernel|select.*|emit_kernel|sort_.*|bank_prepare_kernel|norm_chunk_min_kernel|compact_kernel|mark_invalid_pairs_kernel|knn.*|ip_.*)$'
timeout 300 python tools/perf_knn2.py > gpurun_out/r02q_perf_knn2.log 2>&1; echo "knn2 rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 3000 --csv --log-file gpurun_out/r02q_launches_knn2.csv python tools/perf_knn2.py > gpurun_out/r02q_ncu_knn2.log 2>&1; echo "ncu knn2 list rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:knn_eval_tc_kernel -s 2 -c 1 -f -o gpurun_out/r02q_knn_tc python tools/perf_knn2.py > gpurun_out/r02q_ncu_knn_tc.log 2>&1; echo "ncu knn tc rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:chunk_sweep256_kernel -s 20 -c 1 -f -o gpurun_out/r02q_chunk python bench.py --steps 1 --warmup 1 --no-secondary --no-cpu-baseline > gpurun_out/r02q_ncu_chunk.log 2>&1; echo "ncu chunk rc=$?"
ls -la gpurun_out/r02q*
