timeout 120 python tools/prof_real_sift.py > gpurun_out/r03m_case.log 2>&1; echo "case rc=$?"; tail -1 gpurun_out/r03m_case.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:^sweep_kernel -s 2 -c 2 -f -o gpurun_out/r03m_verify python tools/prof_real_sift.py > gpurun_out/r03m_ncu.log 2>&1; echo "ncu rc=$?"
