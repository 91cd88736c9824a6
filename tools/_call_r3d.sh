#!/bin/bash
# round-3 evidence call: bench with the new secondary rows, then three ncu --set full captures (one launch each)
set -u
O=gpurun_out
python bench.py > $O/r3d_bench.json 2> $O/r3d_bench.err; echo "bench rc=$?" >> $O/r3d_bench.err
NCU="ncu --set full --clock-control none --import-source on -f"
timeout 600 $NCU -k regex:chunk_sweep256 -s 2 -c 1 -o /tmp/r3d_chunk python tools/prof_case.py 33 16384 1 2 > $O/r3d_ncu_chunk.log 2>&1
timeout 600 $NCU -k regex:sweep_kernel -s 1 -c 1 -o /tmp/r3d_verify python tools/prof_real_sift.py 2 > $O/r3d_ncu_verify.log 2>&1
timeout 600 $NCU -k regex:ip_sweep2_kernel -s 2 -c 1 -o /tmp/r3d_lse python tools/prof_float_case.py 256 1024 head > $O/r3d_ncu_lse.log 2>&1
for n in chunk verify lse; do
  if [ -f /tmp/r3d_$n.ncu-rep ]; then
    python tools/ncu_summary.py /tmp/r3d_$n.ncu-rep > $O/r3d_${n}_ncu_summary.txt 2>&1
    ls -la /tmp/r3d_$n.ncu-rep >> $O/r3d_${n}_ncu_summary.txt
    sz=$(stat -c %s /tmp/r3d_$n.ncu-rep); if [ $sz -lt 15000000 ]; then cp /tmp/r3d_$n.ncu-rep $O/; fi
  fi
done
