for v in 0 1 2 0 1 2; do
  cp gpurun_variants/libb200m_v$v.so 3d_reconstruction_b200/lib/libb200m.so
  timeout 300 python bench.py --no-secondary --no-cpu-baseline > gpurun_out/r03n_v$v.log 2>&1
  python - $v <<'PY'
import json,sys
v=sys.argv[1]
try:
    d=json.loads([x for x in open(f"gpurun_out/r03n_v{v}.log") if x.startswith("{")][-1])
    print("variant",v,"value",round(d["value"]),"ms",round(d["ms_per_step"],1),"launch_ms",round(d["roofline"]["launch_ms"],3),"mhz",d["clocks"]["sm_mhz"],"e2e",round(d["e2e"]["value"]),d["config"]["result_digest"])
except Exception as e:
    print("variant",v,"failed",e)
PY
done
