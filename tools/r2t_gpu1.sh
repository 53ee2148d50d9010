#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_dynamics.py -m gpu -q -k "transport or convergence or mass or cone" > gpurun_out/r2t_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2t_pytest.log
python bench.py --workload cone --no-cpu > gpurun_out/r2t_bench_cone.json 2> gpurun_out/r2t_bench_cone.err; echo "cone rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2t_bench_cone.json"))
print(d["value"], d["ms_per_step"], d["roofline"]["avg_launch_ms"], d["roofline"]["kernel"], d["clocks"]["sm_mhz"], d["mass_drift_rel"])
PY
