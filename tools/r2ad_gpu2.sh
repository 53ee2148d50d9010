#!/bin/bash
# two transport stages per launch on strips (N = 2): parity + bit identity, bench with and without
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -s > gpurun_out/r2ad_pytest_multi_n2.log 2>&1; echo "pytest multi rc=$?"; grep -E "FAIL|MGPU|worst|passed|failed|dG1" gpurun_out/r2ad_pytest_multi_n2.log | tail -14
for fz in 1 0; do
  NSDG_TRANSPORT_FUSED=$fz timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29547 bench.py --gpus 2 --main-only --no-e2e --no-parity > gpurun_out/r2ad_bench_n2_fused$fz.json 2> gpurun_out/r2ad_bench_n2_fused$fz.err; echo "bench n2 fused=$fz rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2ad_bench_n2_fused$fz.json"))
print("fused=$fz", round(d["ms_per_step"],3), d["value"], d["config"].get("halo"), d["clocks"]["sm_mhz"], {k:(round(v["ms_total"]/v["launches"],4), v["launches"]) for k,v in d["roofline"]["kernels"].items()})
PY
done
