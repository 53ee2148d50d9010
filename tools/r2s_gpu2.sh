#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi.py -m gpu -q -s > gpurun_out/r2s_pytest_multi_n2.log 2>&1; echo "pytest multi rc=$?"; grep -E "FAIL|PASS|worst|passed|failed" gpurun_out/r2s_pytest_multi_n2.log | tail -6
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --main-only --no-e2e --no-cpu > gpurun_out/r2s_bench_n2.json 2> gpurun_out/r2s_bench_n2.err; echo "bench n2 rc=$?"
python bench.py --main-only --no-e2e --no-cpu > gpurun_out/r2s_bench_n1.json 2> gpurun_out/r2s_bench_n1.err; echo "bench n1 rc=$?"
python - <<PY
import json
for f in ("gpurun_out/r2s_bench_n1.json","gpurun_out/r2s_bench_n2.json"):
    d=json.load(open(f))
    print(f, round(d["ms_per_step"],3), d["clocks"]["sm_mhz"], {k:(round(v["ms_total"]/d["steps"],3), v["launches"]//d["steps"]) for k,v in d["roofline"]["kernels"].items()}, d.get("parity",{}).get("ok"))
PY
