#!/bin/bash
# is the low deferred e2e number of the N = 2 record (0.45 of the roof, below the blocking call) reproducible?
mkdir -p gpurun_out
for i in 1 2; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2955$i bench.py --gpus 2 --main-only --no-parity > gpurun_out/r2ae_bench_n2_run$i.json 2> gpurun_out/r2ae_bench_n2_run$i.err; echo "run $i rc=$?"
  python - <<PY
import json
d=json.load(open("gpurun_out/r2ae_bench_n2_run$i.json")); e=d["e2e"]
print("run $i", round(d["ms_per_step"],3), "e2e", e["value"], e["roof"]["frac"], "full_state", e["full_state"]["value"], "blocking", e["blocking"]["value"], "roundtrip", e["full_roundtrip"]["value"])
PY
done
