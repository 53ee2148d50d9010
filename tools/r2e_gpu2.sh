#!/bin/bash
# 2 GPUs: strip parity, then A/B of the halo variants on the weak-scaling workload
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi.py -m gpu -q -x -s > gpurun_out/r2e_pytest_multi_n2.log 2>&1; echo "pytest multi rc=$?"; grep -E "FAIL|PASS|worst|Error|error" gpurun_out/r2e_pytest_multi_n2.log | tail -8
python -m pytest tests/test_gpu_dynamics.py -m gpu -q -x > gpurun_out/r2e_pytest_dyn.log 2>&1; echo "pytest dyn rc=$?"; tail -3 gpurun_out/r2e_pytest_dyn.log
run() { tag=$1; shift; env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --main-only --no-parity --no-e2e --no-cpu > gpurun_out/r2e_n2_$tag.json 2> gpurun_out/r2e_n2_$tag.err; echo "$tag rc=$?"; python - <<PY
import json
d=json.load(open("gpurun_out/r2e_n2_$tag.json"))
k=d["roofline"]["kernels"]
print("$tag", round(d["ms_per_step"],3), d["config"]["halo"], "march avg", round(d["roofline"]["avg_launch_ms"],4), "launches", d["roofline"]["launches"], "clk", d["clocks"]["sm_mhz"])
PY
}
run peer_cg X=1
run peer_nc NSDG_PEER_LDNC=1
run peer_rows128 NSDG_MEVP_MARCH=128
run peer_rows256 NSDG_MEVP_MARCH=256
run nccl NSDG_HALO_PEER=0
run nccl_rows128 NSDG_HALO_PEER=0 NSDG_MEVP_MARCH=128
