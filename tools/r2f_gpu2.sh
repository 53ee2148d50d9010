#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi.py -m gpu -q -x -s > gpurun_out/r2f_pytest_multi_n2.log 2>&1; echo "pytest multi rc=$?"; grep -E "FAIL|PASS|worst|Error|error" gpurun_out/r2f_pytest_multi_n2.log | tail -8
MGPU_CASES=6,6,6,6,6,6,6,6,3,3,3,3 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 tests/mgpu_worker.py > gpurun_out/r2f_mgpu_repeat.log 2>&1; echo "repeat rc=$?"; grep -E "FAIL|PASS|worst" gpurun_out/r2f_mgpu_repeat.log | tail -5
python -m pytest tests/test_gpu_dynamics.py -m gpu -q > gpurun_out/r2f_pytest_dyn.log 2>&1; echo "pytest dyn rc=$?"; tail -4 gpurun_out/r2f_pytest_dyn.log
