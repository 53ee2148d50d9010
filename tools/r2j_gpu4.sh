#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi.py -m gpu -q -s > gpurun_out/r2j_pytest_multi_n4.log 2>&1; echo "pytest multi rc=$?"; grep -E "FAIL|PASS|worst|passed|failed" gpurun_out/r2j_pytest_multi_n4.log | tail -8
for rep in 1 2 3; do
MGPU_CASES=5,7,8,9,0,6 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29544 tests/mgpu_worker.py > gpurun_out/r2j_mgpu_repeat$rep.log 2>&1; echo "repeat$rep rc=$?"; grep -E "FAIL|PASS|worst" gpurun_out/r2j_mgpu_repeat$rep.log | tail -3
done
bash tools/r2i_gpuN.sh 4 2>&1 | grep -v "pytest\|MGPU\|worst\|passed"
