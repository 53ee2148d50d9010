#!/bin/bash
# final tree on two GPUs: strip parity / bit identity (the peer variant of the marching kernel changed with plane_row_above), default bench line
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -s > gpurun_out/r2aj_pytest_multi_n2.log 2>&1; echo "pytest multi rc=$?"; grep -E "FAIL|MGPU|worst|passed|failed" gpurun_out/r2aj_pytest_multi_n2.log | tail -6
bash tools/r2x_final.sh 2
