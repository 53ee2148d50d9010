#!/bin/bash
# 2 GPUs: strip parity under every halo variant, then the default bench line at N = 2
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2d_topo.txt 2>&1
python -m pytest tests/test_gpu_multi.py -m gpu -q -x -s > gpurun_out/r2d_pytest_multi_n2.log 2>&1; echo "pytest multi rc=$?"; grep -E "FAIL|PASS|worst|Error|error" gpurun_out/r2d_pytest_multi_n2.log | tail -20
python -m pytest tests/test_gpu_dynamics.py -m gpu -q -x > gpurun_out/r2d_pytest_dyn.log 2>&1; echo "pytest dyn rc=$?"; tail -3 gpurun_out/r2d_pytest_dyn.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 > gpurun_out/r2d_bench_n2.json 2> gpurun_out/r2d_bench_n2.err; echo "bench n2 rc=$?"; tail -5 gpurun_out/r2d_bench_n2.err; cut -c1-300 gpurun_out/r2d_bench_n2.json
