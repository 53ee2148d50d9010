#!/bin/bash
# round-1 (z) single-GPU check: GPU tests, default bench, ncu launch list, full capture of the marching kernel
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r1z_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r1z_pytest.log
python bench.py > gpurun_out/r1z_bench.json 2> gpurun_out/r1z_bench.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/r1z_bench.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1z_launches.csv \
   python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e > gpurun_out/r1z_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:march -s 2 -c 1 -o gpurun_out/r1z_march -f \
   python tools/prof_target.py 4096 8 > gpurun_out/r1z_ncu_full.log 2>&1; echo "ncu full rc=$?"
