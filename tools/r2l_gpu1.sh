#!/bin/bash
# round-2 profiles: launch list of the default step under ncu, full captures of the two marching kernels
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --main-only > gpurun_out/r2l_bench_plain.json 2> gpurun_out/r2l_bench_plain.err; echo "plain rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2l_launches.csv \
   python bench.py --steps 2 --warmup 1 --no-cpu --no-e2e --main-only > gpurun_out/r2l_ncu_launch.log 2>&1; echo "ncu launches rc=$?"
python tools/prof_target.py 4096 8 > /dev/null 2>&1; echo "prof_target rc=$?"
ncu --set full --clock-control none --import-source on -k regex:transport_march -s 1 -c 1 -o gpurun_out/r2l_transport_march -f \
   python tools/prof_target.py 4096 8 > gpurun_out/r2l_ncu_full_transport.log 2>&1; echo "ncu full transport rc=$?"
ncu --set full --clock-control none --import-source on -k regex:mevp_march -s 2 -c 1 -o gpurun_out/r2l_mevp_march -f \
   python tools/prof_target.py 4096 8 > gpurun_out/r2l_ncu_full_march.log 2>&1; echo "ncu full march rc=$?"
ls -la gpurun_out/*.ncu-rep
