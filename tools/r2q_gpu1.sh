#!/bin/bash
mkdir -p gpurun_out
NSDG_MEVP_RING=1 ncu --set full --clock-control none --import-source on -k regex:mevp_march -s 2 -c 1 -o gpurun_out/r2q_mevp_march_ring -f \
   python tools/prof_target.py 4096 8 > gpurun_out/r2q_ncu_full_ring.log 2>&1; echo "ncu full ring rc=$?"
