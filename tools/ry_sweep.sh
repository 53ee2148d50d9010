#!/bin/bash
mkdir -p gpurun_out
run() { python bench.py --steps 5 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; f=d['clocks']['sm_mhz']; print(round(d['ms_per_step'],2), r['kernel'], round(r['avg_launch_ms'],4), f, 'Mcyc', round(r['avg_launch_ms']*f/1e3,3))"; }
{
for ry in ${RYS:-64 128 256 512 1024 64}; do echo -n "ry=$ry: "; NSDG_MEVP_MARCH=$ry run; done
} 2>&1 | tee gpurun_out/ry_sweep.txt
