#!/bin/bash
# scratch: marching pair kernel -- bit-identity against the tiled pair kernel, then a sweep of
# band heights / register budgets at 4096^2 (per-launch average from the event profiler)
mkdir -p gpurun_out
{
echo -n "tiled      "; python tools/dyn_digest.py
for mb in 3 2 4; do
  echo -n "march64 mb$mb "; NSDG_MEVP_MARCH=64 NSDG_MARCH_MINB=$mb python tools/dyn_digest.py
done
run() { python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; f=d['clocks']['sm_mhz']; print(round(d['ms_per_step'],2), r['kernel'], round(r['avg_launch_ms'],4), f, 'Mcyc', round(r['avg_launch_ms']*f/1e3,3))"; }
echo -n "tiled 16x16: "; run
for mb in 3 30 2 4 5; do
  echo -n "march ry=64 mb=$mb: "; NSDG_MEVP_MARCH=64 NSDG_MARCH_MINB=$mb run
done
for ry in 32 128; do
  echo -n "march ry=$ry mb=3: "; NSDG_MEVP_MARCH=$ry NSDG_MARCH_MINB=3 run
done
} 2>&1 | tee gpurun_out/march_sweep.txt
