#!/bin/bash
# scratch: fused-kernel tile sweep at 4096^2 (per-launch average from the event profiler)
for t in ${TILES:-32x8 64x4 32x16 64x8 16x16}; do
  echo -n "$t "; NSDG_FUSED_TILE=$t python bench.py --steps 2 --warmup 1 --no-cpu --e2e-steps 1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print(round(d['ms_per_step'],2), r['kernel'], round(r['avg_launch_ms'],4), round(r['frac'],3))"
done
