#!/bin/bash
# scratch: mEVP kernel sweep at 4096^2 (per-launch average from the event profiler)
run() { echo -n "$* : "; env "$@" python bench.py --steps 2 --warmup 1 --no-cpu --e2e-steps 1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print(round(d['ms_per_step'],2), r['kernel'], round(r['avg_launch_ms'],4), round(r['frac'],3))"; }
run NSDG_MEVP_PAIR=0
run NSDG_MEVP_PAIR=1 NSDG_PAIR_TILE=16x16
run NSDG_MEVP_PAIR=1 NSDG_PAIR_TILE=16x16m3
run NSDG_MEVP_PAIR=1 NSDG_PAIR_TILE=16x12
run NSDG_MEVP_PAIR=1 NSDG_PAIR_TILE=16x8
run NSDG_MEVP_PAIR=1 NSDG_PAIR_TILE=32x8
