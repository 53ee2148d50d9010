#!/bin/bash
# scratch: marching pair kernel -- bit-identity against the tiled pair kernel, then a sweep of
# the compiled configurations at 4096^2 (per-launch average from the event profiler)
mkdir -p gpurun_out
CFGS=${CFGS:-0 1 2 3 4 5 6 7}
{
echo -n "tiled      "; python tools/dyn_digest.py
for c in ${DIGEST_CFGS:-0 1 5}; do
  echo -n "march64 cfg$c "; NSDG_MEVP_MARCH=64 NSDG_MARCH_CFG=$c python tools/dyn_digest.py
done
run() { python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; f=d['clocks']['sm_mhz']; print(round(d['ms_per_step'],2), r['kernel'], round(r['avg_launch_ms'],4), f, 'Mcyc', round(r['avg_launch_ms']*f/1e3,3))"; }
echo -n "tiled 16x16: "; run
for c in $CFGS; do
  echo -n "march ry=64 cfg=$c: "; NSDG_MEVP_MARCH=64 NSDG_MARCH_CFG=$c run
done
for ry in ${RYS:-32}; do
  echo -n "march ry=$ry cfg=${RYCFG:-1}: "; NSDG_MEVP_MARCH=$ry NSDG_MARCH_CFG=${RYCFG:-1} run
done
} 2>&1 | tee gpurun_out/march_sweep.txt
if [ -n "$NCU_CFG" ]; then
  NSDG_MEVP_MARCH=64 NSDG_MARCH_CFG=$NCU_CFG ncu --set full --clock-control none --import-source on -k regex:march -s 2 -c 1 \
     -o gpurun_out/march_cfg$NCU_CFG -f python tools/prof_target.py 4096 8 > gpurun_out/ncu_march.log 2>&1
  tail -2 gpurun_out/ncu_march.log
fi
