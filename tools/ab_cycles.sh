#!/bin/bash
# same-box A/B of two builds of libnsdg_gpu.so (build_ab/old.so, build_ab/new.so): ms per step, ms per
# dominant launch, sampled SM clock and the resulting cycles per launch (the boxes are power-capped,
# so the clock differs between runs and milliseconds alone mislead)
for v in old new old new old new; do
  cp build_ab/$v.so nextsimdg_b200/csrc/libnsdg_gpu.so
  python bench.py --steps 6 --warmup 3 --no-cpu --no-e2e | python -c "
import json,sys; d=json.loads(sys.stdin.read()); ms=d['roofline']['avg_launch_ms']; f=d['clocks']['sm_mhz']
print('$v', round(d['ms_per_step'],2), round(ms,4), f, 'Mcycles/launch', round(ms*f/1e3,3))"
done
