#!/bin/bash
mkdir -p gpurun_out
for pf in 0 1 2 4 8; do
echo "== NSDG_TRANSPORT_PREFETCH=$pf"
NSDG_TRANSPORT_PREFETCH=$pf python tools/transport_ab.py 4096 2>&1 | grep "4096x4096 bc=0 march"
done | tee gpurun_out/r2m_transport_prefetch.txt
