#!/bin/bash
# scratch: 2-GPU weak-scaling A/B of the halo overlap and the band height
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 2 --steps 5 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; print(round(d['ms_per_step'],2), r['kernel'], round(r['avg_launch_ms'],4), d['clocks']['sm_mhz'])"; }
{
echo -n "N=2 default (overlap, band 64): "; run 29701
echo -n "N=2 NSDG_HALO_OVERLAP=0: "; NSDG_HALO_OVERLAP=0 run 29702
echo -n "N=2 band 128: "; NSDG_MEVP_MARCH=128 run 29703
echo -n "N=2 band 32: "; NSDG_MEVP_MARCH=32 run 29704
echo -n "N=2 tiled: "; NSDG_MEVP_MARCH=0 run 29705
} 2>&1 | tee gpurun_out/r1z_n2_ab.txt
