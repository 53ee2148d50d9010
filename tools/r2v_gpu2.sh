#!/bin/bash
mkdir -p gpurun_out
out=gpurun_out/r2v_strip_digest.txt; : > $out
python tools/strip_digest.py 8192 2>/dev/null | grep strip_digest >> $out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555"
$TR tools/strip_digest.py 8192 2>/dev/null | grep strip_digest >> $out
$TR tools/strip_digest.py 8192 halo_peer=-1 2>/dev/null | grep strip_digest >> $out
$TR tools/strip_digest.py 8192 march_rows=128 2>/dev/null | grep strip_digest >> $out
cat $out
