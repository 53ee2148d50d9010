#!/bin/bash
# N GPUs (N = $1): strip parity (world sizes up to N), then the default bench line at N
N=$1
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2i_topo_n$N.txt 2>&1
if [ "$N" -le 4 ]; then
python -m pytest tests/test_gpu_multi.py -m gpu -q -s > gpurun_out/r2i_pytest_multi_n$N.log 2>&1; echo "pytest multi rc=$?"; grep -E "FAIL|PASS|worst|passed|failed" gpurun_out/r2i_pytest_multi_n$N.log | tail -8
fi
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N > gpurun_out/r2i_bench_n$N.json 2> gpurun_out/r2i_bench_n$N.err; echo "bench n$N rc=$?"; grep "\[bench\]" gpurun_out/r2i_bench_n$N.err | sort -u | tail -6
python - <<PY
import json
d=json.load(open("gpurun_out/r2i_bench_n$N.json"))
print("weak", round(d["ms_per_step"],3), d["value"], d["config"]["halo"], "march avg", round(d["roofline"]["avg_launch_ms"],4), "clk", d["clocks"]["sm_mhz"], "parity", d["parity"]["ok"], d["parity"]["max_rel"])
s=d["strong_16384sq"]; print("strong", round(s["ms_per_step"],3), s["value"], "march avg", round(s["roofline"]["avg_launch_ms"],4), s["config"].get("band_rows"))
print("e2e", d["e2e"]["value"], d["e2e"]["roof"]["frac"], {k:round(v,1) for k,v in d["e2e"]["host_link_roof_gbs"].items() if isinstance(v,float)})
print("thermo", d["configs"]["thermo_4096sq"]["value"], d["configs"]["thermo_4096sq"]["roofline"]["frac"], "wall", d["bench_wall_s"])
PY
