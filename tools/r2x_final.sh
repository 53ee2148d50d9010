#!/bin/bash
# final records of the round on N GPUs ($1): default bench line (+ GPU tests and reference arm at N = 1)
N=${1:-1}
mkdir -p gpurun_out
if [ "$N" -eq 1 ]; then
  # the no-allocation entry point alone in a fresh process first, then the whole suite (it runs last there)
  python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2x_pytest_noalloc_alone.log 2>&1; echo "pytest noalloc alone rc=$?"; tail -2 gpurun_out/r2x_pytest_noalloc_alone.log
  python -m pytest tests -m gpu -x -q > gpurun_out/r2x_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -3 gpurun_out/r2x_pytest_gpu.log
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2x_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/r2x_smoke.log
  python bench.py > gpurun_out/r2x_bench_n1.json 2> gpurun_out/r2x_bench_n1.err; echo "bench rc=$?"
  python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2x_bench_reference.json 2> gpurun_out/r2x_bench_reference.err; echo "reference rc=$?"; cat gpurun_out/r2x_bench_reference.json | cut -c1-600
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N > gpurun_out/r2x_bench_n$N.json 2> gpurun_out/r2x_bench_n$N.err; echo "bench n$N rc=$?"
fi
grep "\[bench\]" gpurun_out/r2x_bench_n$N.err | sort -u | tail -8
python - <<PY
import json
d=json.load(open("gpurun_out/r2x_bench_n$N.json"))
r=d["roofline"]
print("main", round(d["ms_per_step"],3), d["value"], d["config"].get("halo"), "frac", round(r["frac"],3), "march avg", round(r["avg_launch_ms"],4), "clk", d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
if "parity" in d: print("parity", d["parity"]["ok"], d["parity"]["max_rel"])
s=d.get("strong_16384sq")
if s: print("strong", round(s["ms_per_step"],3), s["value"], "march avg", round(s["roofline"]["avg_launch_ms"],4))
e=d["e2e"]; print("e2e", e["value"], e.get("roof",{}).get("frac"))
c=d.get("configs",{})
if c.get("thermo_4096sq"): print("thermo", c["thermo_4096sq"]["value"], c["thermo_4096sq"]["roofline"]["frac"])
if c.get("cone_1024sq"): print("cone", c["cone_1024sq"]["value"])
print("cpu", (d.get("cpu_baseline") or {}).get("value"), "wall", d.get("bench_wall_s"), "launches", d.get("gpu_launches"))
PY
