#!/bin/bash
mkdir -p gpurun_out
python tools/transport_tile_probe.py 16 > gpurun_out/r2h_transport_tile_probe.txt 2>&1; tail -8 gpurun_out/r2h_transport_tile_probe.txt
python tools/transport_ab.py 4096 > gpurun_out/r2h_transport_ab4096.txt 2>&1; tail -9 gpurun_out/r2h_transport_ab4096.txt
python -m pytest tests -m gpu -q > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2h_pytest.log
python bench.py --main-only --no-e2e --no-cpu > gpurun_out/r2h_bench_main.json 2> gpurun_out/r2h_bench_main.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2h_bench_main.json"))
print(d["ms_per_step"], d["clocks"]["sm_mhz"], {k:(round(v["ms_total"]/v["launches"],4), round(v.get("frac",0),3)) for k,v in d["roofline"]["kernels"].items()})
PY
