#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r2r_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2r_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2r_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/r2r_smoke.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2r_bench_ref.json 2> gpurun_out/r2r_bench_ref.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/r2r_bench_ref.json
