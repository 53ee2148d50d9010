#!/bin/bash
# round-2 (a) single-GPU check after the ABI refactor (tuning struct, release, graphs, peer kernel
# variant compiled in): GPU tests, sanitizer passes on the small target, default bench
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2a_pytest.log
for tool in memcheck racecheck synccheck initcheck; do
  timeout 900 compute-sanitizer --tool $tool --log-file gpurun_out/r2a_sanitizer_$tool.log python tools/sanitize_target.py > gpurun_out/r2a_sanitizer_$tool.out 2>&1
  echo "$tool rc=$?"; tail -3 gpurun_out/r2a_sanitizer_$tool.log
done
python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"; cut -c1-800 gpurun_out/r2a_bench.json
