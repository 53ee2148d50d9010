#!/bin/bash
mkdir -p gpurun_out
python tools/transport_ab.py 4096 > gpurun_out/r2c_transport_ab4096.txt 2>&1; echo "ab rc=$?"; cat gpurun_out/r2c_transport_ab4096.txt | tail -12
python -m pytest tests -m gpu -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2c_pytest.log
python bench.py > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/r2c_bench.err
