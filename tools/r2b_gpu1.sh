#!/bin/bash
mkdir -p gpurun_out
python tools/transport_ab.py 1024 > gpurun_out/r2b_transport_ab.txt 2>&1; echo "ab rc=$?"; cat gpurun_out/r2b_transport_ab.txt | tail -12
python tools/transport_ab.py 4096 > gpurun_out/r2b_transport_ab4096.txt 2>&1; echo "ab rc=$?"; cat gpurun_out/r2b_transport_ab4096.txt | tail -12
python -m pytest tests -m gpu -q --deselect tests/test_gpu_dynamics.py::test_full_size_4096sq_is_the_tiled_small_problem > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2b_pytest.log
