set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r1y_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r1y_pytest.log
for t in 16x16 32x8 16x12 16x16 32x8; do
  echo -n "$t "; NSDG_PAIR_TILE=$t python bench.py --steps 4 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; f=d['clocks']['sm_mhz']; print(round(d['ms_per_step'],2), r['kernel'], round(r['avg_launch_ms'],4), f, 'Mcyc', round(r['avg_launch_ms']*f/1e3,3))"
done 2>&1 | tee gpurun_out/r1y_tile_sweep.txt
