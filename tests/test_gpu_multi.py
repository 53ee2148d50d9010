"""Strip decomposition on real GPUs: N ranks (one per GPU) must reproduce the single-domain
oracle under every halo variant (NVLink peer stores from inside the marching kernel; pack / NCCL
send+recv / unpack with and without overlap; tiled pair kernel), and the variants must agree
bit for bit.  Needs >= 2 GPUs: skipped on the 1-GPU box, run with
`gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu` (log under profiles/);
bench.py --gpus N repeats one overlap case before timing and prints it as its `parity` record."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4])
def test_strips_match_single_domain_oracle(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs at least {world} GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                        f"--nproc-per-node={world}", "--master-addr", "127.0.0.1", "--master-port",
                        str(29611 + world), os.path.join(ROOT, "tests", "mgpu_worker.py")],
                       capture_output=True, text=True, timeout=1200)
    print(r.stdout[-12000:], r.stderr[-3000:])
    assert r.returncode == 0 and "MGPU PASS" in r.stdout
