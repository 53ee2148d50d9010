"""Times the H+A transport step alone (3 SSP-RK3 stage launches, dG2) at n x n cells."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nextsimdg_b200 import synthetic as S
from nextsimdg_b200.gpu import Context, GpuDynamics

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
ctx = Context(0)
F, hx, hy = S.box_case(n, n, 6)
dyn = GpuDynamics(ctx, n, n, hx, hy, 6)
dyn.upload_all(F)
for _ in range(3):
    dyn.transport_step(120.0)
ctx.timer_start()
for _ in range(10):
    dyn.transport_step(120.0)
ms = ctx.timer_stop() / 10
print(f"transport H+A step: {ms:.3f} ms, {n * n * 816 / ms / 1e6:.0f} GB/s algorithmic")
