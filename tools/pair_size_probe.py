"""Per-cell cost of the pair kernel on grids that fit in L2 vs. the 4096^2 grid (same tile count per
wave): separates DRAM latency/bandwidth from SM-internal limits.  python tools/pair_size_probe.py"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nextsimdg_b200 import synthetic as S
from nextsimdg_b200._abi import BC_BOX, MevpParams
from nextsimdg_b200.gpu import Context, GpuDynamics

ctx = Context(0)
for nx, ny in ((592, 512), (1184, 512), (2368, 1024), (4096, 4096)):
    F, hx, hy = S.box_case(nx, ny, 6)
    mp = MevpParams.defaults()
    mp.bc = BC_BOX
    dyn = GpuDynamics(ctx, nx, ny, hx, hy, 6, mevp=mp)
    dyn.upload_all(F)
    dyn.mevp_step(120.0, 20)
    ctx.timer_start()
    dyn.mevp_step(120.0, 200)
    ms = ctx.timer_stop()
    tiles = ((nx + 15) // 16) * ((ny + 15) // 16)
    print(f"{nx}x{ny}: {tiles} tiles = {tiles / 296:.2f} waves, {ms / 100 * 1e3:.1f} us per pair launch, "
          f"{ms / 200 * 1e6 / (nx * ny):.4f} ns per cell and sub-iteration, "
          f"pair-kernel planes {nx * ny * 8 * 32 / 1e6:.0f} MB")
    del dyn
