"""TIMING EXPERIMENT (results are meaningless): pair kernel with parts switched off."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nextsimdg_b200 import synthetic as S
from nextsimdg_b200._abi import BC_PERIODIC, MevpParams
from nextsimdg_b200.gpu import Context, GpuDynamics
ctx = Context(0)
n = 4096
F, hx, hy = S.box_case(n, n, 6)
mp = MevpParams.defaults(); mp.bc = BC_PERIODIC
dyn = GpuDynamics(ctx, n, n, hx, hy, 6, mevp=mp)
dyn.upload_all(F)
dyn.mevp_step(120.0, 10)
ctx.timer_start(); dyn.mevp_step(120.0, 100); ms = ctx.timer_stop()
print(f"abl={os.environ.get('NSDG_PAIR_ABLATE','0')}: {ms/50*1e3:.1f} us per pair launch")
