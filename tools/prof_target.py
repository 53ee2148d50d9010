"""Short profiling target for ncu: 3 thermo sweeps, one H+A transport step (3 RK stages) and a
few mEVP sub-iterations at n x n cells.  python tools/prof_target.py [n] [nsub]"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nextsimdg_b200 import synthetic as S
from nextsimdg_b200._abi import BC_BOX, MevpParams, ThermoParams
from nextsimdg_b200.gpu import Context, GpuDynamics, GpuStep

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nsub = int(sys.argv[2]) if len(sys.argv) > 2 else 6
ctx = Context(0)
nc = n * n
st, fo = S.thermo_inputs(nc)
step = GpuStep(ctx, nc, 1, ThermoParams.defaults())
prog = dict(st)
prog["tice"] = st["tice0"]
step.setInitialData(prog, fo)
step.iterate(600.0, 3)
ctx.synchronize()
del step
F, hx, hy = S.box_case(n, n, 6)
mp = MevpParams.defaults()
mp.bc = BC_BOX
dyn = GpuDynamics(ctx, n, n, hx, hy, 6, mevp=mp)
dyn.upload_all(F)
dyn.step(120.0, nsub)
ctx.synchronize()
print("prof_target ok, launches", ctx.launches)
