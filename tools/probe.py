"""Scratch timing probe (not the bench): per-kernel CUDA-event timings at a given grid size."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from nextsimdg_b200 import synthetic as S
from nextsimdg_b200._abi import MevpParams, ThermoParams, BC_BOX
from nextsimdg_b200.gpu import Context, GpuDynamics, GpuStep

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
ctx = Context(0)
out = {}
# thermo
nc = n * n
st, fo = S.thermo_inputs(nc)
step = GpuStep(ctx, nc, 1, ThermoParams.defaults())
prog = dict(st); prog["tice"] = st["tice0"]
step.setInitialData(prog, fo)
step.iterate(600.0, 2)
ctx.timer_start(); step.iterate(600.0, 10); ms = ctx.timer_stop() / 10
out["thermo_ms"] = ms; out["thermo_GBs"] = nc * 160 / ms / 1e6
del step
F, hx, hy = S.box_case(n, n, 6)
mp = MevpParams.defaults(); mp.bc = BC_BOX
dyn = GpuDynamics(ctx, n, n, hx, hy, 6, mevp=mp)
dyn.upload_all(F)
dyn.step(120.0, 2)
ctx.timer_start(); dyn.transport_step(120.0); ms = ctx.timer_stop()
out["transportHA_ms"] = ms; out["transport_GBs"] = nc * 816 / ms / 1e6
ctx.timer_start(); dyn.mevp_step(120.0, 20); ms = ctx.timer_stop() / 20
out["mevp_substep_ms"] = ms; out["mevp_GBs"] = nc * 360 / ms / 1e6
print(json.dumps(out))
