"""A/B of the dG2 transport stage kernels (marching vs one thread per cell): difference of the
results and time per H+A step.  `python tools/transport_ab.py [n]`"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nextsimdg_b200 import synthetic as S  # noqa: E402
from nextsimdg_b200._abi import BC_BOX, BC_PERIODIC, MevpParams  # noqa: E402
from nextsimdg_b200.gpu import Context, GpuDynamics  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
ctx = Context(0)
for (nx, ny, bc) in ((n, n, BC_PERIODIC), (300, 211, BC_BOX)):
    F, hx, hy = S.box_case(nx, ny, 6, separable=True)
    rng = np.random.default_rng(5)
    F["U"] = 0.05 * rng.standard_normal((ny, nx))
    F["V"] = 0.05 * rng.standard_normal((ny, nx))
    for k in range(1, 6):
        F["H"][k] = 0.01 * rng.standard_normal((ny, nx))
    mp = MevpParams.defaults()
    mp.bc = bc
    out = {}
    for name, tm in (("march", 1), ("tiled", -1)):
        dyn = GpuDynamics(ctx, nx, ny, hx, hy, 6, mevp=mp, tuning={"transport_march": tm, "graph": -1})
        dyn.upload_all(F)
        dyn.transport_step(10.0)
        out[name] = dyn.download("H")
        ctx.synchronize()
        ctx.profile_begin()
        for _ in range(20):
            dyn.transport_step(10.0)
        prof = ctx.profile_end()
        ms = sum(v[0] for v in prof.values()) / 20
        print(f"{nx}x{ny} bc={bc} {name}: {ms:.4f} ms per H+A step", {k: (round(v[0] / v[1], 4), v[1]) for k, v in prof.items()})
        dyn.close()
    d = np.abs(out["march"] - out["tiled"])
    scale = np.abs(out["tiled"]).max(axis=(1, 2))
    print("  max abs diff per component:", d.max(axis=(1, 2)), "relative:", d.max(axis=(1, 2)) / scale,
          "differing elements:", int((d > 0).sum()), "of", d.size)
    if (d > 0).any():
        j, i = np.unravel_index(np.argmax(d[0]), d[0].shape)
        print("  worst mean-coefficient cell (j, i) =", (j, i), "rows with differences:",
              np.flatnonzero((d > 0).any(axis=(0, 2)))[:10], "cols:", np.flatnonzero((d > 0).any(axis=(0, 1)))[:10])
ctx.close()
