"""Where does the marching transport kernel on a tiled big grid differ from the small grid?"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nextsimdg_b200 import synthetic as S  # noqa: E402
from nextsimdg_b200._abi import BC_PERIODIC, MevpParams  # noqa: E402
from nextsimdg_b200.gpu import Context, GpuDynamics  # noqa: E402

ctx = Context(0)
m, reps = 256, int(sys.argv[1]) if len(sys.argv) > 1 else 16
n = m * reps
F, hx, hy = S.box_case(m, m, 6)
r = np.random.default_rng(17)
F["U"] = 0.05 * r.standard_normal((m, m))
F["V"] = 0.05 * r.standard_normal((m, m))
mp = MevpParams.defaults()
mp.bc = BC_PERIODIC
for tm in (1, -1):
    small = GpuDynamics(ctx, m, m, hx, hy, 6, mevp=mp, tuning={"transport_march": tm, "graph": -1})
    small.upload_all(F)
    small.transport_step(120.0)
    a = small.download("H")
    small.close()
    big = GpuDynamics(ctx, n, n, hx, hy, 6, mevp=mp, tuning={"transport_march": tm, "graph": -1})
    big.upload_all({k: np.tile(v, (1, reps, reps) if v.ndim == 3 else (reps, reps)) for k, v in F.items()})
    big.transport_step(120.0)
    b = big.download("H")
    big.close()
    d = b != np.tile(a, (1, reps, reps))
    print("transport_march", tm, "differing:", int(d.sum()), "of", d.size)
    if d.any():
        rows = np.flatnonzero(d.any(axis=(0, 2)))
        cols = np.flatnonzero(d.any(axis=(0, 1)))
        print("  rows:", len(rows), rows[:40], "...")
        print("  cols:", len(cols), cols[:40], "...")
        print("  cols mod 28 histogram:", np.bincount(cols % 28, minlength=28))
        rr = np.flatnonzero(d.any(axis=(0, 2)))
        print("  per-row count (first 12 rows with differences):", [(int(q), int(d[:, q, :].sum())) for q in rr[:12]])
        print("  max abs diff:", np.abs(b - np.tile(a, (1, reps, reps))).max())
ctx.close()
