"""Prints a SHA-256 digest of the dynamics state after a few steps on a fixed synthetic case.
Used for A/B checks of kernel variants selected by environment variables, which must be
bit-identical: `NSDG_PAIR_PIPE=0 python tools/dyn_digest.py` vs `NSDG_PAIR_PIPE=1 ...`."""
import hashlib
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nextsimdg_b200 import synthetic as S  # noqa: E402
from nextsimdg_b200._abi import BC_BOX, BC_PERIODIC, MevpParams  # noqa: E402
from nextsimdg_b200.gpu import Context, GpuDynamics  # noqa: E402

ctx = Context(0)
h = hashlib.sha256()
for nx, ny, bc, nsub in ((300, 211, BC_PERIODIC, 10), (64, 48, BC_BOX, 7), (17, 5, BC_PERIODIC, 4),
                         (1024, 1024, BC_BOX, 20)):
    F, hx, hy = S.box_case(nx, ny, 6)
    rng = np.random.default_rng(5)
    F["U"] = 0.05 * rng.standard_normal((ny, nx))
    F["V"] = 0.05 * rng.standard_normal((ny, nx))
    mp = MevpParams.defaults()
    mp.bc = bc
    dyn = GpuDynamics(ctx, nx, ny, hx, hy, 6, mevp=mp)
    dyn.upload_all(F)
    for _ in range(2):
        dyn.step(120.0, nsub)
    for k in ("H", "A", "U", "V", "S11", "S12", "S22"):
        a = dyn.download(k)
        assert np.isfinite(a).all(), k
        h.update(a.tobytes())
print("digest", h.hexdigest())
ctx.close()
