"""SHA-256 digest of the dynamics state after a few steps on fixed synthetic cases, for A/B checks
of kernel variants that must be bit-identical.  `python tools/dyn_digest.py [knob=value ...]`
(knobs of nsdg_dyn_tuning, e.g. march_rows=16 pair_tile=32x8); tests import digest()."""
import hashlib
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nextsimdg_b200 import synthetic as S  # noqa: E402
from nextsimdg_b200._abi import BC_BOX, BC_PERIODIC, MevpParams  # noqa: E402
from nextsimdg_b200.gpu import Context, GpuDynamics  # noqa: E402

CASES = ((300, 211, BC_PERIODIC, 10), (64, 48, BC_BOX, 7), (17, 5, BC_PERIODIC, 4),
         (1024, 1024, BC_BOX, 20))


def digest(ctx, tuning=None, cases=CASES):
    h = hashlib.sha256()
    for nx, ny, bc, nsub in cases:
        F, hx, hy = S.box_case(nx, ny, 6)
        rng = np.random.default_rng(5)
        F["U"] = 0.05 * rng.standard_normal((ny, nx))
        F["V"] = 0.05 * rng.standard_normal((ny, nx))
        mp = MevpParams.defaults()
        mp.bc = bc
        dyn = GpuDynamics(ctx, nx, ny, hx, hy, 6, mevp=mp, tuning=tuning)
        dyn.upload_all(F)
        for _ in range(2):
            dyn.step(120.0, nsub)
        for k in ("H", "A", "U", "V", "S11", "S12", "S22"):
            a = dyn.download(k)
            assert np.isfinite(a).all(), k
            h.update(a.tobytes())
        dyn.close()
    return h.hexdigest()


if __name__ == "__main__":
    knobs = {}
    for arg in sys.argv[1:]:
        k, _, v = arg.partition("=")
        knobs[k] = v if "x" in v else int(v)
    c = Context(0)
    print("digest", digest(c, knobs))
    c.close()
