"""Small, bounded workload over every kernel of the library for compute-sanitizer runs
(memcheck / racecheck / synccheck / initcheck):
    compute-sanitizer --tool racecheck python tools/sanitize_target.py
Single rank: thermodynamics (with diagnostics, slab ocean, forcing blend), transport dG0/1/2
with the limiter, every sub-iteration variant, host-buffer steps.  Under torchrun (WORLD_SIZE > 1)
it runs one strip case per halo variant instead (peer stores and NCCL)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nextsimdg_b200 import synthetic as S  # noqa: E402
from nextsimdg_b200._abi import BC_BOX, BC_PERIODIC, DYN_SPLIT_MEVP, MevpParams, ThermoParams  # noqa: E402
from nextsimdg_b200.gpu import Context, GpuDynamics, GpuStep  # noqa: E402


def single():
    ctx = Context(0)
    st, fo = S.thermo_inputs(5000)
    prog = dict(st)
    prog["tice"] = prog.pop("tice0")
    step = GpuStep(ctx, 5000, 1, ThermoParams.defaults())
    step.setInitialData(prog, fo)
    step.iterate(600.0, 2, diagnostics=True)
    step.setSlabOcean(True)
    step.setForcingLevels(fo, fo)
    step.blendForcing(0.25)
    step.iterate(600.0, 1)
    step.run_host(600.0, 1, prog, fo)
    step.get_prognostic()
    for nc in (1, 3, 6):
        for (nx, ny, bc) in ((70, 37, BC_PERIODIC), (33, 40, BC_BOX)):
            F, hx, hy = S.box_case(nx, ny, nc)
            rng = np.random.default_rng(3)
            F["U"] = 0.05 * rng.standard_normal((ny, nx))
            F["V"] = 0.05 * rng.standard_normal((ny, nx))
            mp = MevpParams.defaults()
            mp.bc = bc
            variants = [dict(), dict(march_rows=5), dict(march_rows=-1), dict(pair=-1), dict(graph=1),
                        dict(transport_march=-1)]
            for tuning in variants:
                dyn = GpuDynamics(ctx, nx, ny, hx, hy, nc, mevp=mp, limiter=1, tuning=tuning)
                dyn.upload_all(F)
                dyn.step(120.0, 7)
                dyn.step(120.0, 6)
                out = {k: np.empty(dyn.field_shape(k)) for k in ("H", "A", "U", "V", "S11")}
                dyn.step_host(120.0, 4, {k: F[k] for k in ("UATM", "VATM")}, out, deferred=True)
                dyn.step_host(120.0, 4, {k: F[k] for k in ("UATM", "VATM")}, out, deferred=True)
                dyn.host_wait()
                dyn.mass("H")
                assert np.isfinite(out["U"]).all()
                dyn.close()
            dyn = GpuDynamics(ctx, nx, ny, hx, hy, nc, mevp=mp, flags=DYN_SPLIT_MEVP)
            dyn.upload_all(F)
            dyn.step(120.0, 3)
            dyn.close()
    ctx.close()
    print("sanitize target: single rank done")


def multi():
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import mgpu_worker as W
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = Context(local)
    uid = [Context.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ctx.comm_init(rank, world, uid[0])
    cases = [c for c in W.all_cases(world) if c[0] in (48, 300)][:3]
    ok, worst, _ = W.run_cases(ctx, rank, world, cases)
    if rank == 0:
        print("sanitize target: strips", "ok" if ok else "FAIL", worst)
    dist.barrier()
    dist.destroy_process_group()
    ctx.close()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    multi() if int(os.environ.get("WORLD_SIZE", "1")) > 1 else single()
