"""torchrun worker: strip-decomposed dG2 transport + mEVP on WORLD_SIZE GPUs vs the
single-domain CPU oracle (rank 0 checks).  Launched by tests/test_gpu_multi.py and by bench.py's
N > 1 parity record (`run_cases`).  Every case runs under several halo / kernel variants chosen
through nsdg_dyn_set_tuning: all must agree with the oracle to 1e-11 and with one another bit
for bit (peer stores over NVLink vs pack / NCCL / unpack, with and without overlap, marching
kernel band heights, tiled pair kernel)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nextsimdg_b200 import strips, synthetic as S  # noqa: E402
from nextsimdg_b200._abi import (BC_BOX, BC_PERIODIC, DYN_SINGLE_MEVP, DYN_SPLIT_MEVP,  # noqa: E402
                                   MevpParams)
from nextsimdg_b200.gpu import Context, GpuDynamics  # noqa: E402

NAMES = ("H", "A", "U", "V", "S11", "S12", "S22")
VARIANTS = (("peer", {}),
            ("peer-rows16", {"march_rows": 16}),
            ("peer-rows3", {"march_rows": 3}),  # remainder bands of one row in several cases
            ("peer-ring", {"march_ring": 1}),  # ghost rows read by the bulk-copy engine
            ("nccl-overlap", {"halo_peer": -1}),
            ("nccl-overlap-rows16", {"halo_peer": -1, "march_rows": 16}),
            ("nccl-serial", {"halo_peer": -1, "halo_overlap": -1}),
            ("nccl-rows64", {"halo_peer": -1, "march_rows": 64}),
            ("tiled", {"march_rows": -1}),
            ("transport-march", {"transport_march": 1}),
            ("transport-tiled", {"transport_march": -1}),
            # stages 1 + 2 of the transport step in one launch: two-deep ghost rows, one exchange less
            ("transport-fused", {"transport_fused": 1}),
            ("transport-fused-march", {"transport_fused": 1, "transport_march": 1}),
            ("transport-unfused", {"transport_fused": -1}))


def all_cases(world):
    return ((64, 16 * world + 3, BC_BOX, 0, 7),
            (45, 9 * world, BC_PERIODIC, 0, 4),
            (64, 16 * world, BC_BOX, DYN_SPLIT_MEVP, 5),
            (96, 5 * world + 1, BC_BOX, DYN_SINGLE_MEVP, 3),
            (40, 2 * world, BC_PERIODIC, 0, 6),
            # >= 3 bands per strip with 16-row bands: boundary bands first, halo beside the interior
            (48, 49 * world, BC_BOX, 0, 6),
            (37, 40 * world, BC_PERIODIC, 0, 7),
            (48, 40 * world, BC_BOX, 0, 6),
            (48, 49 * world, BC_PERIODIC, 0, 6),
            # wider than one CTA of the marching kernel (4 warps x 28 columns), ragged strips
            (300, 33 * world + 1, BC_BOX, 0, 8))


def make_case(nx, ny, bc):
    F, hx, hy = S.box_case(nx, ny, 6)
    rng = np.random.default_rng(11)
    F["U"] = 0.05 * rng.standard_normal((ny, nx))
    F["V"] = 0.05 * rng.standard_normal((ny, nx))
    if bc == BC_BOX:
        F["U"][0, :] = F["U"][:, 0] = F["V"][0, :] = F["V"][:, 0] = 0.0
    F["S11"] = 50 * rng.standard_normal((3, ny, nx))
    F["S12"] = 50 * rng.standard_normal((3, ny, nx))
    F["S22"] = 50 * rng.standard_normal((3, ny, nx))
    F["A"][0] = rng.uniform(0.8, 1.0, (ny, nx))
    mp = MevpParams.defaults()
    mp.bc = bc
    mp.alpha = mp.beta = 300.0
    return F, hx, hy, mp


def run_strips(ctx, rank, world, case, F, hx, hy, mp, tuning, nsteps=2):
    """-> gathered global fields (on every rank), halo mode actually used"""
    nx, ny, bc, flags, nsub = case
    j0, nl = strips.strip_rows(ny, world, rank)
    dyn = GpuDynamics(ctx, nx, ny, hx, hy, 6, ny_local=nl, j0=j0, mevp=mp, flags=flags, tuning=tuning)
    mode = "peer" if ctx.halo_peer else "nccl"
    dyn.upload_all({k: strips.cut(v, world, rank) for k, v in F.items()})
    for _ in range(nsteps):
        dyn.step(120.0, nsub)
    mine = {k: torch.from_numpy(dyn.download(k)).cuda() for k in NAMES}
    gathered = {}
    for k in NAMES:
        parts = []
        for r in range(world):
            _, nlr = strips.strip_rows(ny, world, r)
            shape = list(mine[k].shape)
            shape[-2] = nlr
            buf = mine[k] if r == rank else torch.empty(shape, dtype=torch.float64, device="cuda")
            dist.broadcast(buf, src=r)
            parts.append(buf.cpu().numpy())
        gathered[k] = np.concatenate(parts, axis=-2)
    dyn.close()
    return gathered, mode


def run_cases(ctx, rank, world, cases, variants=VARIANTS, log=print):
    """-> (ok, max relative error against the oracle, list of per-case records)"""
    ok, worst, records = True, 0.0, []
    for case in cases:
        nx, ny, bc, flags, nsub = case
        F, hx, hy, mp = make_case(nx, ny, bc)
        ref = None
        if rank == 0:
            from oracle import cpu_oracle as O
            st = [F[k] for k in NAMES]
            for _ in range(2):
                st = list(O.dyn_step(mp, *st, F["UATM"], F["VATM"], F["UOCN"], F["VOCN"], hx, hy,
                                     120.0, nsub))
            ref = dict(zip(NAMES, st))
        first = None
        for vname, tuning in variants:
            if flags and tuning:
                continue  # the split / single variants do not use the pair kernels
            got, mode = run_strips(ctx, rank, world, case, F, hx, hy, mp, tuning)
            if first is None:
                first = got
            same = all(np.array_equal(got[k], first[k]) for k in NAMES)
            if rank == 0:
                err = max(np.abs(got[k] - ref[k]).max() / max(np.abs(ref[k]).max(), 1e-300) for k in NAMES)
                worst = max(worst, err)
                good = err < 1e-11 and same
                ok = ok and good
                records.append({"grid": f"{nx}x{ny}", "bc": int(bc), "flags": int(flags),
                                "variant": vname, "halo": mode, "max_rel": float(err),
                                "bit_identical_to_first_variant": bool(same)})
                log(f"grid {nx}x{ny} bc={bc} flags={flags} {vname} [{mode}]: maxrel={err:.3e} "
                    f"same_bits={same} {'ok' if good else 'FAIL'}")
    return ok, worst, records


def run_dg1_transport(ctx, rank, world, log=print):
    """dG1 tracers on strips, transport only (Heun): the one-launch step (tracer and stage buffer
    trade places, two-deep ghost rows) against two stage launches and the single-domain oracle."""
    ok, worst = True, 0.0
    for nx, ny in ((60, 7 * world + 2), (130, 2 * world)):
        F, hx, hy = S.box_case(nx, ny, 3)
        rng = np.random.default_rng(21)
        for k in range(1, 3):
            F["H"][k] = 0.02 * rng.standard_normal((ny, nx))
            F["A"][k] = 0.01 * rng.standard_normal((ny, nx))
        F["U"] = 0.3 * rng.standard_normal((ny, nx))
        F["V"] = 0.3 * rng.standard_normal((ny, nx))
        dt = 0.1 * min(hx, hy) / 1.5
        ref = None
        if rank == 0:
            from oracle import cpu_oracle as O
            ref = {k: F[k] for k in ("H", "A")}
            for _ in range(3):
                ref = {k: O.transport_step(ref[k], F["U"], F["V"], hx, hy, dt) for k in ("H", "A")}
        first = None
        for vname, tuning in (("fused", {"transport_fused": 1}), ("unfused", {"transport_fused": -1}),
                              ("unfused-march", {"transport_fused": -1, "transport_march": 1})):
            j0, nl = strips.strip_rows(ny, world, rank)
            dyn = GpuDynamics(ctx, nx, ny, hx, hy, 3, ny_local=nl, j0=j0, tuning=tuning)
            dyn.upload_all({k: strips.cut(F[k], world, rank) for k in ("H", "A", "U", "V")})
            for _ in range(3):
                dyn.transport_step(dt)
            got = {}
            for k in ("H", "A"):
                mine = torch.from_numpy(dyn.download(k)).cuda()
                parts = []
                for r in range(world):
                    _, nlr = strips.strip_rows(ny, world, r)
                    buf = mine if r == rank else torch.empty((3, nlr, nx), dtype=torch.float64, device="cuda")
                    dist.broadcast(buf, src=r)
                    parts.append(buf.cpu().numpy())
                got[k] = np.concatenate(parts, axis=-2)
            dyn.close()
            if first is None:
                first = got
            same = all(np.array_equal(got[k], first[k]) for k in got)
            if rank == 0:
                err = max(np.abs(got[k] - ref[k]).max() / np.abs(ref[k]).max() for k in got)
                worst = max(worst, err)
                good = err < 1e-11 and same
                ok = ok and good
                log(f"dG1 transport {nx}x{ny} {vname}: maxrel={err:.3e} same_bits={same} {'ok' if good else 'FAIL'}")
    return ok, worst


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = Context(local)
    uid = [Context.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ctx.comm_init(rank, world, uid[0])
    cases = all_cases(world)
    only = os.environ.get("MGPU_CASES")  # e.g. "5,6" while debugging
    if only:
        cases = [cases[int(q)] for q in only.split(",")]
    ok, worst, _ = run_cases(ctx, rank, world, cases)
    if not only:
        ok1, worst1 = run_dg1_transport(ctx, rank, world)
        ok, worst = ok and ok1, max(worst, worst1)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, src=0)
    ok = bool(flag.item())
    if rank == 0:
        print("halo bytes sent by rank 0:", ctx.comm_bytes_sent)
        print(f"worst maxrel {worst:.3e}")
        print("MGPU PASS" if ok else "MGPU FAIL")
    dist.barrier()
    dist.destroy_process_group()
    ctx.close()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
