"""torchrun worker: strip-decomposed dG2 transport + mEVP on WORLD_SIZE GPUs vs the
single-domain CPU oracle (rank 0 checks).  Launched by tests/test_gpu_multi.py."""
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


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = Context(local)
    uid = [Context.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ctx.comm_init(rank, world, uid[0])
    ok = True
    cases = ((64, 16 * world + 3, BC_BOX, 0, 7),
             (45, 9 * world, BC_PERIODIC, 0, 4),
             (64, 16 * world, BC_BOX, DYN_SPLIT_MEVP, 5),
             (96, 5 * world + 1, BC_BOX, DYN_SINGLE_MEVP, 3),
             (40, 2 * world, BC_PERIODIC, 0, 6),
             # >= 3 tile rows per strip: boundary tile rows first, halo
             # exchange overlapped with the interior launch
             (48, 49 * world, BC_BOX, 0, 6),
             (37, 40 * world, BC_PERIODIC, 0, 7),
             (48, 40 * world, BC_BOX, 0, 6),
             (48, 49 * world, BC_PERIODIC, 0, 6))
    only = os.environ.get("MGPU_CASES")  # e.g. "5,6" while debugging
    if only:
        cases = [cases[int(q)] for q in only.split(",")]
    for (nx, ny, bc, flags, nsub) in cases:
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
        j0, nl = strips.strip_rows(ny, world, rank)
        dyn = GpuDynamics(ctx, nx, ny, hx, hy, 6, ny_local=nl, j0=j0, mevp=mp, flags=flags)
        dyn.upload_all({k: strips.cut(v, world, rank) for k, v in F.items()})
        for _ in range(2):
            dyn.step(120.0, nsub)
        names = ("H", "A", "U", "V", "S11", "S12", "S22")
        mine = {k: torch.from_numpy(dyn.download(k)).cuda() for k in names}
        gathered = {}
        for k in names:
            parts = []
            for r in range(world):
                _, nlr = strips.strip_rows(ny, world, r)
                shape = list(mine[k].shape)
                shape[-2] = nlr
                buf = mine[k] if r == rank else torch.empty(shape, dtype=torch.float64, device="cuda")
                dist.broadcast(buf, src=r)
                parts.append(buf.cpu().numpy())
            gathered[k] = np.concatenate(parts, axis=-2)
        if rank == 0:
            from oracle import cpu_oracle as O
            st = [F[k] for k in names]
            for _ in range(2):
                st = list(O.dyn_step(mp, *st, F["UATM"], F["VATM"], F["UOCN"], F["VOCN"], hx, hy,
                                     120.0, nsub))
            for k, r in zip(names, st):
                err = np.abs(gathered[k] - r).max() / max(np.abs(r).max(), 1e-300)
                print(f"grid {nx}x{ny} bc={bc} flags={flags} {k}: maxrel={err:.3e}")
                ok = ok and err < 1e-11
    if rank == 0:
        print("halo bytes sent by rank 0:", ctx.comm_bytes_sent)
        print("MGPU PASS" if ok else "MGPU FAIL")
    dist.barrier()
    dist.destroy_process_group()
    ctx.close()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
