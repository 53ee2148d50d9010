"""Full-size strip parity as a size-independent property: the dG2 transport + mEVP state after a
few steps must be the SAME BITS whether the 4096 x (4096 N0) box test runs on one GPU or is cut
into strips over several (peer-store halo, NCCL halo).  Per-row 64-bit checksums of every plane
are gathered and hashed, so the digest does not depend on the number of ranks:
    python tools/strip_digest.py [ny_global]                      (one GPU, whole domain)
    torchrun --nproc-per-node N tools/strip_digest.py [ny_global] (N strips)
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nextsimdg_b200 import strips, synthetic as S  # noqa: E402
from nextsimdg_b200._abi import BC_BOX, MevpParams  # noqa: E402
from nextsimdg_b200.gpu import Context, GpuDynamics  # noqa: E402


def row_sums(a):
    """[planes][rows][nx] float64 -> [planes][rows] uint64: wrapping sum of the bit patterns"""
    a = np.ascontiguousarray(a).reshape(-1, a.shape[-2], a.shape[-1])
    return a.view(np.uint64).sum(axis=2, dtype=np.uint64)


def main():
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    nx = 4096
    ny = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    tuning = {}
    for arg in sys.argv[2:]:
        k, _, v = arg.partition("=")
        tuning[k] = int(v)
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = Context(local)
    if world > 1:
        uid = [Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(rank, world, uid[0])
    j0, nl = strips.strip_rows(ny, world, rank)
    L, Ly = 512e3, 512e3 * ny / nx
    mp = MevpParams.defaults()
    mp.bc = BC_BOX
    dyn = GpuDynamics(ctx, nx, ny, L / nx, Ly / ny, 6, ny_local=nl, j0=j0, mevp=mp, tuning=tuning)
    for c0 in range(0, nl, 1024):
        n = min(1024, nl - c0)
        Fc, _, _ = S.box_case(nx, ny, 6, L=L, Ly=Ly, rows=(j0 + c0, n), separable=True, state=False)
        for k, v in Fc.items():
            dyn.upload_rows(k, c0, v)
    for _ in range(2):
        dyn.step(120.0, 10)
    sums = []
    for k in ("H", "A", "U", "V", "S11", "S12", "S22"):
        parts = [row_sums(dyn.download_rows(k, c0, min(1024, nl - c0))) for c0 in range(0, nl, 1024)]
        sums.append(np.concatenate(parts, axis=1))
    mine = np.concatenate(sums, axis=0)  # [planes][nl]
    halo = ("peer" if ctx.halo_peer else "nccl") if world > 1 else "single domain"
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        full = np.concatenate(gathered, axis=1)
    else:
        full = mine
    if rank == 0:
        assert full.shape[1] == ny
        print(f"strip_digest nx={nx} ny={ny} ranks={world} halo={halo} tuning={tuning}: "
              f"{hashlib.sha256(full.tobytes()).hexdigest()}")
    dyn.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()


if __name__ == "__main__":
    main()
