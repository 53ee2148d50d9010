"""Strip decomposition of the structured grid over ranks (rows of the slow index y).

Host-side logic shared by the GPU path (bench.py, GpuDynamics) and the gloo CPU tests: which
rows a rank owns, who its ring neighbours are, and how a global array is cut into strips.
The exchange itself is done by the C ABI (NCCL send/recv, csrc/comm.cu) on the GPU.
"""
import numpy as np


def strip_rows(ny, nranks, rank):
    """(j0, ny_local) of `rank`: contiguous rows, remainder spread over the first ranks."""
    if not (0 <= rank < nranks):
        raise ValueError("rank out of range")
    if ny < 2 * nranks:
        raise ValueError(f"need at least 2 rows per rank: ny={ny}, nranks={nranks}")
    base, rem = divmod(ny, nranks)
    j0 = rank * base + min(rank, rem)
    return j0, base + (1 if rank < rem else 0)


def neighbours(nranks, rank):
    """(down, up) ring neighbours: `down` owns the rows below this strip (periodic in y)."""
    return (rank - 1) % nranks, (rank + 1) % nranks


def cut(global_array, nranks, rank):
    """Rows of this rank from an array whose last two axes are (ny, nx)."""
    ny = global_array.shape[-2]
    j0, nl = strip_rows(ny, nranks, rank)
    return np.ascontiguousarray(global_array[..., j0:j0 + nl, :])


def halo_layout(nx, nplanes):
    """Packed halo message: plane-major, one row of nx doubles per plane.  Returns the number
    of doubles and the (start, stop) slice of every plane; bit-exact contract of csrc/comm.cu
    pack_rows/unpack_rows."""
    return nplanes * nx, [(p * nx, (p + 1) * nx) for p in range(nplanes)]


def pad(local):
    """Add the two ghost rows (row 0 and row ny_local+1) around a strip (..., ny_local, nx)."""
    shape = list(local.shape)
    shape[-2] += 2
    out = np.zeros(shape, dtype=local.dtype)
    out[..., 1:-1, :] = local
    return out


def exchange_ghosts(padded, send_recv, which=3):
    """Fill ghost rows of a padded strip through `send_recv(to_up, to_down) -> (from_down,
    from_up)` where every message is the packed layout of halo_layout().  Used by the gloo
    tests to emulate exactly what nsdg_halo_exchange does on the GPU."""
    planes = padded.reshape(-1, padded.shape[-2], padded.shape[-1])
    to_up = np.ascontiguousarray(planes[:, -2, :]).ravel()  # top owned row -> up's lower ghost
    to_down = np.ascontiguousarray(planes[:, 1, :]).ravel()  # bottom owned row -> down's upper ghost
    from_down, from_up = send_recv(to_up, to_down)
    nx = padded.shape[-1]
    if which & 1:
        planes[:, 0, :] = from_down.reshape(-1, nx)
    if which & 2:
        planes[:, -1, :] = from_up.reshape(-1, nx)
    return padded
