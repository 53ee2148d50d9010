"""nsdg_dyn_mass_noalloc (include/nsdg_gpu.h): the mass diagnostic with its scratch in a fixed
__device__ array of the library instead of cudaMalloc'ed memory (nextsimdg_b200/csrc/noalloc.cu).

No counterpart in the reference.  Checked against the existing entry point nsdg_dyn_mass through
the C ABI, on two calls in a row that share the scratch array.  The file name sorts last on
purpose: a failure here cannot hide another test.  The tests use
their own context and pass alone in a fresh process as well as after the rest of the suite.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL = 1e-11  # the tolerance of the DG-coefficient parity tests (tests/test_gpu_dynamics.py)
MAX_ROWS = 16384  # NSDG_NOALLOC_MAX_ROWS


@pytest.fixture()
def setup():
    import torch
    from nextsimdg_b200 import _abi
    from nextsimdg_b200.gpu import Context, GpuDynamics
    ctx = Context(0)
    nx, ny, hx, hy = 96, 64, 4000.0, 3000.0
    dyn = GpuDynamics(ctx, nx, ny, hx, hy, ncomp=6)
    r = np.random.default_rng(2024)
    fields = {}
    for k, (name, lo, hi) in enumerate((("H", 0.2, 2.0), ("A", 0.5, 1.0))):
        a = 0.05 * r.standard_normal((6, ny, nx))
        a[0] = r.uniform(lo, hi, (ny, nx))
        dyn.upload(name, a)
        fields[name] = a
    yield dict(torch=torch, lib=_abi.load(), abi=_abi, ctx=ctx, dyn=dyn, nx=nx, ny=ny, hx=hx, hy=hy,
               fields=fields)
    ctx.close()


def interior_ptr(s, name):
    """device pointer of the first interior row of component 0 (two ghost rows come first)"""
    p = ctypes.c_void_p()
    s["abi"].check(s["ctx"].h, s["lib"].nsdg_dyn_field_ptr(s["ctx"].h, ["H", "A"].index(name), ctypes.byref(p)))
    return p.value + 2 * s["nx"] * 8


def noalloc(s, ptr, nx, ny, pitch, out):
    return s["lib"].nsdg_dyn_mass_noalloc(s["ctx"].h, ptr, nx, ny, pitch, s["hx"], s["hy"], out.data_ptr())


def test_two_calls_in_a_row_match_the_allocating_entry_point(setup):
    s = setup
    torch, lib, ctx = s["torch"], s["lib"], s["ctx"]
    nx, ny = s["nx"], s["ny"]
    out = torch.full((4,), -1.0, dtype=torch.float64, device="cuda:0")  # the caller's own memory
    torch.cuda.synchronize()
    want = {k: s["dyn"].mass(k) for k in ("H", "A")}  # existing entry point (cudaMalloc'ed scratch)
    for k in ("H", "A"):
        exact = s["fields"][k][0].sum() * s["hx"] * s["hy"]
        assert abs(want[k] - exact) <= RTOL * abs(exact)

    # call 1: H, all rows
    assert noalloc(s, interior_ptr(s, "H"), nx, ny, nx, out[0:1]) == 0
    assert lib.nsdg_synchronize(ctx.h) == 0
    free1 = torch.cuda.mem_get_info(0)[0]
    # call 2 straight after: A, all rows (the scratch still holds H's row sums)
    assert noalloc(s, interior_ptr(s, "A"), nx, ny, nx, out[1:2]) == 0
    # calls 3, 4: fewer rows and a narrower window of H than the scratch was last filled for, then
    # call 1 again: stale entries must not leak and the result must repeat bit for bit
    assert noalloc(s, interior_ptr(s, "H") + (5 * nx + 7) * 8, nx - 20, ny - 24, nx, out[2:3]) == 0
    assert noalloc(s, interior_ptr(s, "H"), nx, ny, nx, out[3:4]) == 0
    assert lib.nsdg_synchronize(ctx.h) == 0
    free2 = torch.cuda.mem_get_info(0)[0]
    got = out.cpu().numpy()

    assert abs(got[0] - want["H"]) <= RTOL * abs(want["H"])
    assert abs(got[1] - want["A"]) <= RTOL * abs(want["A"])
    window = s["fields"]["H"][0, 5:5 + ny - 24, 7:7 + nx - 20].sum() * s["hx"] * s["hy"]
    assert abs(got[2] - window) <= RTOL * abs(window)
    assert got[3] == got[0]  # fixed summation order
    assert free2 >= free1, "device memory was allocated between two calls"


def test_full_scratch_and_rejections(setup):
    s = setup
    torch, lib, ctx = s["torch"], s["lib"], s["ctx"]
    out = torch.zeros(1, dtype=torch.float64, device="cuda:0")
    col = torch.rand(MAX_ROWS, 3, dtype=torch.float64, device="cuda:0")
    torch.cuda.synchronize()
    # the largest size: every entry of the scratch in use, pitch > nx
    assert noalloc(s, col.data_ptr(), 2, MAX_ROWS, 3, out) == 0
    assert lib.nsdg_synchronize(ctx.h) == 0
    want = col[:, :2].sum().item() * s["hx"] * s["hy"]
    assert abs(out.item() - want) <= RTOL * abs(want)
    # one row more than the scratch holds: refused, nothing launched, the context stays usable
    launches = lib.nsdg_launch_count(ctx.h)
    rc = noalloc(s, col.data_ptr(), 1, MAX_ROWS + 1, 1, out)
    assert rc == -1  # NSDG_EINVAL
    msg = lib.nsdg_last_error(ctx.h).decode()
    assert "NSDG_NOALLOC_MAX_ROWS" in msg and str(MAX_ROWS) in msg
    assert lib.nsdg_dyn_mass_noalloc(ctx.h, None, 4, 4, 4, 1.0, 1.0, out.data_ptr()) == -1
    assert lib.nsdg_dyn_mass_noalloc(ctx.h, col.data_ptr(), 4, 4, 4, 1.0, 1.0, None) == -1
    assert noalloc(s, col.data_ptr(), 4, 4, 3, out) == -1  # pitch < nx
    assert noalloc(s, col.data_ptr(), 0, 4, 3, out) == -1
    assert lib.nsdg_launch_count(ctx.h) == launches
    assert noalloc(s, col.data_ptr(), 2, 100, 3, out) == 0
    assert lib.nsdg_synchronize(ctx.h) == 0
    want = col[:100, :2].sum().item() * s["hx"] * s["hy"]
    assert abs(out.item() - want) <= RTOL * abs(want)
