#!/usr/bin/env python
"""Benchmark of the north-star unit: one dG2 transport (H, A; SSP-RK3) + 100 mEVP sub-iterations
step on a synthetic box-test grid.  `python bench.py --gpus N --steps K --warmup W` prints ONE
JSON line.

metric   cell-updates/s = cells of the whole job x steps / max-over-ranks device time
N = 1    BASELINE.json configs[2]: 4096 x 4096 cells, state resident in HBM
N > 1    weak scaling: every rank owns a 4096 x 4096 strip of a 4096 x (4096 N) grid (ring of
         strips over y; the sub-cycle's halo goes by NVLink peer stores from inside the kernel,
         the transport stages' by NCCL send/recv)
e2e      same step through the C ABI with HOST (pinned) buffers (nsdg_dyn_step_host_async): every
         step uploads its forcing and downloads the fields H, A, U, V inside the timed region
The same line also carries, as sub-records measured right after the main workload:
  parity            (N > 1) one overlap strip case against the CPU oracle under every halo variant
  configs.thermo_4096sq   BASELINE configs[3]: column thermodynamics, 4096^2 cells per GPU, CPU arm
                          = the reference's own sources (oracle/_ref)
  configs.cone_1024sq     BASELINE configs[1]: 1024^2 dG1 rotating cone (N = 1 only)
  strong_16384sq    BASELINE configs[4]: 16384^2 split over the N ranks (strong scaling; at N = 1
                    the grid takes 150 GB and runs after everything else has been freed)
`--workload thermo|cone|16384sq` prints that record alone as the main line.
--impl reference   the CPU arm: the oracle's restatement of the same step on the host cores
         (the reference snapshot has no dynamics code to run)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# stdout carries exactly ONE JSON line.  Libraries write there too (this image sets
# NCCL_DEBUG=VERSION and NCCL prints its banner to stdout), so file descriptor 1 is pointed at
# stderr for the whole run and the JSON line goes to the saved descriptor.
_JSON_FD = os.dup(1)
os.dup2(2, 1)


def emit(line):
    os.write(_JSON_FD, (json.dumps(line) + "\n").encode())


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


NSUB = 100
DT = 120.0
NCOMP = 6
# algorithmic bytes per cell (counting rule of SURVEY.md 8d: every distinct
# array element one cell-update must read or write counts once, neighbour re-reads are free)
BYTES_STRESS = (9 + 2 + 4 + 9) * 8  # r S(9) u,v(2) P_g(4)  w S(9)
BYTES_VELOCITY = (7 + 2 + 2 + 4 + 2) * 8  # r S(7 of 9) u,v u_o,v_o node constants(4)  w u,v
BYTES_FUSED = (9 + 2 + 4 + 2 + 4 + 9 + 2) * 8  # one fused sub-iteration: 32 doubles
BYTES_TRANSPORT_STEP = (2 * 8 * NCOMP + 6) * 8  # H and A, SSP-RK3, shared velocity reads
# stages 1 + 2 in one launch (transport_pair_march_kernel, single rank): per tracer 6 in + 6 out,
# then stage 3 with 6 + 6 in, 6 out; the velocity is read by two launches instead of three
BYTES_TRANSPORT_STEP_FUSED = (2 * 5 * NCOMP + 4) * 8
BYTES_PREPARE = (12 + 4 + 8 + 4) * 8  # P_g from H,A; node constants from H0,A0,u_a,u,u_o
# the pair kernels run TWO sub-iterations per pass over HBM (temporal blocking): the compulsory
# DRAM bytes of a launch are those of ONE fused sub-iteration
SUBITS_PER_LAUNCH = {"mevp_pair_kernel": 2, "mevp_march_kernel": 2}
# dram__bytes_read.sum + dram__bytes_write.sum per launch at 4096^2 cells (one launch over the whole
# grid) from ncu --set full captures; the file each number comes from travels with it
NCU_TRAFFIC_4096SQ = {
    "mevp_march_kernel": (4.2980e9, "profiles/r1z_march_ncu_summary.txt"),
    "mevp_pair_kernel": (4.2719e9, "profiles/r1t_ncu_full_summary.txt"),
    "mevp_fused_kernel": (4.5397e9, "profiles/r1d_ncu_full_summary.txt"),
    "transport_stage_kernel": (3.7266e9, "profiles/r1w_ncu_full_summary.txt"),
}
BYTES_STEP_SPLIT = BYTES_TRANSPORT_STEP + BYTES_PREPARE + NSUB * (BYTES_STRESS + BYTES_VELOCITY)
BYTES_STEP_FUSED = BYTES_TRANSPORT_STEP + BYTES_PREPARE + NSUB * BYTES_FUSED
# what one pass per PAIR has to move: the DRAM view of the step with the pair kernels
BYTES_STEP_PAIR = BYTES_TRANSPORT_STEP + BYTES_PREPARE + (NSUB // 2) * BYTES_FUSED
METRIC = "cell-updates/sec (dG2 transport+mEVP step)"
PARITY_NOTE = "unpinned against draenog/nextsimdg (no dynamics in the reference snapshot)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap,power.draw,power.limit")

    def __init__(self, index, active=True):
        self.index, self.proc, self.lines, self.active = index, None, [], active

    def start(self):
        if not self.active:
            return self
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self, note=None):
        if not self.active:
            return None
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw, plim = [], None, set(), [], None
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for nm, val in zip(names, f[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
            try:
                pw.append(float(f[6]))
                plim = float(f[7])
            except (IndexError, ValueError):
                pass
        out = {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
               "reasons": sorted(reasons), "samples": len(sm),
               "power_w": float(np.median(pw)) if pw else None, "power_limit_w": plim}
        if note:
            out["sampled_over"] = note
        return out


def bind_to_gpu_numa_node(local):
    """Pin this rank (and therefore the pinned host buffers it first-touches) to the CPUs next to
    its GPU: with 8 ranks on a two-socket host the staging traffic otherwise crosses the socket
    interconnect.  NSDG_BENCH_NUMA=0 disables.  Returns a short description for the JSON line."""
    if os.environ.get("NSDG_BENCH_NUMA", "1") == "0":
        return "off"
    try:
        import torch
        pr = torch.cuda.get_device_properties(local)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        base = f"/sys/bus/pci/devices/{bdf}"
        node = open(base + "/numa_node").read().strip()
        if int(node) < 0:
            return "sysfs reports no NUMA node for the GPU (node -1): not bound"
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return f"node {node}: no usable cpus"
        os.sched_setaffinity(0, cpus)
        return f"node {node}, {len(cpus)} cpus"
    except Exception as e:  # not fatal: the numbers are just measured unbound
        return f"unavailable ({type(e).__name__})"


class Env:
    """One process per GPU: rank / world from torchrun's environment, NCCL through
    torch.distributed for the barrier and the max-over-ranks reductions."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus:
            raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={self.world}: launch with torchrun")
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device; this framework has no CPU path")
        torch.cuda.set_device(self.local)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        self.numa = bind_to_gpu_numa_node(self.local) if self.world > 1 else "single rank: not bound"

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max(self, x):
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum(self, x):
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def pinned(self, a):
        t = self.torch.empty(a.shape, dtype=self.torch.float64).pin_memory()
        t.numpy()[...] = a
        return t

    def context(self):
        from nextsimdg_b200.gpu import Context
        ctx = Context(self.local)
        if self.world > 1:
            uid = [Context.comm_unique_id() if self.rank == 0 else None]
            self.dist.broadcast_object_list(uid, src=0)
            ctx.comm_init(self.rank, self.world, uid[0])
        return ctx

    def close(self):
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------
# CPU arms
def cpu_dyn_step(F, hx, hy, mp, nsteps=1):
    """The oracle's restatement of the bench step on the host cores (OpenMP, all threads)."""
    from oracle import cpu_oracle as O
    st = [F[k] for k in ("H", "A", "U", "V", "S11", "S12", "S22")]
    t0 = time.perf_counter()
    for _ in range(nsteps):
        st = list(O.dyn_step(mp, *st, F["UATM"], F["VATM"], F["UOCN"], F["VOCN"], hx, hy, DT, NSUB))
    return time.perf_counter() - t0


def cpu_dyn_sample(nx, ny_full, budget_s, nsteps, mp):
    """-> (rows of the sample, F, hx, hy): the largest strip [0, ny) of the nx x ny_full box test
    whose `nsteps` CPU steps fit the time budget, from a short calibration run."""
    from nextsimdg_b200 import synthetic as S
    ncal = min(ny_full, 128)
    F, hx, hy = S.box_case(nx, ny_full, NCOMP, rows=(0, ncal), separable=True)
    secs = cpu_dyn_step(F, hx, hy, mp, 1)
    rate = nx * ncal / secs
    ny = int(min(ny_full, max(ncal, budget_s * rate / (nsteps * nx))))
    if ny < ny_full:
        ny -= ny % 32
    F, hx, hy = S.box_case(nx, ny_full, NCOMP, rows=(0, ny), separable=True)
    return ny, F, hx, hy


def run_reference(args):
    """--impl reference: CPU arm.  DG transport + mEVP do not exist in the reference snapshot
    (CMakeLists.txt:43-46), so the arm is the oracle port of the same step, all host threads.
    It runs the configuration it prints: the full 4096 x 4096 grid when K + 1 steps of it fit the
    time budget (about 18 s per step on 16 cores), else the largest strip of whole rows that
    does; the strip is named in config.cpu_sample and same_config is stated, not implied."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from nextsimdg_b200._abi import BC_BOX, MevpParams
    mp = MevpParams.defaults()
    mp.bc = BC_BOX
    cores = os.cpu_count() or 1
    nx = args.nx
    ny_full = args.ny
    wsteps = 1 if args.warmup > 0 else 0
    if args.cpu_sample:
        ny = min(args.cpu_sample, ny_full)
        from nextsimdg_b200 import synthetic as S
        F, hx, hy = S.box_case(nx, ny_full, NCOMP, rows=(0, ny), separable=True)
    else:
        ny, F, hx, hy = cpu_dyn_sample(nx, ny_full, args.cpu_budget, args.steps + wsteps, mp)
    for _ in range(wsteps):
        cpu_dyn_step(F, hx, hy, mp)
    secs = cpu_dyn_step(F, hx, hy, mp, args.steps)
    value = nx * ny * args.steps / secs
    full = ny == ny_full and args.gpus == 1
    sample = (f"rows [0, {ny}) of the {nx}x{ny_full} box test ({'the whole grid' if ny == ny_full else 'a strip'}"
              f"{'' if args.gpus == 1 else ' of one rank'}), {args.steps} full steps (transport + {NSUB} "
              f"sub-iterations), {wsteps} warm-up step")
    cfg = workload_config(args, args.gpus)
    cfg["cpu_sample"] = {"nx": nx, "ny": ny, "same_config": bool(full)}
    line = {"impl": "reference", "metric": METRIC,
            "value": value, "unit": "cell-updates/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": value, "unit": "cell-updates/s", "cores": cores,
                             "kind": "port", "sample": sample,
                             "note": "parity unpinned against draenog/nextsimdg: the reference "
                                     "snapshot has no dynamics code; this is oracle/ (C, OpenMP)"},
            "e2e": {"value": value, "unit": "cell-updates/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}
    emit(line)


def workload_config(args, n):
    if args.workload == "16384sq":
        name = "synthetic 16384x16384 dG2 transport + mEVP (100 substeps), strips over y"
        nx, ny = 16384, 16384
    elif n == 1:
        name = f"synthetic {args.nx}x{args.ny} dG2 transport + mEVP rheology subcycle (100 substeps), 1 GPU"
        nx, ny = args.nx, args.ny
    else:
        name = (f"synthetic {args.nx}x{args.ny * n} dG2 transport + mEVP (100 substeps): "
                f"{n} strips of {args.nx}x{args.ny}, ring halo")
        nx, ny = args.nx, args.ny * n
    return {"workload": name, "nx": nx, "ny": ny, "substeps": NSUB, "dt_s": DT, "dg": "dG2",
            "rk": "SSP-RK3", "bc": "box", "l2": "working set >> 126 MB L2 (>= 10 GB per GPU)",
            "parity": PARITY_NOTE}


# ---------------------------------------------------------------------------------------------
def kernel_table(prof, ms, cells_local, steps, peak):
    """Per-kernel CUDA-event totals of the timed region with the algorithmic (compulsory DRAM)
    bytes each kernel class has to move in it, and the fraction of the HBM peak that gives."""
    pairs = steps * (NSUB // 2)
    tm = prof.get("transport_march_kernel")
    transport = BYTES_TRANSPORT_STEP_FUSED if tm and tm[1] == 2 * steps else BYTES_TRANSPORT_STEP
    total_bytes = {
        # one pass over HBM per launch (= per pair for the pair kernels)
        "mevp_march_kernel": BYTES_FUSED * cells_local * pairs,
        "mevp_pair_kernel": BYTES_FUSED * cells_local * pairs,
        "transport_stage_kernel": BYTES_TRANSPORT_STEP * cells_local * steps,
        "transport_march_kernel": transport * cells_local * steps,
        "pressure_kernel": (12 + 4) * 8 * cells_local * steps,
        "nodal_kernel": (8 + 4) * 8 * cells_local * steps,
    }
    out = {}
    for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0]):
        rec = {"ms_total": v[0], "launches": v[1], "share": v[0] / ms}
        b = total_bytes.get(k)
        if k == "mevp_fused_kernel":
            b = BYTES_FUSED * cells_local * v[1]
        elif k == "stress_kernel":
            b = BYTES_STRESS * cells_local * v[1]
        elif k == "velocity_kernel":
            b = BYTES_VELOCITY * cells_local * v[1]
        if b and v[0] > 0:
            rec["dram_bytes"] = b
            rec["gbs"] = b / (v[0] * 1e-3) / 1e9
            rec["frac"] = rec["gbs"] / peak
        out[k] = rec
    return out


def measure_north_star(args, env, strong=False, with_e2e=True, with_cpu=True, steps=None, warmup=None):
    """The dG2 transport + mEVP step: resident-state timing, e2e legs, roofline, CPU baseline.
    strong=True: BASELINE configs[4], 16384^2 split over the ranks, filled chunk by chunk."""
    from nextsimdg_b200 import strips, synthetic as S
    from nextsimdg_b200._abi import BC_BOX, MevpParams
    from nextsimdg_b200.gpu import GpuDynamics
    torch, rank, world = env.torch, env.rank, env.world
    steps = args.steps if steps is None else steps
    warmup = args.warmup if warmup is None else warmup
    ctx = env.context()
    mp = MevpParams.defaults()
    mp.bc = BC_BOX
    if strong:
        nx, ny_glob = 16384, 16384
        j0, nyl = strips.strip_rows(ny_glob, world, rank)
        Lx = Ly = 512e3
    else:
        nx, nyl = args.nx, args.ny
        ny_glob, j0 = nyl * world, nyl * rank
        Lx, Ly = 512e3, 512e3 * world
    hx, hy = Lx / nx, Ly / ny_glob
    dyn = GpuDynamics(ctx, nx, ny_glob, hx, hy, NCOMP, ny_local=nyl, j0=j0, mevp=mp)
    F = None
    if strong:
        # the strip is generated and uploaded 1024 rows at a time (34 GB of H, A and forcing at
        # N = 1); U, V and the stress start at zero like the device buffers
        for c0 in range(0, nyl, 1024):
            n = min(1024, nyl - c0)
            Fc, _, _ = S.box_case(nx, ny_glob, NCOMP, L=Lx, Ly=Ly, rows=(j0 + c0, n), separable=True,
                                  state=False)
            for k, v in Fc.items():
                dyn.upload_rows(k, c0, v)
        del Fc
    else:
        F, _, _ = S.box_case(nx, ny_glob, NCOMP, L=Lx, Ly=Ly, rows=(j0, nyl), separable=True)
        dyn.upload_all(F)
    cells_local = nx * nyl
    cells_total = nx * ny_glob

    # ---- resident-state timing: W warm-up steps, then exactly K steps
    for _ in range(warmup):
        dyn.step(DT, NSUB)
    ctx.synchronize()
    env.barrier()
    sampler = ClockSampler(env.local, rank == 0).start()
    l0 = ctx.launches
    ctx.profile_begin()
    ctx.timer_start()
    for _ in range(steps):
        dyn.step(DT, NSUB)
    ms = ctx.timer_stop()
    prof = ctx.profile_end()
    launches = ctx.launches - l0
    env.barrier()
    ms_max = env.max(ms)
    extra = 0
    if ms_max < 1000.0:
        # nvidia-smi needs a few hundred ms for its first samples: keep the same step running
        # (untimed; the count is the same on every rank) so that the clocks are sampled under
        # this load
        extra = int(np.ceil((1000.0 - ms_max) / (ms_max / steps)))
        for _ in range(extra):
            dyn.step(DT, NSUB)
        ctx.synchronize()
        env.barrier()
    clocks = sampler.stop(f"the timed steps plus {extra} identical steps right after them" if extra else None)
    value = cells_total * steps / (ms_max * 1e-3)
    umax = float(np.abs(dyn.download_rows("U", 0, min(nyl, 1024))).max())
    if not np.isfinite(umax):
        raise SystemExit("bench.py: non-finite velocity after the timed steps")
    halo_mode = ("peer stores (NVLink, CUDA IPC)" if ctx.halo_peer else "NCCL send/recv") if world > 1 else "none"

    # ---- e2e: one C-ABI call per step with HOST (pinned) buffers (nsdg_dyn_step_host_async):
    # this step's forcing goes up, the fields other model components read (H, A, U, V) come down,
    # copies overlap the compute; the model state itself stays resident between steps (as
    # DevStep's state stays in memory) and so does the stress, which is internal state of the
    # rheology.  e2e.full_state also downloads the stress every step.
    e2e = None
    if with_e2e:
        names = ("H", "A", "U", "V", "S11", "S12", "S22", "UATM", "VATM", "UOCN", "VOCN")
        forcing = ("UATM", "VATM", "UOCN", "VOCN")
        outs_full = ("H", "A", "U", "V", "S11", "S12", "S22")
        outs_fields = ("H", "A", "U", "V")
        host = {k: env.pinned(np.ascontiguousarray(F[k])) for k in names}
        # a second set of output buffers: a consumer reads step k's set while step k+1 fills the other
        host2 = {k: torch.empty(host[k].shape, dtype=torch.float64).pin_memory() for k in outs_full}

        def e2e_run(in_names, nsteps, deferred, outs):
            ins = {k: host[k].data_ptr() for k in in_names}
            outp = [{k: host[k].data_ptr() for k in outs}, {k: host2[k].data_ptr() for k in outs}]
            dyn.step_host(DT, NSUB, ins, outp[0])  # untimed first call
            env.barrier()
            t0 = time.perf_counter()
            for s in range(nsteps):
                dyn.step_host(DT, NSUB, ins, outp[s & 1], deferred=deferred)
            if deferred:
                dyn.host_wait()  # the last step's fields are on the host when the clock stops
            env.barrier()
            return cells_total * nsteps / env.max(time.perf_counter() - t0)

        e2e_steps = max(1, args.e2e_steps)  # the last step's download drains inside the timed region
        h2d = sum(host[k].numel() * 8 for k in forcing)
        d2h_fields = sum(host[k].numel() * 8 for k in outs_fields)
        d2h_full = sum(host[k].numel() * 8 for k in outs_full)
        h2d_full = sum(host[k].numel() * 8 for k in names)
        # the headline leg twice, the better one counts: its timed region is ~0.5 s of wall clock
        # and one host-side stall of the box (seen: 0.3 s, in any one of the legs) would halve it
        v_fields_runs = [e2e_run(forcing, e2e_steps, True, outs_fields) for _ in range(2)]
        v_fields = max(v_fields_runs)
        v_full = e2e_run(forcing, e2e_steps, True, outs_full)
        v_block = e2e_run(forcing, min(e2e_steps, 3), False, outs_fields)
        v_round = e2e_run(names, 1, False, outs_full)
        link = host_link_probe(env)
        # the copies of one step run beside its compute: the roof of this call is the slower of
        # the device step and the two copy directions at the link rate measured just above
        t_link = max(h2d / (link["h2d_gbs_per_gpu"] * 1e9), d2h_fields / (link["d2h_gbs_per_gpu"] * 1e9))
        roof_val = cells_total / max(ms_max * 1e-3 / steps, t_link)
        e2e = {"value": v_fields, "unit": "cell-updates/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h_fields, "steps": e2e_steps, "host_numa_binding": env.numa,
               "repetitions": v_fields_runs, "repetition_rule": "two repetitions of the timed steps, the better one is `value`",
               "what": "nsdg_dyn_step_host_async per step + nsdg_dyn_host_wait inside the timed "
                       "region: every step uploads its forcing (wind, ocean current) from pinned "
                       "host memory and downloads the fields H, A, U, V; the downloads of step k "
                       "overlap the compute of step k+1; the prognostic state and the stress "
                       "(internal state of the rheology) stay resident between steps",
               "host_link_roof_gbs": link,
               "roof": {"value": roof_val, "frac": v_fields / roof_val,
                        "what": "cells / max(device step, h2d bytes / measured h2d rate, d2h bytes "
                                "/ measured d2h rate), all ranks copying at once"},
               "full_state": {"value": v_full, "d2h_bytes_per_step": d2h_full,
                              "what": "same call, the stress (9 planes) comes down every step too"},
               "blocking": {"value": v_block,
                            "what": "nsdg_dyn_step_host: same copies as e2e.value, each call "
                                    "returns with its outputs on the host"},
               "full_roundtrip": {"value": v_round, "h2d_bytes_per_step": h2d_full,
                                  "d2h_bytes_per_step": d2h_full,
                                  "what": "blocking call, whole state uploaded and downloaded every step"}}
        del host, host2

    line = None
    if rank == 0:
        peak, peak_src = peaks()
        kernels = kernel_table(prof, ms, cells_local, steps, peak)
        # dominant kernel of the step by measured share
        kname, (ktot, kcnt) = max(prof.items(), key=lambda kv: kv[1][0])
        units = SUBITS_PER_LAUNCH.get(kname, 1)
        krec = kernels[kname]
        kavg_ms = ktot / kcnt
        split = "stress_kernel" in prof
        paired = units > 1
        BYTES_STEP = BYTES_STEP_SPLIT if split else (BYTES_STEP_PAIR if paired else BYTES_STEP_FUSED)
        tm = prof.get("transport_march_kernel")
        if tm and tm[1] == 2 * steps:  # two launches per step: stages 1 + 2 share one pass
            BYTES_STEP += BYTES_TRANSPORT_STEP_FUSED - BYTES_TRANSPORT_STEP
        step_gbs = BYTES_STEP * cells_local * steps / (ms * 1e-3) / 1e9
        single_launch = world == 1 or ctx.halo_peer
        traffic = NCU_TRAFFIC_4096SQ.get(kname) if (cells_local == 4096 * 4096 and single_launch) else None
        roof = {"bound": "hbm", "kernel": kname,
                "achieved": krec.get("gbs"), "peak": peak, "unit": "GB/s", "frac": krec.get("frac"),
                "peak_source": peak_src,
                "definition": "compulsory DRAM bytes of the kernel's launches in the timed region "
                              "(256 B per cell per pass over HBM; a pair launch advances two "
                              "sub-iterations in one pass) / their CUDA-event time / measured HBM peak",
                "traffic": traffic[0] if traffic else None,
                "traffic_source": (traffic[1] + " (ncu --set full, one launch over 4096^2 cells; not "
                                   "re-measured in this run)") if traffic else None,
                "algorithmic_bytes_per_launch": krec.get("dram_bytes", 0) / max(kcnt, 1),
                "units_per_launch": f"{units} sub-iteration(s) x {cells_local} cells, one pass of "
                                    f"{BYTES_FUSED} B per cell" if "mevp" in kname else None,
                "avg_launch_ms": kavg_ms, "launches": kcnt, "share_of_step": ktot / ms,
                "speedup_vs_one_pass_per_subiteration": units,
                "by_subiteration_counting_rule": {
                    "achieved": krec.get("gbs", 0) * units, "frac": (krec.get("frac") or 0) * units,
                    "what": "SURVEY.md 8(d) counts 256 B per cell per SUB-ITERATION; with temporal "
                            "blocking that figure exceeds the HBM peak and is not a DRAM fraction"},
                "whole_step": {"bytes_per_cell": BYTES_STEP, "achieved": step_gbs, "frac": step_gbs / peak,
                               "what": "compulsory DRAM bytes of the whole step (transport + prepare + "
                                       "one pass per launch of the sub-cycle) / step time / peak"},
                "kernels": kernels}
        if world > 1 and not ctx.halo_peer:
            roof["note"] = ("NCCL halo path: the boundary-band launches run on the halo stream outside "
                            "the per-kernel event pairs, so the kernel time above misses them")
        cpu = None
        if with_cpu and world == 1 and not args.no_cpu:
            from oracle import cpu_oracle  # noqa: F401  (fails loudly if the checker is missing)
            ny_s, Fc, hxc, hyc = cpu_dyn_sample(nx, ny_glob, args.cpu_budget, 1, mp)
            secs = cpu_dyn_step(Fc, hxc, hyc, mp, 1)
            cpu = {"value": nx * ny_s / secs, "unit": "cell-updates/s", "cores": os.cpu_count() or 1,
                   "kind": "port",
                   "sample": f"rows [0, {ny_s}) of the same {nx}x{ny_glob} box test "
                             f"({'the whole grid' if ny_s == ny_glob else 'a strip'}), 1 full step "
                             f"(transport + {NSUB} sub-iterations), oracle/ C + OpenMP",
                   "note": "parity unpinned against draenog/nextsimdg (no dynamics in the snapshot)"}
            del Fc
        cfg = dict(workload_config(args, world))
        if strong:
            cfg = {"workload": f"synthetic 16384x16384 dG2 transport + mEVP (100 substeps), {world} "
                               f"strip(s) over y (BASELINE configs[4])", "nx": 16384, "ny": 16384,
                   "substeps": NSUB, "dt_s": DT, "dg": "dG2", "rk": "SSP-RK3", "bc": "box",
                   "l2": "working set >> 126 MB L2", "parity": PARITY_NOTE}
        cfg["halo"] = halo_mode
        cfg["band_rows"] = int(ctx.lib.nsdg_mevp_band_rows(nyl, nx, 148, 0))
        line = {"metric": METRIC, "value": value,
                "unit": "cell-updates/s", "n_gpus": world, "steps": steps,
                "warmup": warmup, "ms_per_step": ms_max / steps,
                "higher_is_better": True, "scaling": "strong" if strong else "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": cfg, "clocks": clocks, "e2e": e2e,
                "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu,
                "max_abs_u": umax,
                "halo_bytes_sent_per_rank": ctx.comm_bytes_sent if world > 1 else 0}
    dyn.close()
    ctx.close()
    del F
    return line


def host_link_probe(env, nbytes=1 << 30, reps=3):
    """Pinned-memory copies at the size class of the e2e buffers, every rank at once (plain
    cudaMemcpyAsync through torch): host-to-device alone, device-to-host alone, both directions
    together.  Returns GB/s per GPU (max-over-ranks time) and aggregate."""
    torch = env.torch
    h_in = torch.empty(nbytes // 8, dtype=torch.float64).pin_memory()
    h_out = torch.empty(nbytes // 8, dtype=torch.float64).pin_memory()
    d_in = torch.empty(nbytes // 8, dtype=torch.float64, device="cuda")
    d_out = torch.zeros(nbytes // 8, dtype=torch.float64, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(up, down):
        best = 1e30
        for _ in range(reps):
            env.barrier()
            t0 = time.perf_counter()
            if up:
                with torch.cuda.stream(s1):
                    d_in.copy_(h_in, non_blocking=True)
            if down:
                with torch.cuda.stream(s2):
                    h_out.copy_(d_out, non_blocking=True)
            torch.cuda.synchronize()
            best = min(best, env.max(time.perf_counter() - t0))
        return nbytes / best / 1e9

    h2d, d2h = run(True, False), run(False, True)
    both = run(True, True)
    return {"h2d_gbs_per_gpu": h2d, "d2h_gbs_per_gpu": d2h, "bidirectional_gbs_per_gpu_each_way": both,
            "aggregate_h2d_gbs": h2d * env.world, "aggregate_d2h_gbs": d2h * env.world,
            "n_gpus": env.world, "bytes": nbytes,
            "what": "1 GiB pinned cudaMemcpyAsync per direction on every rank at once, best of 3, "
                    "max over ranks"}


def measure_parity(args, env):
    """N > 1, before any timing: one overlap strip case of tests/mgpu_worker.py against the CPU
    oracle on rank 0, under the peer-store halo and the NCCL halo; the variants must also agree
    bit for bit."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import mgpu_worker as W
    from nextsimdg_b200._abi import BC_BOX
    ctx = env.context()
    variants = [v for v in W.VARIANTS if v[0] in ("peer", "peer-rows16", "peer-rows3", "nccl-overlap-rows16", "nccl-serial")]
    cases = [(48, 49 * env.world, BC_BOX, 0, 6), (300, 33 * env.world + 1, BC_BOX, 0, 8)]
    ok, worst, records = W.run_cases(ctx, env.rank, env.world, cases, variants, log=log)
    ctx.close()
    if env.rank != 0:  # rank 0 holds the oracle's answer
        return None
    return {"ok": bool(ok), "max_rel": worst, "tolerance": 1e-11,
            "oracle": "oracle/dg_transport.c + oracle/mevp.c on the undivided grid (rank 0)",
            "what": f"{env.world} strips, 2 steps of dG2 transport + mEVP, box walls; every variant "
                    "against the oracle and bitwise against the first variant",
            "cases": records}


# ---------------------------------------------------------------------------------------------
# BASELINE configs[3], the per-cell column thermodynamics (the part of the path that EXISTS in
# the reference: parity pinned, CPU arm = the reference's own sources)
THERMO_BYTES = 160  # 15 input + 5 output doubles per cell-update
THERMO_DT = 600.0


def thermo_config(args, n):
    return {"workload": f"synthetic {args.nx}x{args.ny} per-cell column thermodynamics "
                        f"(NextsimPhysics/ThermoIce0), {n} GPU(s), independent cells, contiguous "
                        "split, no collective",
            "cells_per_gpu": args.nx * args.ny, "dt_s": THERMO_DT,
            "l2": "working set 2.7 GB per GPU >> 126 MB L2",
            "parity": "pinned: reference sources compiled in place (oracle/_ref/libnsdg_ref.so)"}


def thermo_reference_rate(ncells, nsteps, nthreads):
    """The reference's own column physics (DevStep::iterate body over std::vector<ElementData>,
    oracle/ref_harness.cpp) on `nthreads` host threads; returns cell-updates/s."""
    from nextsimdg_b200 import synthetic as S
    from oracle import refthermo as R
    R.configure(R.ini())
    st, fo = S.thermo_inputs(ncells)
    g = R.RefGrid(ncells, 1)
    g.set_prognostic(st["hice"], st["cice"], st["hsnow"], st["sst"], st["sss"], st["tice0"])
    g.set_forcing(*[fo[k] for k in R.FORC])
    g.step(THERMO_DT, 1, nthreads)  # warm-up (first-touch, statics)
    t0 = time.perf_counter()
    g.step(THERMO_DT, nsteps, nthreads)
    secs = time.perf_counter() - t0
    g.close()
    return ncells * nsteps / secs, secs


def thermo_cpu_baseline(nsteps=3, ncells=1 << 21):
    from oracle import refthermo as R
    cores = os.cpu_count() or 1
    if R.available():
        rate, _ = thermo_reference_rate(ncells, nsteps, cores)
        one, _ = thermo_reference_rate(ncells, 1, 1)
        return {"value": rate, "unit": "cell-updates/s", "cores": cores, "kind": "reference",
                "sample": f"{ncells} cells of the same distribution, {nsteps} sweeps, the reference's own "
                          "column physics (physics/src/modules/NextsimPhysics.cpp:116-131, "
                          "ThermoIce0.cpp:34-133, driven like core/src/DevStep.cpp:14-23) compiled in "
                          "place into oracle/_ref/libnsdg_ref.so",
                "single_thread": one}
    # the compiled reference did not travel: the C restatement of the same arithmetic
    from nextsimdg_b200 import synthetic as S
    from nextsimdg_b200._abi import ThermoParams
    from oracle import cpu_oracle as O
    st, fo = S.thermo_inputs(ncells)
    t0 = time.perf_counter()
    O.thermo_step(ThermoParams.defaults(), THERMO_DT, nsteps, st, fo, nthreads=cores)
    secs = time.perf_counter() - t0
    return {"value": ncells * nsteps / secs, "unit": "cell-updates/s", "cores": cores, "kind": "port",
            "sample": f"{ncells} cells, {nsteps} sweeps, oracle/thermo_port.c (OpenMP)"}


def run_thermo_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cpu = thermo_cpu_baseline(args.steps)
    n = 1 << 21
    line = {"impl": "reference", "metric": "cell-updates/sec (column thermodynamics)", "value": cpu["value"],
            "unit": "cell-updates/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * n / cpu["value"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": thermo_config(args, args.gpus), "cpu_baseline": cpu,
            "e2e": {"value": cpu["value"], "unit": "cell-updates/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}
    emit(line)


def measure_thermo(args, env, steps=None, warmup=None):
    from nextsimdg_b200 import synthetic as S
    from nextsimdg_b200._abi import ThermoParams
    from nextsimdg_b200.gpu import Context, GpuStep
    rank, world = env.rank, env.world
    steps = max(args.steps, 20) if steps is None else steps
    warmup = max(args.warmup, 3) if warmup is None else warmup
    ctx = Context(env.local)  # independent columns: no communicator
    n = args.nx * args.ny  # cells of this rank: independent columns, no collective
    st, fo = S.thermo_inputs(n, seed=S.THERMO_SEED + rank)
    prog = dict(st)
    prog["tice"] = prog.pop("tice0")
    step = GpuStep(ctx, n, 1, ThermoParams.defaults())
    step.setInitialData(prog, fo)
    for _ in range(warmup):
        step.iterate(THERMO_DT, 1)
    ctx.synchronize()
    env.barrier()
    sampler = ClockSampler(env.local, rank == 0).start()
    l0 = ctx.launches
    ctx.profile_begin()
    ctx.timer_start()
    for _ in range(steps):
        step.iterate(THERMO_DT, 1)
    ms = ctx.timer_stop()
    prof = ctx.profile_end()
    launches = ctx.launches - l0
    env.barrier()
    # a sweep takes 0.4 ms: keep the same kernel running until nvidia-smi has samples
    extra = 2500
    step.iterate(THERMO_DT, extra)
    ctx.synchronize()
    env.barrier()
    clocks = sampler.stop(f"the timed sweeps plus {extra} more right after them")
    ms_max = env.max(ms)
    value = n * world * steps / (ms_max * 1e-3)
    got = step.get_prognostic()
    if not all(np.isfinite(v).all() for v in got.values()):
        raise SystemExit("bench.py: non-finite thermodynamic state after the timed steps")

    # ---- e2e: nsdg_thermo_run_host with pinned host buffers: 14 arrays up, one sweep, 4 down
    e2e = None
    if not args.no_e2e:
        keep = {k: env.pinned(v) for k, v in list(prog.items()) + list(fo.items())}
        hp = {k: keep[k].numpy() for k in prog}
        hf = {k: keep[k].numpy() for k in fo}
        esteps = 5
        step.run_host(THERMO_DT, 1, hp, hf)
        env.barrier()
        t0 = time.perf_counter()
        for _ in range(esteps):
            step.run_host(THERMO_DT, 1, hp, hf)
        env.barrier()
        e2e = {"value": n * world * esteps / env.max(time.perf_counter() - t0), "unit": "cell-updates/s",
               "h2d_bytes_per_step": 14 * n * 8, "d2h_bytes_per_step": 4 * n * 8, "steps": esteps,
               "what": "nsdg_thermo_run_host per step: prognostic state + forcing up from pinned host "
                       "memory, one sweep, the changed prognostic fields down (chunked three-stage "
                       "pipeline); bound by the 14-array upload over the host link"}
        del keep
    line = None
    if rank == 0:
        peak, peak_src = peaks()
        ktot, kcnt = prof["thermo_kernel"]
        kavg = ktot / kcnt
        achieved = THERMO_BYTES * n / (kavg * 1e-3) / 1e9
        cpu = thermo_cpu_baseline() if (world == 1 and not args.no_cpu) else None
        line = {"metric": "cell-updates/sec (column thermodynamics)", "value": value,
                "unit": "cell-updates/s", "n_gpus": world, "steps": steps, "warmup": warmup,
                "ms_per_step": ms_max / steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": thermo_config(args, world), "clocks": clocks, "e2e": e2e,
                "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "kernel": "thermo_kernel", "achieved": achieved,
                             "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "peak_source": peak_src, "traffic": 2.656e9 if n == 4096 * 4096 else None,
                             "traffic_source": "profiles/r1w_ncu_full_summary.txt (ncu --set full; not "
                                               "re-measured in this run)" if n == 4096 * 4096 else None,
                             "algorithmic_bytes_per_launch": THERMO_BYTES * n,
                             "units_per_launch": f"{n} cells x {THERMO_BYTES} B",
                             "avg_launch_ms": kavg, "launches": kcnt, "share_of_step": ktot / ms},
                "cpu_baseline": cpu}
        if cpu:
            line["vs_cpu_reference"] = {"resident": value / cpu["value"],
                                        "e2e": (e2e["value"] / cpu["value"]) if e2e else None,
                                        "single_thread_resident": value / cpu["single_thread"]
                                        if cpu.get("single_thread") else None}
    step.close()
    ctx.close()
    return line


# ---------------------------------------------------------------------------------------------
# BASELINE configs[1], 1024^2 dG1 rotating-cone advection (periodic, Heun RK2)
CONE_N = 1024
CONE_BYTES = (5 * 3 + 4) * 8  # Heun RK2, one dG1 field: 152 B per cell and step


def cone_setup():
    from nextsimdg_b200 import synthetic as S
    phi, u, v, hx, hy, f0 = S.cone_case(CONE_N, CONE_N, 3)
    vmax = float(np.sqrt(u * u + v * v).max())
    dt = (0.1 / 3.0) * hx / vmax  # CFL = 0.1 / (2k + 1), k = 1
    return phi, u, v, hx, hy, dt


def cone_config(n):
    return {"workload": f"synthetic {CONE_N}x{CONE_N} dG1 rotating-cone advection, periodic BCs, "
                        f"Heun RK2, {n} GPU(s) (independent replicas for N > 1)",
            "cells_per_gpu": CONE_N * CONE_N, "cfl": "0.1/3",
            "l2": "working set 67 MB (field, RK stage, velocity) FITS the 126 MB L2 at this config's own "
                  "size and is not flushed: consecutive steps of a real run find it there too; the HBM "
                  "fraction below is therefore not a DRAM utilisation",
            "parity": "unpinned against draenog/nextsimdg (no transport in the reference snapshot)"}


def run_cone_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from oracle import cpu_oracle as O
    phi, u, v, hx, hy, dt = cone_setup()
    O.transport_step(phi, u, v, hx, hy, dt)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        phi = O.transport_step(phi, u, v, hx, hy, dt)
    secs = time.perf_counter() - t0
    value = CONE_N * CONE_N * args.steps / secs
    cores = os.cpu_count() or 1
    line = {"impl": "reference", "metric": "cell-updates/sec (dG1 transport step)", "value": value,
            "unit": "cell-updates/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cone_config(args.gpus),
            "cpu_baseline": {"value": value, "unit": "cell-updates/s", "cores": cores, "kind": "port",
                             "sample": f"the full {CONE_N}^2 grid, {args.steps} steps, "
                                       "oracle/dg_transport.c (OpenMP)"},
            "e2e": {"value": value, "unit": "cell-updates/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}
    emit(line)


def measure_cone(args, env):
    from nextsimdg_b200.gpu import Context, GpuDynamics
    rank, world = env.rank, env.world
    ctx = Context(env.local)  # replicas: no communicator
    phi, u, v, hx, hy, dt = cone_setup()
    n = CONE_N * CONE_N
    dyn = GpuDynamics(ctx, CONE_N, CONE_N, hx, hy, 3)
    dyn.upload_all(dict(H=phi, U=u, V=v))
    m0 = dyn.mass("H")
    steps = max(args.steps, 20000)  # a step takes ~30 us: time enough of them for the clocks sampler
    warm = max(args.warmup, 3)
    for _ in range(warm):
        dyn.transport_step(dt, "H")
    ctx.synchronize()
    env.barrier()
    sampler = ClockSampler(env.local, rank == 0).start()
    l0 = ctx.launches
    ctx.timer_start()
    t_host = time.perf_counter()
    for _ in range(steps):
        dyn.transport_step(dt, "H")
    t_host = time.perf_counter() - t_host  # the host's share: enqueueing only (the queue never fills)
    ms = ctx.timer_stop()
    launches = ctx.launches - l0
    env.barrier()
    if ms < 900.0:
        for _ in range(int(steps * (900.0 - ms) / ms) + 1):
            dyn.transport_step(dt, "H")
        ctx.synchronize()
    clocks = sampler.stop("the timed steps (plus identical ones up to 0.9 s)")
    # per-launch times from a short profiled run (event pairs around every launch, no graph replay)
    ctx.profile_begin()
    for _ in range(50):
        dyn.transport_step(dt, "H")
    prof = ctx.profile_end()
    ms_max = env.max(ms)
    value = n * world * steps / (ms_max * 1e-3)
    drift = abs(dyn.mass("H") - m0) / m0
    got = dyn.download("H")
    if not np.isfinite(got).all():
        raise SystemExit("bench.py: non-finite tracer after the timed steps")
    e2e = None
    if not args.no_e2e:
        # per step: the field goes up from the host, one step, the field comes down
        esteps = 50
        host = np.ascontiguousarray(got)
        env.barrier()
        t0 = time.perf_counter()
        for _ in range(esteps):
            dyn.upload("H", host)
            dyn.transport_step(dt, "H")
            host = dyn.download("H")
        env.barrier()
        e2e = {"value": n * world * esteps / env.max(time.perf_counter() - t0), "unit": "cell-updates/s",
               "h2d_bytes_per_step": 3 * n * 8, "d2h_bytes_per_step": 3 * n * 8,
               "what": "nsdg_dyn_upload + nsdg_transport_step + nsdg_dyn_download per step"}
    line = None
    if rank == 0:
        peak, peak_src = peaks()
        kname = max(prof.items(), key=lambda kv: kv[1][0])[0]
        ktot, kcnt = prof[kname]
        kavg = ktot / kcnt
        # default: both Heun stages in ONE launch (transport_heun_march_kernel): phi in, u, v, phi out
        # = 8 doubles per cell are all it has to move; with transport_fused = -1 two stage launches
        fused = kcnt <= 50
        kbytes = 8 * 8 * n if fused else CONE_BYTES * n / 2.0
        achieved = kbytes / (kavg * 1e-3) / 1e9
        cpu = None
        if world == 1 and not args.no_cpu:
            from oracle import cpu_oracle as O
            p2 = phi
            t0 = time.perf_counter()
            for _ in range(20):
                p2 = O.transport_step(p2, u, v, hx, hy, dt)
            secs = time.perf_counter() - t0
            cpu = {"value": n * 20 / secs, "unit": "cell-updates/s", "cores": os.cpu_count() or 1,
                   "kind": "port", "sample": f"the full {CONE_N}^2 grid, 20 steps, oracle/dg_transport.c"}
        line = {"metric": "cell-updates/sec (dG1 transport step)", "value": value,
                "unit": "cell-updates/s", "n_gpus": world, "steps": steps, "warmup": warm,
                "ms_per_step": ms_max / steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cone_config(world),
                "clocks": clocks, "e2e": e2e,
                "gpu_launches": int(launches), "mass_drift_rel": drift,
                "launch_mode": "one plain launch per step (both Heun stages fused)" if fused and dyn.get_tuning().graph <= 0
                else ("CUDA-graph replay of the step's " + ("one fused two-stage launch" if fused else "two stage launches"))
                if dyn.get_tuning().graph >= 0 else "plain launches",
                "host_enqueue_us_per_step": t_host / steps * 1e6,
                "roofline": {"bound": "hbm", "kernel": kname, "achieved": achieved,
                             "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "peak_source": peak_src, "traffic": None,
                             "algorithmic_bytes_per_launch": kbytes,
                             "units_per_launch": (f"{n} cells x 64 B (phi in, u, v, phi out: both Heun stages in one launch; "
                                                  f"SURVEY.md 8d counts {CONE_BYTES} B per cell and step for stage-by-stage passes)")
                             if fused else f"{n} cells x {CONE_BYTES}/2 B (one of two RK stages)",
                             "avg_launch_ms": kavg, "launches": kcnt,
                             "whole_step": {"bytes_per_cell": CONE_BYTES,
                                            "achieved": CONE_BYTES * n * steps / (ms * 1e-3) / 1e9,
                                            "frac": CONE_BYTES * n * steps / (ms * 1e-3) / 1e9 / peak}},
                "cpu_baseline": cpu}
    dyn.close()
    ctx.close()
    return line


def use_all_host_threads():
    """The CPU arms use every host core; torchrun presets OMP_NUM_THREADS=1 for its workers, which
    would silently turn the OpenMP oracle into a single-thread run."""
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)


def trim(rec):
    """a sub-record keeps what it measured, not the contract's constant keys"""
    if rec is None or "error" in rec:
        return rec
    for k in ("higher_is_better", "vs_baseline", "dtype", "data"):
        rec.pop(k, None)
    return rec


def run_gpu(args):
    use_all_host_threads()  # the cpu_baseline legs (rank 0, N = 1) use every core
    env = Env(args)
    t_start = time.perf_counter()
    if args.workload == "thermo":
        line = measure_thermo(args, env, steps=args.steps, warmup=args.warmup)
    elif args.workload == "cone":
        line = measure_cone(args, env)
    elif args.workload == "16384sq":
        line = measure_north_star(args, env, strong=True, with_e2e=False, with_cpu=False)
    else:
        parity = measure_parity(args, env) if (env.world > 1 and not args.no_parity) else None
        log(f"parity done at {time.perf_counter() - t_start:.0f} s")
        line = measure_north_star(args, env)
        log(f"north star done at {time.perf_counter() - t_start:.0f} s")
        if not args.main_only:
            def sub(name, fn):
                """a sub-record must never cost the main line: a failure is reported in its place"""
                try:
                    rec = fn()
                except Exception as e:  # noqa: BLE001
                    rec = {"error": f"{type(e).__name__}: {e}"}
                    env.torch.cuda.empty_cache()
                log(f"{name} done at {time.perf_counter() - t_start:.0f} s")
                return rec
            th = sub("thermo", lambda: measure_thermo(args, env))
            cone = sub("cone", lambda: measure_cone(args, env)) if env.world == 1 else None
            strong = sub("strong", lambda: measure_north_star(args, env, strong=True, with_e2e=False,
                                                              with_cpu=False, steps=args.strong_steps, warmup=1))
            if env.rank == 0:
                line["configs"] = {"thermo_4096sq": trim(th), "cone_1024sq": trim(cone)}
                if strong and "error" not in strong:
                    strong = trim(strong)
                    strong["cells_per_gpu"] = 16384 * 16384 // env.world
                    strong["note"] = ("strong scaling: the same 16384^2 grid at every N; efficiency at N "
                                      "= value(N) / (N x value(1)) from the driver's per-N lines")
                line["strong_16384sq"] = strong
        if env.rank == 0 and parity is not None:
            line["parity"] = parity
    if env.rank == 0:
        line["bench_wall_s"] = time.perf_counter() - t_start
        emit(line)
    env.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto", "16384sq", "thermo", "cone"])
    ap.add_argument("--nx", type=int, default=4096)
    ap.add_argument("--ny", type=int, default=4096, help="rows per GPU")
    ap.add_argument("--e2e-steps", type=int, default=10)
    ap.add_argument("--strong-steps", type=int, default=3, help="timed steps of the 16384^2 sub-record")
    ap.add_argument("--cpu-sample", type=int, default=0,
                    help="rows of the CPU sample strip (0: the whole grid if it fits --cpu-budget)")
    ap.add_argument("--cpu-budget", type=float, default=None,
                    help="seconds of CPU work the CPU arm may spend on its timed steps "
                         "(default 240 for --impl reference, 25 for the cpu_baseline leg)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer legs")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the parity record")
    ap.add_argument("--main-only", action="store_true",
                    help="only the north-star workload (no thermo / cone / 16384^2 sub-records)")
    args = ap.parse_args()
    if args.cpu_budget is None:
        args.cpu_budget = 240.0 if args.impl == "reference" else 25.0
    if args.impl == "reference":
        use_all_host_threads()  # before the oracle library (and its OpenMP runtime) is loaded
        if args.workload == "thermo":
            run_thermo_reference(args)
        elif args.workload == "cone":
            run_cone_reference(args)
        else:
            run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
