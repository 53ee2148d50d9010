#!/usr/bin/env python
"""Benchmark of the north-star unit: one dG2 transport (H, A; SSP-RK3) + 100 mEVP sub-iterations
step on a synthetic box-test grid.  `python bench.py --gpus N --steps K --warmup W`.

metric   cell-updates/s = cells of the whole job x steps / max-over-ranks device time
N = 1    BASELINE.json configs[2]: 4096 x 4096 cells, state resident in HBM
N > 1    weak scaling: every rank owns a 4096 x 4096 strip of a 4096 x (4096 N) grid (ring of
         strips over y, one-row halo per sub-iteration via NCCL send/recv); `--workload 16384sq`
         runs BASELINE configs[4] (16384^2 split over N >= 2 ranks, strong scaling) instead
e2e      same step through the C ABI with HOST (pinned) buffers (nsdg_dyn_step_host_async): every
         step uploads its forcing and downloads the full state inside the timed region
--impl reference   the CPU arm: the oracle's restatement of the same step on the host cores
         (the reference snapshot has no dynamics code to run), bounded sample
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

NSUB = 100
DT = 120.0
NCOMP = 6
# algorithmic bytes per cell (counting rule of SURVEY.md 8d: every distinct
# array element one cell-update must read or write counts once, neighbour re-reads are free)
BYTES_STRESS = (9 + 2 + 4 + 9) * 8  # r S(9) u,v(2) P_g(4)  w S(9)
BYTES_VELOCITY = (7 + 2 + 2 + 4 + 2) * 8  # r S(7 of 9) u,v u_o,v_o node constants(4)  w u,v
BYTES_FUSED = (9 + 2 + 4 + 2 + 4 + 9 + 2) * 8  # one fused sub-iteration: 32 doubles
BYTES_TRANSPORT_STEP = (2 * 8 * NCOMP + 6) * 8  # H and A, SSP-RK3, shared velocity reads
BYTES_PREPARE = (12 + 4 + 8 + 4) * 8  # P_g from H,A; node constants from H0,A0,u_a,u,u_o
# the pair kernel runs TWO sub-iterations per launch: its units per launch are 2 x cells, its
# compulsory DRAM bytes per launch (temporal blocking) are those of ONE fused sub-iteration
SUBITS_PER_LAUNCH = {"mevp_pair_kernel": 2, "mevp_march_kernel": 2}
# dram__bytes_read.sum + dram__bytes_write.sum per launch at 4096^2 cells from the ncu --set full
# captures summarised in profiles/r1t_ncu_full_summary.txt (r1d_... for the fused kernel,
# r1z_march_ncu_summary.txt for the marching kernel)
NCU_TRAFFIC_4096SQ = {"mevp_pair_kernel": 4.2719e9, "mevp_march_kernel": 4.2980e9,
                      "mevp_fused_kernel": 4.5397e9,
                      "transport_stage_kernel": 3.7266e9}
BYTES_STEP_SPLIT = BYTES_TRANSPORT_STEP + BYTES_PREPARE + NSUB * (BYTES_STRESS + BYTES_VELOCITY)
BYTES_STEP_FUSED = BYTES_TRANSPORT_STEP + BYTES_PREPARE + NSUB * BYTES_FUSED


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

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
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
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm),
                "power_w": float(np.median(pw)) if pw else None, "power_limit_w": plim}


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


def cpu_step(F, hx, hy, mp, nsteps=1):
    """The oracle's restatement of the bench step on the host cores (OpenMP, all threads)."""
    from oracle import cpu_oracle as O
    st = [F[k] for k in ("H", "A", "U", "V", "S11", "S12", "S22")]
    t0 = time.perf_counter()
    for _ in range(nsteps):
        st = list(O.dyn_step(mp, *st, F["UATM"], F["VATM"], F["UOCN"], F["VOCN"], hx, hy, DT, NSUB))
    return time.perf_counter() - t0


def run_reference(args):
    """--impl reference: CPU arm.  DG transport + mEVP do not exist in the reference snapshot
    (CMakeLists.txt:43-46), so the arm is the oracle port of the same step, all host threads,
    on a bounded sample (a 768^2 window of the same box test) per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from nextsimdg_b200 import synthetic as S
    from nextsimdg_b200._abi import BC_BOX, MevpParams
    n = args.cpu_sample
    F, hx, hy = S.box_case(n, n, NCOMP)
    mp = MevpParams.defaults()
    mp.bc = BC_BOX
    cores = os.cpu_count() or 1
    for _ in range(args.warmup if args.warmup < 2 else 1):
        cpu_step(F, hx, hy, mp)
    secs = cpu_step(F, hx, hy, mp, args.steps)
    value = n * n * args.steps / secs
    sample = f"{n}x{n}-cell window of the box test, {args.steps} full steps (transport + {NSUB} sub-iterations)"
    line = {"impl": "reference", "metric": "cell-updates/sec (dG2 transport+mEVP step)",
            "value": value, "unit": "cell-updates/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": workload_config(args, args.gpus),
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
        name = "synthetic 4096x4096 dG2 transport + mEVP rheology subcycle (100 substeps), 1 GPU"
        nx, ny = args.nx, args.ny
    else:
        name = (f"synthetic {args.nx}x{args.ny * n} dG2 transport + mEVP (100 substeps): "
                f"{n} strips of {args.nx}x{args.ny}, ring halo")
        nx, ny = args.nx, args.ny * n
    return {"workload": name, "nx": nx, "ny": ny, "substeps": NSUB, "dt_s": DT, "dg": "dG2",
            "rk": "SSP-RK3", "bc": "box", "l2": "working set >> 126 MB L2 (>= 10 GB per GPU)",
            "parity": "unpinned against draenog/nextsimdg (no dynamics in the reference snapshot)"}


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from nextsimdg_b200 import strips, synthetic as S
    from nextsimdg_b200._abi import BC_BOX, MevpParams
    from nextsimdg_b200.gpu import Context, GpuDynamics

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this framework has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    numa = bind_to_gpu_numa_node(local) if world > 1 else "single rank: not bound"
    ctx = Context(local)
    # ---- grid of this rank
    if args.workload == "16384sq":
        nx, ny_glob = 16384, 16384
        j0, nyl = strips.strip_rows(ny_glob, world, rank)
        Lx = Ly = 512e3
    else:
        nx, nyl = args.nx, args.ny
        ny_glob, j0 = nyl * world, nyl * rank
        Lx, Ly = 512e3, 512e3 * world
    if world > 1:
        uid = [Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(rank, world, uid[0])
    F, hx, hy = S.box_case(nx, ny_glob, NCOMP, L=Lx, Ly=Ly, rows=(j0, nyl))
    mp = MevpParams.defaults()
    mp.bc = BC_BOX
    dyn = GpuDynamics(ctx, nx, ny_glob, hx, hy, NCOMP, ny_local=nyl, j0=j0, mevp=mp)
    dyn.upload_all(F)
    cells_local = nx * nyl
    cells_total = nx * ny_glob

    # ---- resident-state timing: W warm-up steps, then exactly K steps
    for _ in range(args.warmup):
        dyn.step(DT, NSUB)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ctx.launches
    ctx.profile_begin()
    ctx.timer_start()
    for _ in range(args.steps):
        dyn.step(DT, NSUB)
    ms = ctx.timer_stop()
    prof = ctx.profile_end()
    launches = ctx.launches - l0
    barrier()
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    extra = 0
    if ms_max < 1000.0:
        # nvidia-smi needs a few hundred ms for its first samples: keep the same step running
        # (untimed; the count is the same on every rank) so that the clocks are sampled under
        # this load
        extra = int(np.ceil((1000.0 - ms_max) / (ms_max / args.steps)))
        for _ in range(extra):
            dyn.step(DT, NSUB)
        ctx.synchronize()
        barrier()
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None and extra:
        clocks["sampled_over"] = f"the timed steps plus {extra} identical steps right after them"
    value = cells_total * args.steps / (ms_max * 1e-3)
    umax = float(np.abs(dyn.download("U")).max())
    if not np.isfinite(umax):
        raise SystemExit("bench.py: non-finite velocity after the timed steps")

    # ---- e2e: one C-ABI call per step with HOST (pinned) buffers, nsdg_dyn_step_host: this
    # step's forcing goes up, the full prognostic state comes down, copies overlap the compute;
    # the model state itself stays resident between steps (as DevStep's state stays in memory)
    e2e_value = e2e_blocking = e2e_full = e2e_fields = None
    h2d = d2h = h2d_full = d2h_fields = e2e_steps = 0
    if not args.no_e2e:
        names = ("H", "A", "U", "V", "S11", "S12", "S22", "UATM", "VATM", "UOCN", "VOCN")
        forcing = ("UATM", "VATM", "UOCN", "VOCN")
        outs = ("H", "A", "U", "V", "S11", "S12", "S22")
        host = {}
        for k in names:
            a = np.ascontiguousarray(F[k])
            tpin = torch.empty(a.shape, dtype=torch.float64).pin_memory()
            tpin.numpy()[...] = a
            host[k] = tpin

        # a second set of output buffers: a consumer reads step k's set while step k+1 fills the other
        host2 = {k: torch.empty(host[k].shape, dtype=torch.float64).pin_memory() for k in outs}

        def e2e_run(in_names, nsteps, deferred, outs=outs):
            ins = {k: host[k].data_ptr() for k in in_names}
            outp = [{k: host[k].data_ptr() for k in outs}, {k: host2[k].data_ptr() for k in outs}]
            dyn.step_host(DT, NSUB, ins, outp[0])  # untimed first call
            barrier()
            t0 = time.perf_counter()
            for s in range(nsteps):
                dyn.step_host(DT, NSUB, ins, outp[s & 1], deferred=deferred)
            if deferred:
                dyn.host_wait()  # the last step's state is on the host when the clock stops
            barrier()
            te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
            if world > 1:
                dist.all_reduce(te, op=dist.ReduceOp.MAX)
            return cells_total * nsteps / float(te.item())

        e2e_steps = max(1, args.e2e_steps)  # the last step's download drains inside the timed region
        h2d = sum(host[k].numel() * 8 for k in forcing)
        d2h = sum(host[k].numel() * 8 for k in outs)
        e2e_value = e2e_run(forcing, e2e_steps, True)
        # the same with the stress left on the device (it is internal state of the rheology; the
        # fields other model components read are H, A, U, V)
        outs_fields = ("H", "A", "U", "V")
        d2h_fields = sum(host[k].numel() * 8 for k in outs_fields)
        e2e_fields = e2e_run(forcing, e2e_steps, True, outs_fields)
        e2e_blocking = e2e_run(forcing, min(e2e_steps, 3), False)
        # for transparency: the blocking call with the WHOLE state uploaded every step as well
        h2d_full = sum(host[k].numel() * 8 for k in names)
        e2e_full = e2e_run(names, 1, False)

    if rank == 0:
        peak, peak_src = peaks()
        # dominant kernel of the step by measured share
        dom = max(prof.items(), key=lambda kv: kv[1][0])
        per_kernel_bytes = {"stress_kernel": BYTES_STRESS, "velocity_kernel": BYTES_VELOCITY,
                            "mevp_fused_kernel": BYTES_FUSED, "mevp_pair_kernel": BYTES_FUSED,
                            "mevp_march_kernel": BYTES_FUSED,
                            "transport_stage_kernel": BYTES_TRANSPORT_STEP / 3.0}
        kname, (ktot, kcnt) = dom
        units = SUBITS_PER_LAUNCH.get(kname, 1)
        kbytes = per_kernel_bytes.get(kname, 0) * units * cells_local
        kavg_ms = ktot / kcnt
        achieved = kbytes / (kavg_ms * 1e-3) / 1e9
        split = "stress_kernel" in prof
        BYTES_STEP = BYTES_STEP_SPLIT if split else BYTES_STEP_FUSED
        step_gbs = BYTES_STEP * cells_local * args.steps / (ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                "traffic": args.traffic if args.traffic is not None else (
                    NCU_TRAFFIC_4096SQ.get(kname) if cells_local == 4096 * 4096 else None),
                "algorithmic_bytes_per_launch": kbytes,
                "units_per_launch": f"{units} sub-iteration(s) x {cells_local} cells x "
                                    f"{per_kernel_bytes.get(kname, 0)} B",
                "avg_launch_ms": kavg_ms,
                "launches": kcnt, "share_of_step": ktot / ms,
                "whole_step": {"bytes_per_cell": BYTES_STEP, "achieved": step_gbs,
                               "frac": step_gbs / peak},
                "kernels": {k: {"ms_total": v[0], "launches": v[1], "share": v[0] / ms}
                            for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])}}
        if units > 1:
            # temporal blocking: the launch must move only one sub-iteration's bytes; its own
            # DRAM-bound time is therefore half of what `frac` is measured against
            roof["temporal_blocking"] = {
                "compulsory_dram_bytes_per_launch": kbytes // units,
                "dram_gbs": kbytes / units / (kavg_ms * 1e-3) / 1e9,
                "dram_frac_of_peak": kbytes / units / (kavg_ms * 1e-3) / 1e9 / peak,
                "note": "frac > dram_frac_of_peak because two sub-iterations share one pass "
                        "over HBM (profiles/: marching kernel 79 % of the measured HBM "
                        "peak, FP64 pipe 55 %)"}
        cpu = None
        if args.gpus == 1 and not args.no_cpu:
            n = args.cpu_sample
            Fc, hxc, hyc = S.box_case(n, n, NCOMP)
            secs = cpu_step(Fc, hxc, hyc, mp, 1)
            cpu = {"value": n * n / secs, "unit": "cell-updates/s", "cores": os.cpu_count() or 1,
                   "kind": "port",
                   "sample": f"{n}x{n}-cell window of the same box test, 1 full step "
                             f"(transport + {NSUB} sub-iterations), oracle/ C + OpenMP",
                   "note": "parity unpinned against draenog/nextsimdg (no dynamics in the snapshot)"}
        line = {"metric": "cell-updates/sec (dG2 transport+mEVP step)", "value": value,
                "unit": "cell-updates/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_max / args.steps,
                "higher_is_better": True,
                "scaling": "strong" if args.workload == "16384sq" else "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": workload_config(args, world), "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": "cell-updates/s", "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": d2h, "steps": e2e_steps, "host_numa_binding": numa,
                        "what": "nsdg_dyn_step_host_async per step + nsdg_dyn_host_wait inside "
                                "the timed region: every step uploads its forcing (wind, ocean "
                                "current) from pinned host memory and downloads the full prognostic "
                                "state (H, A, U, V, S); the downloads of step k overlap the compute "
                                "of step k+1; the state stays resident between steps",
                        "fields_only": {"value": e2e_fields, "d2h_bytes_per_step": d2h_fields,
                                        "what": "same call, downloads H, A, U, V only (the stress "
                                                "stays resident as internal state of the rheology)"},
                        "blocking": {"value": e2e_blocking,
                                     "what": "nsdg_dyn_step_host: same copies, each call returns "
                                             "with its outputs on the host"},
                        "full_roundtrip": {"value": e2e_full, "h2d_bytes_per_step": h2d_full,
                                           "d2h_bytes_per_step": d2h,
                                           "what": "blocking call, whole state uploaded every step too"}},
                "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cpu,
                "max_abs_u": umax,
                "halo_bytes_sent_per_rank": ctx.comm_bytes_sent if world > 1 else 0}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()


# ---------------------------------------------------------------------------------------------
# --workload thermo: BASELINE configs[3], the per-cell column thermodynamics (the part of the
# path that EXISTS in the reference: parity pinned, CPU arm = the reference's own sources)
THERMO_BYTES = 160  # 15 input + 5 output doubles per cell-update
THERMO_DT = 600.0


def thermo_config(args, n):
    return {"workload": f"synthetic {args.nx}x{args.ny} per-cell column thermodynamics "
                        f"(NextsimPhysics/ThermoIce0), {n} GPU(s), independent cells",
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


def run_thermo_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from oracle import refthermo as R
    cores = os.cpu_count() or 1
    n = 1 << 21
    if R.available():
        value, secs = thermo_reference_rate(n, args.steps, cores)
        one, _ = thermo_reference_rate(n, 1, 1)
        kind, note = "reference", f"single thread (as the reference runs): {one:.4g} cell-updates/s"
    else:  # the compiled reference did not travel: the C restatement of the same arithmetic
        from oracle import cpu_oracle as O
        from nextsimdg_b200 import synthetic as S
        from nextsimdg_b200._abi import ThermoParams
        st, fo = S.thermo_inputs(n)
        cores = os.cpu_count() or 1
        t0 = time.perf_counter()
        O.thermo_step(ThermoParams.defaults(), THERMO_DT, args.steps, st, fo, nthreads=cores)
        secs = time.perf_counter() - t0
        value, kind, note = n * args.steps / secs, "port", "oracle/thermo_port.c (OpenMP)"
    line = {"impl": "reference", "metric": "cell-updates/sec (column thermodynamics)", "value": value,
            "unit": "cell-updates/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": thermo_config(args, args.gpus),
            "cpu_baseline": {"value": value, "unit": "cell-updates/s", "cores": cores, "kind": kind,
                             "sample": f"{n} cells of the same synthetic distribution, "
                                       f"{args.steps} sweeps", "note": note},
            "e2e": {"value": value, "unit": "cell-updates/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}
    emit(line)


def run_thermo(args):
    import torch
    import torch.distributed as dist
    from nextsimdg_b200 import synthetic as S
    from nextsimdg_b200._abi import ThermoParams
    from nextsimdg_b200.gpu import Context, GpuStep

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this framework has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ctx = Context(local)
    n = args.nx * args.ny  # cells of this rank: independent columns, no collective
    st, fo = S.thermo_inputs(n, seed=S.THERMO_SEED + rank)
    prog = dict(st)
    prog["tice"] = prog.pop("tice0")
    step = GpuStep(ctx, n, 1, ThermoParams.defaults())
    step.setInitialData(prog, fo)
    for _ in range(args.warmup):
        step.iterate(THERMO_DT, 1)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ctx.launches
    ctx.profile_begin()
    ctx.timer_start()
    for _ in range(args.steps):
        step.iterate(THERMO_DT, 1)
    ms = ctx.timer_stop()
    prof = ctx.profile_end()
    launches = ctx.launches - l0
    barrier()
    extra = 0
    if ms < 400.0:  # a sweep takes 0.4 ms: keep the same kernel running until nvidia-smi has samples
        extra = 1000
        step.iterate(THERMO_DT, extra)
        barrier()
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None and extra:
        clocks["sampled_over"] = f"the timed sweeps plus {extra} more right after them"
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = n * world * args.steps / (ms_max * 1e-3)
    got = step.get_prognostic()
    if not all(np.isfinite(v).all() for v in got.values()):
        raise SystemExit("bench.py: non-finite thermodynamic state after the timed steps")

    # ---- e2e: nsdg_thermo_run_host with pinned host buffers: 14 arrays up, one sweep, 6 down
    e2e_value, h2d, d2h = None, 0, 0
    if not args.no_e2e:
        def pin(a):
            tpin = torch.empty(a.shape, dtype=torch.float64).pin_memory()
            tpin.numpy()[...] = a
            return tpin
        keep = {k: pin(v) for k, v in list(prog.items()) + list(fo.items())}
        hp = {k: keep[k].numpy() for k in prog}
        hf = {k: keep[k].numpy() for k in fo}
        step.run_host(THERMO_DT, 1, hp, hf)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step.run_host(THERMO_DT, 1, hp, hf)
        barrier()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e_value = n * world * args.steps / float(te.item())
        h2d, d2h = 14 * n * 8, 6 * n * 8
    if rank == 0:
        peak, peak_src = peaks()
        ktot, kcnt = prof["thermo_kernel"]
        kavg = ktot / kcnt
        achieved = THERMO_BYTES * n / (kavg * 1e-3) / 1e9
        cpu = None
        if args.gpus == 1 and not args.no_cpu:
            from oracle import refthermo as R
            if R.available():
                cores = os.cpu_count() or 1
                m = 1 << 21
                rate, _ = thermo_reference_rate(m, 3, cores)
                one, _ = thermo_reference_rate(m, 1, 1)
                cpu = {"value": rate, "unit": "cell-updates/s", "cores": cores, "kind": "reference",
                       "sample": f"{m} cells of the same distribution, 3 sweeps, the reference's own "
                                 "column physics compiled in place (oracle/_ref)",
                       "single_thread": one}
        line = {"metric": "cell-updates/sec (column thermodynamics)", "value": value,
                "unit": "cell-updates/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": thermo_config(args, world), "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": "cell-updates/s", "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": d2h,
                        "what": "nsdg_thermo_run_host per step: prognostic state + forcing up from "
                                "pinned host memory, one sweep, prognostic state down"},
                "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "kernel": "thermo_kernel", "achieved": achieved,
                             "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "peak_source": peak_src, "traffic": None,
                             "algorithmic_bytes_per_launch": THERMO_BYTES * n,
                             "units_per_launch": f"{n} cells x {THERMO_BYTES} B",
                             "avg_launch_ms": kavg, "launches": kcnt, "share_of_step": ktot / ms},
                "cpu_baseline": cpu}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()


# ---------------------------------------------------------------------------------------------
# --workload cone: BASELINE configs[1], 1024^2 dG1 rotating-cone advection (periodic, Heun RK2)
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


def run_cone(args):
    import torch
    import torch.distributed as dist
    from nextsimdg_b200.gpu import Context, GpuDynamics

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this framework has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ctx = Context(local)
    phi, u, v, hx, hy, dt = cone_setup()
    n = CONE_N * CONE_N
    dyn = GpuDynamics(ctx, CONE_N, CONE_N, hx, hy, 3)
    dyn.upload_all(dict(H=phi, U=u, V=v))
    m0 = dyn.mass("H")
    steps = max(args.steps, 2000)  # a step takes ~30 us: time enough of them for the clocks sampler
    for _ in range(max(args.warmup, 3)):
        dyn.transport_step(dt, "H")
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = ctx.launches
    ctx.timer_start()
    for _ in range(steps):
        dyn.transport_step(dt, "H")
    ms = ctx.timer_stop()
    launches = ctx.launches - l0
    barrier()
    # per-launch times from a short profiled run (the event pairs would perturb the timed one)
    ctx.profile_begin()
    for _ in range(50):
        dyn.transport_step(dt, "H")
    prof = ctx.profile_end()
    time.sleep(0.3)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = n * world * steps / (ms_max * 1e-3)
    drift = abs(dyn.mass("H") - m0) / m0
    got = dyn.download("H")
    if not np.isfinite(got).all():
        raise SystemExit("bench.py: non-finite tracer after the timed steps")
    e2e_value, h2d, d2h = None, 0, 0
    if not args.no_e2e:
        # per step: the field goes up from the host, one step, the field comes down
        esteps = 50
        host = np.ascontiguousarray(got)
        barrier()
        t0 = time.perf_counter()
        for _ in range(esteps):
            dyn.upload("H", host)
            dyn.transport_step(dt, "H")
            host = dyn.download("H")
        barrier()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e_value = n * world * esteps / float(te.item())
        h2d = d2h = 3 * n * 8
    if rank == 0:
        peak, peak_src = peaks()
        ktot, kcnt = prof["transport_stage_kernel"]
        kavg = ktot / kcnt
        kbytes = CONE_BYTES * n / 2.0  # two stage launches per step
        achieved = kbytes / (kavg * 1e-3) / 1e9
        cpu = None
        if args.gpus == 1 and not args.no_cpu:
            from oracle import cpu_oracle as O
            p2 = phi
            t0 = time.perf_counter()
            for _ in range(20):
                p2 = O.transport_step(p2, u, v, hx, hy, dt)
            secs = time.perf_counter() - t0
            cpu = {"value": n * 20 / secs, "unit": "cell-updates/s", "cores": os.cpu_count() or 1,
                   "kind": "port", "sample": f"the full {CONE_N}^2 grid, 20 steps, oracle/dg_transport.c"}
        line = {"metric": "cell-updates/sec (dG1 transport step)", "value": value,
                "unit": "cell-updates/s", "n_gpus": world, "steps": steps, "warmup": max(args.warmup, 3),
                "ms_per_step": ms_max / steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cone_config(world),
                "clocks": clocks,
                "e2e": {"value": e2e_value, "unit": "cell-updates/s", "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": d2h,
                        "what": "nsdg_dyn_upload + nsdg_transport_step + nsdg_dyn_download per step"},
                "gpu_launches": int(launches), "mass_drift_rel": drift,
                "roofline": {"bound": "hbm", "kernel": "transport_stage_kernel", "achieved": achieved,
                             "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "peak_source": peak_src, "traffic": None,
                             "algorithmic_bytes_per_launch": kbytes,
                             "units_per_launch": f"{n} cells x {CONE_BYTES}/2 B (one of two RK stages)",
                             "avg_launch_ms": kavg, "launches": kcnt,
                             "whole_step": {"bytes_per_cell": CONE_BYTES,
                                            "achieved": CONE_BYTES * n * steps / (ms * 1e-3) / 1e9,
                                            "frac": CONE_BYTES * n * steps / (ms * 1e-3) / 1e9 / peak}},
                "cpu_baseline": cpu}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()


def use_all_host_threads():
    """The CPU arms use every host core; torchrun presets OMP_NUM_THREADS=1 for its workers, which
    would silently turn the OpenMP oracle into a single-thread run."""
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)


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
    ap.add_argument("--cpu-sample", type=int, default=768, help="edge of the CPU sample window")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true",
                    help="skip the host-buffer legs (the 16384^2 grid at N = 1 leaves no room for them)")
    ap.add_argument("--traffic", type=float, default=None,
                    help="dram bytes per launch of the dominant kernel from an ncu capture")
    args = ap.parse_args()
    if args.impl == "reference":
        use_all_host_threads()  # before the oracle library (and its OpenMP runtime) is loaded
    if args.workload == "thermo":
        run_thermo_reference(args) if args.impl == "reference" else run_thermo(args)
    elif args.workload == "cone":
        run_cone_reference(args) if args.impl == "reference" else run_cone(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
