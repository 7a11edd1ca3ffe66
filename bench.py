#!/usr/bin/env python
"""Headline benchmark: steady.3D reaction-diffusion to steady state (BASELINE.json config C5).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU algorithm (oracle port)

A *step* is one complete solve to steady state from the same initial state: the Newton loop of
src/sparse.f:114-188 (grouped FD Jacobian, linear solve, update, convergence tests).
  value     K such steps timed with y0 already resident in HBM, per-launch profiling OFF (structure / grouping /
            row-wise pattern were built once by rs_solver_create and are reported as setup_s);
  e2e       the reference-facing C-ABI call rs_stsparse() with HOST buffers: host->device copy of y0, set-up,
            Newton loop, device->host copy of y all inside the timed region;
  kernels   per-kernel-class CUDA-event times from a SECOND, separate pass with profiling on;
  parity    the reference's acceptance quantity max|f/ewt| (src/sparse.f:126-129) at the returned state, f evaluated
            by the oracle's HOST model; for N > 1 also max|y_N - y_1| / (atol + rtol|y|) against a single-GPU solve
            made by every rank before the communicator exists;
  same_config  both arms -- this library end to end, and the oracle (the reference's algorithm, Yale LDU, one core)
            -- on the SAME inputs, for every BASELINE configuration or smaller instance of it the oracle finishes in
            about a minute, with the oracle's phase times and the fitted growth exponent of its run time.  The
            headline 256^3 cannot be run by the reference at all (32-bit work array, fill ~ N^(4/3)); a speed-up
            is only ever quoted from this table.
One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PARMS = [1.0, 1.0, 1.0, 1.0, 1.0, 1.0, 0.025, 1.0]   # Dx Dy Dz dx dy dz r BND (man/steady.3D.Rd:105-140)
P_MONOD = [1.0, 1.0, 1.0, 1.0, 0.25, 1.0, 1.0]        # config C1 / C3
P_RD3D = [1.0, 0.5, 1.0, 1.0, 1.0, 0.01, 0.2, 10.0, 1.0, 0.0]   # config C4
P_RT1D = [1.0, 1.0, 1.0, 0.5, 0.5, 1.0, 0.1]          # config C2
TOL = 1e-10
METRIC = "steady.3D newton throughput (unknowns x Newton iterations per second of time-to-steady-state)"
UNIT = "unknown-iterations/s"


def case(kind, size):
    """One workload: what both arms need to run it.  kind: '3D' (C5 model), '2D' (C1/C3 model), '3Dx2' (C4), '1Dx5' (C2)."""
    if kind == "3D":
        S = size
        return dict(name=f"steady.3D {S}^3 diffusion-reaction (C5 model)", n=S ** 3, model="diff3d", parms=PARMS, st="3D",
                    nnz=[1, S, S, S, 0, 0, 0], ipar=[1, S, S, S, 0], y0=np.full(S ** 3, 10.0))
    if kind == "2D":
        S = size
        return dict(name=f"steady.2D {S}x{S} diffusion + Monod (C1/C3 model)", n=S * S, model="diff2d", parms=P_MONOD, st="2D",
                    nnz=[1, S, S, 0, 0], ipar=[1, S, S, 1, 1], y0=np.full(S * S, 10.0))
    if kind == "3Dx2":
        S = size
        nc = S ** 3
        return dict(name=f"steady.3D {S}^3 two-species reaction-diffusion (C4 model)", n=2 * nc, model="rd3d", parms=P_RD3D, st="3D",
                    nnz=[2, S, S, S, 0, 0, 0], ipar=[2, S, S, S, 0], y0=np.concatenate([np.full(nc, 10.0), np.full(nc, 1.0)]))
    if kind == "1Dx5":
        nb = size
        return dict(name=f"steady.1D 5 species x {nb} boxes reaction-transport (C2 model)", n=5 * nb, model="rt1d", parms=P_RT1D,
                    st="1D", nnz=[5, nb, 0], ipar=[5, nb, 1, 1, 0], y0=np.full(5 * nb, 1.0))
    raise ValueError(kind)


# sizes the oracle finishes in about a minute in total; the 3-D series carries the growth exponent
SAME_CONFIG = [("2D", 50), ("2D", 100), ("2D", 200), ("3D", 16), ("3D", 24), ("3D", 32), ("3Dx2", 16), ("1Dx5", 200000)]
SAME_CONFIG_QUICK = [("2D", 50), ("3D", 16), ("3D", 24)]
REF_ROW = ("3D", 24)       # the row the reference arm (--impl reference) times


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except ValueError:
                continue
            for nm, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def oracle_lrw(cs):
    n = cs["n"]
    return int(min(2 ** 31 - 1, 60 * n * max(8.0, math.log2(n)) if cs["st"] == "3D" else 60 * n * max(4.0, math.log2(n) / 2)))


def oracle_solve(cs):
    """the reference's algorithm on this workload: one core, Yale ordering + LDU; returns (seconds, result)"""
    from oracle import oracle as O
    t0 = time.perf_counter()
    r = O.stsparse(cs["y0"], O.model(cs["model"], cs["parms"]), sparsetype=cs["st"], nnz=cs["nnz"], atol=TOL, rtol=TOL, ctol=TOL,
                   lrw=oracle_lrw(cs), positive=True, ipar=cs["ipar"])
    return time.perf_counter() - t0, r


def host_accept(cs, y):
    """max |f / ewt| at y with f from the oracle's host model (src/sparse.f:126-129)"""
    from oracle import oracle as O
    func = C.cast(O.model(cs["model"], cs["parms"]), O.DERIV)
    n = len(y)
    f = np.zeros(n)
    ip = np.concatenate([[0, 0, 3 + len(cs["ipar"])], cs["ipar"]]).astype(np.int32)
    neq, t = C.c_int(n), C.c_double(0.0)
    func(C.byref(neq), C.byref(t), np.ascontiguousarray(y).ctypes.data_as(C.POINTER(C.c_double)),
         f.ctypes.data_as(C.POINTER(C.c_double)), np.zeros(1).ctypes.data_as(C.POINTER(C.c_double)),
         ip.ctypes.data_as(C.POINTER(C.c_int)))
    return float(np.max(np.abs(f) / (TOL * np.abs(y) + TOL)))


# ------------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference's own algorithm on the host cores: the C oracle port of dsparse + Yale LDU
    (the reference itself needs R + Fortran, neither exists on this image).  Single-threaded, as the
    reference is.  Each step is a bounded sample of the workload: the same model and tolerances on the
    largest grid of the same shape that finishes in seconds -- the row of this library's `same_config`
    table named in config.same_config_row; a ratio of the two arms is only meaningful on that row."""
    rank, _, _ = dist_env()
    if rank != 0:
        return
    kind, S = REF_ROW if args.ref_size is None else ("3D", args.ref_size)
    cs = case(kind, S)
    n = cs["n"]
    for _ in range(args.warmup):
        oracle_solve(cs)
    times, iters, r = [], 0, None
    for _ in range(args.steps):
        dt, r = oracle_solve(cs)
        times.append(dt)
        iters = len(r["precis"])
    total = sum(times)
    value = n * iters * args.steps / total
    tm = r["timing"]
    sample = (f"{cs['name']} (N={n}) same model/tolerances as the {args.size}^3 headline, full solve incl. structure + "
              f"ordering, Yale direct LDU; phases of the last solve: structure {tm['struct']:.3f} s, ordering {tm['order']:.3f} s, "
              f"FD Jacobian {tm['jac']:.3f} s, LDU factor {tm['factor']:.3f} s, solves {tm['solve']:.3f} s")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"steady.3D {args.size}^3 reaction-diffusion, nspec=1 (BASELINE config C5), stodes path",
                   "bounded_sample": f"{S}^3 (the Yale direct solve cannot reach {args.size}^3: 32-bit work array, ~N^2 time)",
                   "same_config": False, "same_config_row": cs["name"],
                   "note": "not the same inputs as the 256^3 line of the other arm: compare with its same_config row of this name",
                   "newton_iters": iters, "atol=rtol=ctol": TOL},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "time_to_steady_state_s": total / args.steps,
        "oracle_phases_s": tm,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
class GpuCase:
    """A workload packed for the C ABI with pinned host buffers (what steady.xD -> stodes builds)."""

    def __init__(self, cs, host, _capi, torch):
        self.cs, self.capi = cs, _capi
        n = cs["n"]
        self.y0_host = torch.from_numpy(cs["y0"].copy()).pin_memory()
        self.yout_host = torch.empty(n, dtype=torch.float64).pin_memory()
        self.P = host.pack_args(self.y0_host.numpy(), 0.0, cs["model"], cs["parms"], TOL, TOL, TOL, cs["st"], False, cs["nnz"], None,
                                None, None, True, 100, None, "rs_models", cs["parms"], None, cs["ipar"], 0, "yale", None)
        self.P.args.y = C.cast(self.y0_host.data_ptr(), _capi.c_double_p)
        self.precis = np.zeros(100)

    def result(self, with_y):
        R = self.capi.StsparseResult()
        R.precis = self.precis.ctypes.data_as(self.capi.c_double_p)
        if with_y:
            R.y = C.cast(self.yout_host.data_ptr(), self.capi.c_double_p)
        return R


def kernel_table(L, h, steps):
    cap = 40
    names = (C.c_char_p * cap)()
    kms = (C.c_double * cap)()
    kl = (C.c_longlong * cap)()
    kb = (C.c_double * cap)()
    nk = L.rs_solver_kernel_stats(h, cap, names, kms, kl, kb)
    raw = {names[i].decode(): {"ms": kms[i], "launches": int(kl[i]), "bytes_per_launch": kb[i]} for i in range(nk)}
    tab = {k: {"ms_per_step": v["ms"] / steps, "launches_per_step": v["launches"] / steps,
               "algorithmic_bytes_per_launch": v["bytes_per_launch"],
               "GBps": (v["bytes_per_launch"] * v["launches"] / (v["ms"] * 1e-3) / 1e9) if v["ms"] > 0 else None}
           for k, v in raw.items()}
    return raw, tab


def gpu_e2e(L, G, torch, reps, barrier=None):
    """rs_stsparse with host buffers, best-effort mean of `reps` calls after one warm-up; returns (seconds, result)"""
    _capi = G.capi
    R = G.result(True)
    _capi.check(L.rs_stsparse(C.byref(G.P.args), C.byref(R)))
    if barrier:
        barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        _capi.check(L.rs_stsparse(C.byref(G.P.args), C.byref(R)))
    torch.cuda.synchronize()
    if barrier:
        barrier()
    return (time.perf_counter() - t0) / reps, R


def profiled_config(L, host, _capi, torch, cs, steps=2):
    """One BASELINE configuration other than the headline: solve time (profiling off) and its kernel table (second pass)."""
    G = GpuCase(cs, host, _capi, torch)
    h = C.c_void_p()
    _capi.check(L.rs_solver_create(C.byref(G.P.args), C.byref(h)))
    y0_dev = torch.from_numpy(cs["y0"]).cuda()
    R = G.result(False)

    def step():
        _capi.check(L.rs_solver_set_state(h, y0_dev.data_ptr(), 1))
        _capi.check(L.rs_solver_run(h, C.byref(R)))
    step()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(steps):
        step()
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / steps
    _capi.check(L.rs_solver_set_profiling(h, 1))
    for _ in range(steps):
        step()
    _, tab = kernel_table(L, h, steps)
    _capi.check(L.rs_solver_set_profiling(h, 0))
    y = np.zeros(cs["n"])
    _capi.check(L.rs_solver_get_state(h, y.ctypes.data_as(_capi.c_double_p)))
    L.rs_solver_destroy(h)
    prec = {0: "none", 1: "jacobi", 2: "block-jacobi", 3: "mg", 4: "btd"}.get(R.precond)
    return {"workload": cs["name"], "unknowns": cs["n"], "ms_per_solve": ms, "newton_iters": R.niter, "krylov_iters": R.krylov_iters,
            "steady": bool(R.steady[0]), "precond": prec, "krylov_method": "cg" if R.krylov_method else "bicgstab",
            "parity_max_f_over_ewt": host_accept(cs, y), "kernels": tab}


# ------------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist
    from rootsolve_b200 import _capi
    import importlib
    host = importlib.import_module("rootsolve_b200.stodes")

    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- rootsolve_b200 has no CPU path to measure")
    torch.cuda.set_device(local_rank)
    # libraries (NCCL's version banner) write to stdout: everything but the JSON line goes to stderr
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    L = _capi.lib()

    S = args.size
    cs = case("3D", S)
    n = cs["n"]
    Z = n * 7 - 2 * 3 * S * S
    G = GpuCase(cs, host, _capi, torch)

    # N > 1: every rank first solves the whole problem alone -- the single-GPU state the sharded one is checked against
    y_single = None
    if world > 1:
        _, R1 = gpu_e2e(L, G, torch, 1)
        y_single = G.yout_host.numpy().copy()
        from rootsolve_b200 import parallel
        parallel.init_comm(dist)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(v):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident arm, profiling off --------------------------------------------------------
    t0 = time.perf_counter()
    h = C.c_void_p()
    _capi.check(L.rs_solver_create(C.byref(G.P.args), C.byref(h)))
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t0
    y0_dev = torch.from_numpy(cs["y0"]).cuda()
    R = G.result(False)

    def step():
        _capi.check(L.rs_solver_set_state(h, y0_dev.data_ptr(), 1))
        _capi.check(L.rs_solver_run(h, C.byref(R)))

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    launches = kry = newton = 0
    for _ in range(args.steps):
        step()
        launches += R.kernel_launches
        kry += R.krylov_iters
        newton = R.niter
    ev1.record()
    barrier()
    ms_total = allmax(ev0.elapsed_time(ev1))
    clocks = sampler.stop() if rank == 0 else None
    steady = bool(R.steady[0])
    ngp = int(R.steady[2])
    precond = {0: "none", 1: "jacobi", 2: "block-jacobi", 3: "mg", 4: "btd"}.get(R.precond)
    method = "cg" if R.krylov_method else "bicgstab"

    # ---- second pass: per-kernel CUDA events (single GPU: graphs are off while profiling) -------------------------
    # (sharded: rank 0's table; a kernel's time then includes its waits for the neighbours' halos and partial sums)
    raw, ktab, psteps = {}, {}, max(1, min(args.steps, 3))
    _capi.check(L.rs_solver_set_profiling(h, 1))
    for _ in range(psteps):
        step()
    raw, ktab = kernel_table(L, h, psteps)
    _capi.check(L.rs_solver_set_profiling(h, 0))
    L.rs_solver_destroy(h)

    # ---- end-to-end arm: the reference-facing call with host buffers ------------------------------
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    e2e_s, Re = gpu_e2e(L, G, torch, e2e_steps, barrier)
    e2e_s = allmax(e2e_s)
    e2e_newton, e2e_setup, e2e_loop = Re.niter, Re.t_setup_s, Re.t_newton_s
    y_final = G.yout_host.numpy().copy()
    parity = {"max_abs_f_over_ewt_host_model": None, "sharded_vs_single_gpu": None}
    if rank == 0:
        parity["max_abs_f_over_ewt_host_model"] = host_accept(cs, y_final)
        parity["accepted_by_reference_test"] = parity["max_abs_f_over_ewt_host_model"] <= 1.0
    if y_single is not None:
        parity["sharded_vs_single_gpu"] = allmax(float(np.max(np.abs(y_final - y_single) / (TOL + TOL * np.abs(y_single)))))

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- single GPU only: the other BASELINE configurations and the same-input table ----------------------
    other, table, exponent = {}, [], None
    if world == 1 and not args.headline_only:
        for key, kc in (("C4 steady.3D 128^3 x 2 species", case("3Dx2", 128)), ("C3 steady.2D 2048^2", case("2D", 2048)),
                        ("C2 steady.1D 5 x 200000", case("1Dx5", 200000)), ("C1 steady.2D 50^2", case("2D", 50))):
            other[key] = profiled_config(L, host, _capi, torch, kc)
    if world == 1 and not args.no_cpu_baseline:
        rows = SAME_CONFIG_QUICK if args.quick else SAME_CONFIG
        for kind, size in rows:
            kc = case(kind, size)
            Gk = GpuCase(kc, host, _capi, torch)
            g_s, Rk = gpu_e2e(L, Gk, torch, 3)
            yk = Gk.yout_host.numpy().copy()
            o_s, ro = oracle_solve(kc)
            err = float(np.max(np.abs(yk - ro["y"]) / (TOL + TOL * np.abs(ro["y"]))))
            table.append({"workload": kc["name"], "unknowns": kc["n"], "gpu_e2e_s": g_s, "gpu_setup_s": Rk.t_setup_s,
                          "gpu_newton_loop_s": Rk.t_newton_s, "gpu_newton_iters": Rk.niter, "gpu_krylov_iters": Rk.krylov_iters,
                          "oracle_total_s": o_s, "oracle_newton_iters": len(ro["precis"]),
                          "oracle_phases_s": {k: ro["timing"][k] for k in ("struct", "order", "jac", "factor", "solve")},
                          "speedup_e2e": o_s / g_s, "max_err_over_tol_vs_oracle": err, "same_inputs": True})
        pts = [(r["unknowns"], r["oracle_total_s"]) for r in table if r["workload"].startswith("steady.3D") and "C5" in r["workload"]]
        if len(pts) >= 2:
            lx, ly = np.log([p[0] for p in pts]), np.log([p[1] for p in pts])
            exponent = float(np.polyfit(lx, ly, 1)[0])

    pk, pk_src = peaks()
    # roofline of the dominant kernel: the level-0 smoothing pass of the multigrid cycle when the cycle ran, else the SpMV
    traffic, prof_note = None, None
    if "mg_smooth_l0" in raw and raw["mg_smooth_l0"]["ms"] >= raw.get("spmv_grid", {"ms": 0.0})["ms"]:
        dom = "mg_smooth_l0"
        dom_desc = ("mg_smooth_l0: Chebyshev smoothing passes of the finest multigrid level (slot-major single-precision matrix "
                    "4*7N + dinv 4N bytes + 2-4 vectors per pass: 48N / 53N / 61N; mg_pre / mg_post_first / mg_post_second kernels)")
        prof = os.path.join(ROOT, "profiles", "r02_ncu_mg_level_kernels_full.json")
    else:
        dom = "spmv_grid" if raw.get("spmv_grid", {}).get("launches", 0) > 0 else "spmv_csr"
        dom_desc = dom + (" (implicit columns: 8Z+4(N+1)+16N bytes, <p,Ap> fused)" if dom == "spmv_grid" else " (12Z+4(N+1)+16N bytes)")
        prof = os.path.join(ROOT, "profiles", "r01_ncu_spmv_grid_full.json" if dom == "spmv_grid" else "r01_ncu_spmv_tma_full.json")
    if S == 256 and world == 1 and os.path.exists(prof):
        try:
            pj = json.load(open(prof))
            traffic = pj.get("dram_bytes_per_launch_avg") if isinstance(pj, dict) else None
            if traffic is None and isinstance(pj, list):
                unit = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
                tot = []
                for r_ in pj:
                    t_ = 0.0
                    for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                        v_, u_ = r_[key].split()
                        t_ += float(v_) * unit[u_]
                    tot.append(t_)
                traffic = sum(tot) / len(tot)
            prof_note = os.path.relpath(prof, ROOT)
        except Exception:
            traffic = None
    sp = raw.get(dom, {"ms": 0.0, "launches": 0, "bytes_per_launch": 0.0})
    ach = (sp["bytes_per_launch"] * sp["launches"] / (sp["ms"] * 1e-3) / 1e9) if sp["ms"] > 0 else 0.0
    ms_per_step = ms_total / args.steps
    value = n * newton / (ms_per_step * 1e-3)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"steady.3D {S}^3 reaction-diffusion, nspec=1 (BASELINE config C5), stodes path",
                   "unknowns": n, "nnz": Z, "ngp": ngp, "newton_iters": newton, "krylov_iters_per_step": kry // args.steps,
                   "linear_solver": f"{method} + {precond}", "atol=rtol=ctol": TOL,
                   "cache": "inputs_exceed_l2 (matrix values 8Z bytes + vectors >> 126 MB)",
                   "parallelism": f"slab{world}" if world > 1 else "single-gpu", "steady": steady,
                   "timing": "value: profiling off; kernels: separate profiled pass"},
        "time_to_steady_state_s": ms_per_step * 1e-3, "newton_iters_per_s": newton / (ms_per_step * 1e-3),
        "setup_s": setup_s,
        "roofline": {"bound": "hbm", "kernel": dom_desc, "achieved": ach, "peak": pk["hbm_gbs"], "unit": "GB/s",
                     "frac": ach / pk["hbm_gbs"] if pk["hbm_gbs"] else None, "traffic": traffic, "traffic_source": prof_note,
                     "algorithmic_bytes": sp["bytes_per_launch"], "peak_source": pk_src,
                     "avg_launch_ms": sp["ms"] / sp["launches"] if sp["launches"] else None,
                     "share_of_step": (sp["ms"] / psteps) / ms_per_step if sp["ms"] > 0 else None},
        "kernels": ktab,
        "e2e": {"value": n * e2e_newton / e2e_s, "unit": UNIT, "h2d_bytes_per_step": 8 * n, "d2h_bytes_per_step": 8 * n,
                "s_per_step": e2e_s, "setup_s": e2e_setup, "newton_loop_s": e2e_loop, "steps": e2e_steps},
        "parity": parity,
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    if other:
        line["other_configs"] = other
    if table:
        ref_name = case(*REF_ROW)["name"]
        ref_row = next((r for r in table if r["workload"] == ref_name), None)
        line["same_config"] = {"rows": table, "oracle_time_growth_exponent_3d": exponent,
                               "note": "gpu_e2e_s: rs_stsparse with host buffers; oracle: the reference's algorithm (Yale LDU), 1 core; "
                                       "same model, inputs and tolerances in both columns",
                               "reference_arm_row": ref_name,
                               "speedup_on_reference_arm_row": ref_row["speedup_e2e"] if ref_row else None}
        big = max((r for r in table if r["workload"].startswith("steady.3D") and "C5" in r["workload"]), key=lambda r: r["unknowns"],
                  default=None)
        if big:
            line["cpu_baseline"] = {"value": big["unknowns"] * big["oracle_newton_iters"] / big["oracle_total_s"], "unit": UNIT,
                                    "cores": 1, "kind": "port",
                                    "sample": f"{big['workload']} (N={big['unknowns']}), one full solve = {big['oracle_total_s']:.2f} s "
                                              f"(FD Jacobian {big['oracle_phases_s']['jac']:.3f} s, LDU factor "
                                              f"{big['oracle_phases_s']['factor']:.2f} s, solves {big['oracle_phases_s']['solve']:.2f} s); "
                                              f"the reference cannot run {S}^3"}
    sys.stdout.flush()
    os.dup2(saved_stdout, 1)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--size", type=int, default=256, help="grid edge S of the S^3 workload")
    ap.add_argument("--ref-size", type=int, default=None, help="grid edge of the reference arm's bounded sample (default 24)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the same-input table (no oracle runs)")
    ap.add_argument("--quick", action="store_true", help="short same-input table")
    ap.add_argument("--headline-only", action="store_true", help="skip the other BASELINE configurations")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
