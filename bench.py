#!/usr/bin/env python
"""Headline benchmark: steady.3D reaction-diffusion to steady state (BASELINE.json config C5).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU algorithm (oracle port)

A *step* is one complete solve to steady state from the same initial state: the
Newton loop of src/sparse.f:114-188 (grouped FD Jacobian, linear solve, update,
convergence tests).  `value` times K such steps with y0 already resident in HBM
(structure/grouping/CSR were built once by rs_solver_create, reported as
setup_s); `e2e` times the reference-facing C-ABI call rs_stsparse() with HOST
buffers -- host->device copy of y0, set-up, Newton loop, device->host copy of y
all inside the timed region.  One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
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
TOL = 1e-10
METRIC = "steady.3D newton throughput (unknowns x Newton iterations per second of time-to-steady-state)"
UNIT = "unknown-iterations/s"


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


# ------------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference's own algorithm on the host cores: the C oracle port of dsparse + Yale LDU
    (the reference itself needs R + Fortran, neither exists on this image).  Single-threaded, as the
    reference is.  Each step is a bounded sample of the workload: the same model on an S^3 grid."""
    rank, _, _ = dist_env()
    if rank != 0:
        return
    from oracle import oracle as O
    S = args.ref_size
    n = S ** 3
    y0 = np.full(n, 10.0)
    f = O.model("diff3d", PARMS)
    lrw = int(min(2 ** 30, 40 * n * max(8.0, np.log2(n))))

    def step():
        t0 = time.perf_counter()
        r = O.stsparse(y0, f, sparsetype="3D", nnz=[1, S, S, S, 0, 0, 0], atol=TOL, rtol=TOL, ctol=TOL, lrw=lrw,
                       positive=True, ipar=[1, S, S, S, 0])
        return time.perf_counter() - t0, r
    for _ in range(args.warmup):
        step()
    times, iters = [], 0
    for _ in range(args.steps):
        dt, r = step()
        times.append(dt)
        iters = len(r["precis"])
    total = sum(times)
    value = n * iters * args.steps / total
    sample = f"steady.3D {S}^3 (N={n}) same model/tolerances, full solve incl. structure+ordering, Yale direct LDU"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"steady.3D {args.size}^3 reaction-diffusion, nspec=1 (BASELINE config C5), stodes path",
                   "bounded_sample": f"{S}^3 (the Yale direct solve cannot reach {args.size}^3: 32-bit work array, ~N^2 time)",
                   "newton_iters": iters, "atol=rtol=ctol": TOL},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "time_to_steady_state_s": total / args.steps,
    }
    print(json.dumps(line), flush=True)


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
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    L = _capi.lib()
    if world > 1:
        from rootsolve_b200 import parallel
        parallel.init_comm(dist)

    S = args.size
    n = S ** 3
    Z = n * 7 - 2 * 3 * S * S
    ipar = [1, S, S, S, 0]
    # build the argument block exactly as steady.3D -> stodes does, then point y at pinned host memory
    y0_host = torch.full((n,), 10.0, dtype=torch.float64).pin_memory()
    yout_host = torch.empty(n, dtype=torch.float64).pin_memory()
    P = host.pack_args(y0_host.numpy(), 0.0, "diff3d", PARMS, TOL, TOL, TOL, "3D", False, [1, S, S, S, 0, 0, 0], None,
                       None, None, True, 100, None, "rs_models", PARMS, None, ipar, 0, "yale", None)
    P.args.y = C.cast(y0_host.data_ptr(), _capi.c_double_p)
    precis = np.zeros(100)

    def new_result(with_y):
        R = _capi.StsparseResult()
        R.precis = precis.ctypes.data_as(_capi.c_double_p)
        if with_y:
            R.y = C.cast(yout_host.data_ptr(), _capi.c_double_p)
        return R

    # ---- device-resident arm -----------------------------------------------------------------
    t0 = time.perf_counter()
    h = C.c_void_p()
    _capi.check(L.rs_solver_create(C.byref(P.args), C.byref(h)))
    torch.cuda.synchronize()
    setup_s = time.perf_counter() - t0
    y0_dev = torch.full((n,), 10.0, dtype=torch.float64, device="cuda")

    def step(R):
        _capi.check(L.rs_solver_set_state(h, y0_dev.data_ptr(), 1))
        _capi.check(L.rs_solver_run(h, C.byref(R)))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    R = new_result(False)
    for _ in range(args.warmup):
        step(R)
    # per-kernel CUDA events over the timed region (N = 1: the roofline line).  Sharded runs replay
    # captured CUDA graphs, which cannot carry per-launch events, so their kernel table is left empty.
    _capi.check(L.rs_solver_set_profiling(h, 1 if world == 1 else 0))
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    launches = kry = newton = 0
    for _ in range(args.steps):
        step(R)
        launches += R.kernel_launches
        kry += R.krylov_iters
        newton = R.niter
    ev1.record()
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    steady = bool(R.steady[0])
    ngp = int(R.steady[2])
    # per-kernel device time over the timed region
    cap = 32
    names = (C.c_char_p * cap)()
    kms = (C.c_double * cap)()
    kl = (C.c_longlong * cap)()
    kb = (C.c_double * cap)()
    nk = L.rs_solver_kernel_stats(h, cap, names, kms, kl, kb)
    kstats = {names[i].decode(): {"ms": kms[i], "launches": int(kl[i]), "bytes_per_launch": kb[i]} for i in range(nk)}
    _capi.check(L.rs_solver_set_profiling(h, 0))
    L.rs_solver_destroy(h)

    # ---- end-to-end arm: the reference-facing call with host buffers ------------------------------
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    Re = new_result(True)
    _capi.check(L.rs_stsparse(C.byref(P.args), C.byref(Re)))   # warm-up
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        _capi.check(L.rs_stsparse(C.byref(P.args), C.byref(Re)))
    barrier()
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_newton = Re.niter
    e2e_setup, e2e_loop = Re.t_setup_s, Re.t_newton_s

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    pk, pk_src = peaks()
    # DRAM traffic of the SpMV kernel per launch from the committed `ncu --set full` capture (256^3 only)
    traffic = None
    # the solver runs the implicit-column SpMV on this grid (kernel class "spmv_grid"); "spmv_csr" is the general kernel
    spname = "spmv_grid" if kstats.get("spmv_grid", {}).get("launches", 0) > 0 else "spmv_csr"
    prof = os.path.join(ROOT, "profiles", "r01_ncu_spmv_grid_full.json" if spname == "spmv_grid" else "r01_ncu_spmv_tma_full.json")
    if S == 256 and world == 1 and os.path.exists(prof):
        try:
            unit = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
            rows = json.load(open(prof))
            tot = []
            for r_ in rows:
                t_ = 0.0
                for key in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                    v_, u_ = r_[key].split()
                    t_ += float(v_) * unit[u_]
                tot.append(t_)
            traffic = sum(tot) / len(tot)
        except Exception:
            traffic = None
    sp = kstats.get(spname, {"ms": 0.0, "launches": 0, "bytes_per_launch": 0.0})
    ach = (sp["bytes_per_launch"] * sp["launches"] / (sp["ms"] * 1e-3) / 1e9) if sp["ms"] > 0 else 0.0
    ms_per_step = ms_total / args.steps
    value = n * newton / (ms_per_step * 1e-3)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"steady.3D {S}^3 reaction-diffusion, nspec=1 (BASELINE config C5), stodes path",
                   "unknowns": n, "nnz": Z, "ngp": ngp, "newton_iters": newton, "krylov_iters_per_step": kry // args.steps,
                   "atol=rtol=ctol": TOL, "cache": "inputs_exceed_l2 (matrix values 8Z bytes + vectors >> 126 MB)",
                   "parallelism": f"slab{world}" if world > 1 else "single-gpu", "steady": steady},
        "time_to_steady_state_s": ms_per_step * 1e-3, "newton_iters_per_s": newton / (ms_per_step * 1e-3),
        "setup_s": setup_s,
        "roofline": {"bound": "hbm", "kernel": spname + (" (implicit columns: 8Z+4(N+1)+16N bytes, <p,Ap> fused)" if spname == "spmv_grid"
                                                        else " (12Z+4(N+1)+16N bytes, dots fused)"), "achieved": ach, "peak": pk["hbm_gbs"], "unit": "GB/s",
                     "frac": ach / pk["hbm_gbs"] if pk["hbm_gbs"] else None, "traffic": traffic, "algorithmic_bytes": sp["bytes_per_launch"],
                     "peak_source": pk_src,
                     "avg_launch_ms": sp["ms"] / sp["launches"] if sp["launches"] else None},
        "kernels": {k: {"ms_per_step": v["ms"] / args.steps, "launches_per_step": v["launches"] / args.steps,
                        "GBps": (v["bytes_per_launch"] * v["launches"] / (v["ms"] * 1e-3) / 1e9) if v["ms"] > 0 else None}
                    for k, v in kstats.items()},
        "e2e": {"value": n * e2e_newton / e2e_s, "unit": UNIT, "h2d_bytes_per_step": 8 * n, "d2h_bytes_per_step": 8 * n,
                "s_per_step": e2e_s, "setup_s": e2e_setup, "newton_loop_s": e2e_loop, "steps": e2e_steps},
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    if world == 1 and not args.no_cpu_baseline:
        from oracle import oracle as O
        Sb = args.ref_size
        nb = Sb ** 3
        t0 = time.perf_counter()
        r = O.stsparse(np.full(nb, 10.0), O.model("diff3d", PARMS), sparsetype="3D", nnz=[1, Sb, Sb, Sb, 0, 0, 0],
                       atol=TOL, rtol=TOL, ctol=TOL, lrw=int(min(2 ** 30, 40 * nb * max(8.0, np.log2(nb)))), positive=True,
                       ipar=[1, Sb, Sb, Sb, 0])
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": nb * len(r["precis"]) / dt, "unit": UNIT, "cores": 1, "kind": "port",
                                "sample": f"steady.3D {Sb}^3 (N={nb}), same model/tolerances, one full solve = {dt:.2f} s "
                                          f"(Yale LDU factor {r['timing']['factor']:.2f} s + solves {r['timing']['solve']:.2f} s)"}
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
    ap.add_argument("--ref-size", type=int, default=24, help="grid edge of the bounded CPU sample")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
