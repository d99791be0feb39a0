#!/usr/bin/env python
"""Benchmark of the hot path (see DESIGN.md "Measurement").

    python bench.py --gpus N --steps K --warmup W            # B200 engine (one rank per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K ...  # CPU reference arm (rank 0 only)

Workload (config.workload = "C4"): one trajectory of the reference's Ga/Al geometry at n_sea = 27 (N = 28
spins, 4.3 GB state vector, Fibonacci shell of dipolar_ensemble_with_rare.py:232-247), sea detuning
50 kHz (+12.5 kHz per extra rank: independent sweep points, no collective), rare drive ON, the sweep's
tolerances atol 1e-10 / rtol 1e-9.  A "step" is one Dormand-Prince 5(4) step attempt = 6 H-applies.
The N = 7 detuning sweep (C2, CI-sized window) is timed as well and reported under "sweep".
"""
from __future__ import annotations

import argparse
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

G_SEA, G_RARE, B0_T, F1A = 8.1812e7, 6.976e7, 3.0, 50_000.0


def sweep_params(n_sea, delta_hz, variant="center_on", t_final=3.0e-2, steps=2000, f1a=F1A):
    """DipolarRareParams as sweep_sea_detuning.py:414-668 derives them."""
    from quantumsimulations_b200 import DipolarRareParams

    f_az = G_SEA * B0_T / (2 * np.pi)
    f1r = (f1a**2 + f1a**2) ** 0.5
    b0 = 2 * np.pi * f_az / G_SEA
    f_rz = G_RARE * b0 / (2 * np.pi)
    return DipolarRareParams(
        n_sea=n_sea, gamma_sea=G_SEA, gamma_rare=G_RARE, B0_sea=b0, B0_rare=b0,
        B1_sea=2 * np.pi * f1a / G_SEA, B1_rare=2 * np.pi * f1r / G_RARE,
        omega_rf_sea=2 * np.pi * (f_az - delta_hz), omega_rf_rare=2 * np.pi * f_rz,
        phi_sea=np.pi / 2.0, phi_rare=np.pi / 2.0, dipolar_scale=1.0e-7 * 1.054571817e-34,
        shell_scale=0.282393e-9, t_final=t_final, steps=steps, drive_sea=True,
        drive_rare=(variant == "center_on"), is_spin_three_half=False, is_center_rare=(variant != "shell_off"),
        solver_atol=1e-10, solver_rtol=1e-9, solver_nsteps=10_000_000, solver_max_step=1e-5,
    )


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.marks = index, [], None, []

    def mark(self):
        """Rows between the first and the second mark are the timed region."""
        self.marks.append(len(self.rows))

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.index)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._pump, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *exc):
        if self.proc:
            self.proc.terminate()
            self.th.join(timeout=2)

    def summary(self):
        rows = self.rows
        if len(self.marks) == 2 and self.marks[1] > self.marks[0]:
            rows = rows[self.marks[0]:self.marks[1]]
        elif len(self.marks) == 2 and self.rows:          # region shorter than one sample: the nearest sample
            rows = rows[max(0, self.marks[0] - 1):self.marks[0] + 1]
        sm = [float(r[1]) for r in rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = sorted({n for r in rows if len(r) >= 8 for n, v in zip(names, r[4:8]) if v == "Active"})
        pw = []
        for r in rows:
            try:
                pw.append(float(r[3]))
            except (ValueError, IndexError):
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm), "power_w": statistics.median(pw) if pw else None}


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """Per-launch DRAM bytes of the dominant kernel from the committed ncu capture, if any."""
    path = os.path.join(ROOT, "profiles", "ncu_summary.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f)
    return None


# --------------------------------------------------------------------------------------------------------
# CPU baseline (oracle port): the row-gather product a CSR matvec performs, matrix-free, OpenMP on all cores
# --------------------------------------------------------------------------------------------------------
class CpuPort:
    """oracle/c on the C4 table with every host core (the team size is forced: torchrun exports OMP_NUM_THREADS=1)."""

    def __init__(self, n_sea, delta_hz):
        from oracle.c import build as oc
        from quantumsimulations_b200 import compile_table

        self.oc = oc
        self.cores = oc.set_threads(os.cpu_count() or 1)
        table = compile_table(sweep_params(n_sea, delta_hz))
        self.bm = table.bit_major()
        self.n = table.n
        self.psi = np.zeros(1 << self.n, dtype=np.complex128)
        rng = np.random.default_rng(0)
        self.psi[: 1 << 20] = rng.normal(size=1 << 20) + 1j * rng.normal(size=1 << 20)
        self.psi[table.psi0_index] = 1.0

    def seconds(self, row0, rows):
        t0 = time.perf_counter()
        self.oc.apply_h(self.bm["h"], self.bm["J"], self.psi, row0, row0 + rows)
        return time.perf_counter() - t0

    def rows_for(self, budget_s):
        """Largest power-of-two row count (at most the whole vector) whose H-apply fits budget_s."""
        self.seconds(0, 1 << 14)                      # first touch / thread start-up
        per_row = self.seconds(0, 1 << 18) / (1 << 18)
        rows = int(min(1 << self.n, max(1 << 18, budget_s / per_row)))
        return 1 << (rows.bit_length() - 1)

    def sample_text(self, rows, secs):
        whole = rows == 1 << self.n
        what = (f"one COMPLETE matrix-free H-apply (all 2^{self.n} rows)" if whole else
                f"EXTRAPOLATED from rows [0,2^{rows.bit_length() - 1}) of one matrix-free H-apply at N={self.n}")
        return (f"{what}, {np.mean(secs):.1f} s each x{len(secs)}, scaled to the 6 H-applies of a DP5 step attempt; "
                "the stage axpys / error norm of the step are not charged to the CPU")


def cpu_port_step_rate(n_sea, delta_hz, budget_s=25.0, samples=1):
    port = CpuPort(n_sea, delta_hz)
    rows = port.rows_for(budget_s / samples)
    secs = [port.seconds((s * rows) % (1 << port.n), rows) for s in range(samples)]
    rates = [1.0 / (6.0 * dt * (1 << port.n) / rows) for dt in secs]   # 6 H-applies per DP5 step attempt
    return {"value": float(np.mean(rates)), "unit": "ODE steps/s", "cores": port.cores, "kind": "port",
            "sample": port.sample_text(rows, secs), "extrapolated": rows != 1 << port.n,
            "rows_timed": rows, "seconds_per_sample": secs}


def cpu_zvode_rate(window=2.0e-3):
    """Reference-equivalent integrator (SciPy zvode-Adams as QuTiP drives it) on one N=7 sweep point, 1 core."""
    from oracle import dense, solvers
    from quantumsimulations_b200 import compile_table

    p = sweep_params(6, 50e3, t_final=window, steps=21)
    table = compile_table(p)
    bm = table.bit_major()
    H = dense.csr_from_table(bm["h"], bm["J"])
    psi0 = np.zeros(table.dim, complex)
    psi0[table.psi0_index] = 1.0
    t = np.linspace(0, p.t_final, p.steps)
    t0 = time.perf_counter()
    _, _, _, st = solvers.evolve_csr(H, psi0, t, [], backend="zvode", atol=p.solver_atol, rtol=p.solver_rtol,
                                     nsteps=p.solver_nsteps, max_step=p.solver_max_step)
    dt = time.perf_counter() - t0
    return {"steps_per_s_1core": st["n_steps"] / dt, "rhs_per_s_1core": st["n_rhs"] / dt, "window_s": window,
            "n_steps": st["n_steps"], "kind": "port (SciPy zvode-Adams, QuTiP's default integrator)"}


def _zvode_one(args):
    """One N=7 sweep point on the reference-equivalent integrator over a bounded window (worker of the all-cores leg)."""
    delta, variant, window = args
    from oracle import dense, solvers
    from quantumsimulations_b200 import compile_table

    p = sweep_params(6, delta, variant, t_final=window, steps=5)
    table = compile_table(p)
    bm = table.bit_major()
    H = dense.csr_from_table(bm["h"], bm["J"])
    psi0 = np.zeros(table.dim, complex)
    psi0[table.psi0_index] = 1.0
    _, _, _, st = solvers.evolve_csr(H, psi0, np.linspace(0, p.t_final, p.steps), [], backend="zvode", atol=p.solver_atol,
                                     rtol=p.solver_rtol, nsteps=p.solver_nsteps, max_step=p.solver_max_step)
    return st["n_steps"], st["n_rhs"]


def cpu_zvode_sweep_all_cores(window=1.0e-3):
    """SURVEY 8(d) CPU baseline (ii): the 39-trajectory C2 sweep as independent processes on ALL host cores, each
    trajectory integrated by SciPy zvode-Adams (what QuTiP drives) over a bounded window; the full-size wall is the
    window's wall scaled by 30 s / window."""
    cores = os.cpu_count() or 1
    env = dict(os.environ, OMP_NUM_THREADS="1", OPENBLAS_NUM_THREADS="1", MKL_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    t0 = time.perf_counter()
    procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--zvode-worker", f"{i}/{cores}/{window}"],
                              stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, env=env) for i in range(cores)]
    outs = [p.communicate()[0] for p in procs]
    wall = time.perf_counter() - t0
    steps, count = 0.0, 0
    for o in outs:
        for ln in o.strip().splitlines():
            if ln.startswith("{"):
                d = json.loads(ln)
                steps += d["steps"]
                count += d["trajectories"]
    return {"cores": cores, "trajectories": count, "window_s": window, "wall_s": wall, "steps": steps,
            "steps_per_s_all_cores": steps / wall, "extrapolated_full_sweep_wall_s": wall * 30.0 / window,
            "kind": "port (SciPy zvode-Adams, one process per host core, trajectories dealt round-robin; includes "
                    "interpreter start-up)"}


def zvode_worker(spec):
    i, cores, window = spec.split("/")
    dets = np.linspace(0.0, 3.0 * F1A, 13)
    jobs = [(d, v, float(window)) for d in dets for v in ("center_off", "center_on", "shell_off")][int(i)::int(cores)]
    res = [_zvode_one(j) for j in jobs]
    print(json.dumps({"trajectories": len(jobs), "steps": float(sum(r[0] for r in res))}), flush=True)


def c1_spin32_leg(device):
    """The sweep's N = 7 point (SW:1201-1252) with the J = 3/2 rare spin the reference declares (DE:22-25, DE:377) and
    cannot run: 8 index bits, rotated frame + Pauli-string observables."""
    import dataclasses

    from quantumsimulations_b200 import simulate_rare

    p = dataclasses.replace(sweep_params(6, 50e3, t_final=30.0, steps=20000), is_spin_three_half=True)
    simulate_rare(dataclasses.replace(p, t_final=3e-3, steps=300), device=device)          # warm-up
    t0 = time.perf_counter()
    _, obs, st = simulate_rare(p, device=device, return_stats=True, raw_norm=True)
    wall = time.perf_counter() - t0
    return {"workload": "C1 (i) with is_spin_three_half=True: n_sea=6 + a four-level rare site (8 index bits), delta=50 kHz, "
                        "rare ON, t_final=30 s, 20000 outputs, atol 1e-10 / rtol 1e-9",
            "method": "propagator columns + squarings as C1 (i), state-storing time-split chain, Pauli-string reduction "
                      "(qsb_propagate_expect); not an ODE step rate",
            "wall_s": wall, "kernel_ms": st.device_ms, "path": st.path, "column_h_applies": st.n_rhs,
            "max_raw_norm_deviation": float(np.max(np.abs(obs["raw_norm"] - 1.0)))}


def c1_leg(device, with_cpu):
    """BASELINE config 1: ONE trajectory at the reference's own sizes, through simulate_rare.
    (i) the sweep's N = 7 point (SW:1201-1252: 30 s, 20000 outputs); (ii) the dataclass-default N = 13 icosahedron
    (DE:332-384 with is_spin_three_half=False: 0.02 s, 2000 outputs, QuTiP default tolerances)."""
    from oracle import dense, solvers
    from quantumsimulations_b200 import DipolarRareParams, compile_table, simulate_rare

    out = {}
    p7 = sweep_params(6, 50e3, t_final=30.0, steps=20000)
    simulate_rare(sweep_params(6, 50e3, t_final=3e-3, steps=200), device=device)           # warm the context
    t0 = time.perf_counter()
    _, _, st = simulate_rare(p7, device=device, return_stats=True)
    w_auto = time.perf_counter() - t0
    # the stepped integrators on a 3 ms window of the same trajectory (the full 30 s is ~10^9 steps on ANY stepper)
    p7w = sweep_params(6, 50e3, t_final=3e-3, steps=200)
    legs = {}
    for m in ("taylor", "dp5"):
        t0 = time.perf_counter()
        _, _, sm = simulate_rare(p7w, device=device, method=m, return_stats=True)
        w = time.perf_counter() - t0
        legs[m] = {"wall_s": w, "kernel_ms": sm.device_ms, "steps": sm.n_accepted + sm.n_rejected, "h_applies": sm.n_rhs,
                   "steps_per_s": (sm.n_accepted + sm.n_rejected) / (sm.device_ms / 1e3),
                   "h_applies_per_s": sm.n_rhs / (sm.device_ms / 1e3)}
    out["n7_sweep_point"] = {
        "workload": "C1 (i): N=7, delta=50 kHz, rare ON, t_final=30 s, 20000 outputs, atol 1e-10 / rtol 1e-9 (SW:1201-1252)",
        "wall_s": w_auto, "kernel_ms": st.device_ms, "path": st.path, "column_h_applies": st.n_rhs,
        "method": "auto (propagator: not an ODE step rate)", "stepped_3ms_window": legs}
    try:
        out["n7_sweep_point_spin_three_half"] = c1_spin32_leg(device)
    except Exception as e:  # noqa: BLE001  a side leg must not take the bench line down
        out["n7_sweep_point_spin_three_half"] = {"error": repr(e)}
    p13 = DipolarRareParams(is_spin_three_half=False)
    simulate_rare(p13, device=device)                                                      # plan + dense buffers warm
    t0 = time.perf_counter()
    _, _, s13 = simulate_rare(p13, device=device, return_stats=True)
    w13 = time.perf_counter() - t0
    os.environ["QSB200_DENSE_OUTPUT"] = "0"          # round-1 behaviour: step ONTO each of the 2000 output times
    try:
        t0 = time.perf_counter()
        _, _, s13s = simulate_rare(p13, device=device, return_stats=True)
        w13s = time.perf_counter() - t0
    finally:
        del os.environ["QSB200_DENSE_OUTPUT"]
    t0 = time.perf_counter()
    _, _, s13d = simulate_rare(p13, device=device, method="dp5", return_stats=True)
    w13d = time.perf_counter() - t0
    out["n13_dataclass_default"] = {
        "workload": "C1 (ii): DipolarRareParams(is_spin_three_half=False): N=13 icosahedron, 0.02 s, 2000 outputs, "
                    "options None (QuTiP defaults atol 1e-8 / rtol 1e-6)",
        "method": "auto = Taylor(20) with dense output: unclipped steps, interior outputs read off the step's polynomial",
        "wall_s": w13, "kernel_ms": s13.device_ms, "path": s13.path, "outputs": p13.steps,
        "steps": s13.n_accepted + s13.n_rejected, "h_applies": s13.n_rhs, "outputs_per_s": p13.steps / w13,
        "steps_per_s": (s13.n_accepted + s13.n_rejected) / w13, "h_applies_per_s": s13.n_rhs / w13,
        "stepping_onto_every_output": {"wall_s": w13s, "steps": s13s.n_accepted + s13s.n_rejected, "h_applies": s13s.n_rhs,
                                       "steps_per_s": (s13s.n_accepted + s13s.n_rejected) / w13s,
                                       "h_applies_per_s": s13s.n_rhs / w13s},
        "dp5": {"wall_s": w13d, "steps": s13d.n_accepted + s13d.n_rejected, "h_applies": s13d.n_rhs,
                "steps_per_s": (s13d.n_accepted + s13d.n_rejected) / w13d}}
    # the sweep's parameters at the reference sweep FUNCTION's default size, n_sea = 12 (SW:356-376), full length:
    # "auto" takes the HBM-resident propagator there (stepping is ~10^9 H-applies)
    p13s = sweep_params(12, 50e3, t_final=30.0, steps=20000)
    t0 = time.perf_counter()
    _, o13s, s13p = simulate_rare(p13s, device=device, return_stats=True, raw_norm=True)
    w13p = time.perf_counter() - t0
    p13w = sweep_params(12, 50e3, t_final=3.0e-4, steps=3)
    t0 = time.perf_counter()
    _, _, s13w = simulate_rare(p13w, device=device, method="taylor", return_stats=True)
    w13w = time.perf_counter() - t0
    out["n13_sweep_point_full_size"] = {
        "workload": "N=13 (n_sea=12, the default of run_sweep_sea_detuning), delta=50 kHz, rare ON, t_final=30 s, 20000 outputs, "
                    "atol 1e-10 / rtol 1e-9",
        "method": "auto -> HBM-resident propagator (qsb_propagate_hbm): 2^13 columns integrated at once as a 26-qubit "
                  "tiled-path state, cuBLAS Zgemm squarings + time-split chain; not an ODE step rate",
        "wall_s": w13p, "kernel_ms": s13p.device_ms, "path": s13p.path, "column_h_applies": s13p.n_rhs,
        "max_raw_norm_deviation": float(np.max(np.abs(o13s["raw_norm"] - 1.0))),
        "stepped_3e-4s_window": {"wall_s": w13w, "steps": s13w.n_accepted + s13w.n_rejected, "h_applies": s13w.n_rhs,
                                 "steps_per_s": (s13w.n_accepted + s13w.n_rejected) / w13w,
                                 "extrapolated_full_trajectory_s": w13w * 30.0 / 3.0e-4}}
    del o13s
    if with_cpu:
        table = compile_table(p13)
        bm = table.bit_major()
        H = dense.csr_from_table(bm["h"], bm["J"])
        psi0 = np.zeros(table.dim, complex)
        psi0[table.psi0_index] = 1.0
        t = np.linspace(0, p13.t_final, p13.steps)
        t0 = time.perf_counter()
        _, _, _, zs = solvers.evolve_csr(H, psi0, t, [], backend="zvode", atol=1e-8, rtol=1e-6, nsteps=2500, max_step=0.0)
        wz = time.perf_counter() - t0
        out["n13_dataclass_default"]["cpu_zvode_1core"] = {"wall_s": wz, "steps": zs["n_steps"], "rhs": zs["n_rhs"],
                                                          "steps_per_s": zs["n_steps"] / wz}
        out["n7_sweep_point"]["cpu_zvode_1core"] = cpu_zvode_rate()
    return out


# --------------------------------------------------------------------------------------------------------
def workload_text(n, method):
    """config.workload shared by both arms (the driver compares the two lines on the same config)."""
    return (f"C4: single trajectory per GPU, N={n} spins (n_sea={n - 1}, Ga/Al constants, "
            f"delta=50 kHz + 12.5 kHz*rank, rare drive ON), {method} step attempts")


def bench_config(n, method, world):
    """config shared VERBATIM by both arms (the driver compares the two lines on the same config)."""
    return {"workload": workload_text(n, method), "n_qubits": n, "state_bytes": 16.0 * (1 << n), "method": method,
            "atol": 1e-10, "rtol": 1e-9, "l2": "inputs larger than L2 (4.3 GB vectors vs 126 MB)",
            "parallelism": f"independent sweep points x{world}"}


def run_reference(args, rank, world):
    """The reference arm: the CPU port of the CSR row product on all host cores, K timed samples after W warm-ups,
    sized so that the whole run stays within ~2.5 minutes (a complete H-apply per step when that fits)."""
    if rank != 0:
        return
    port = CpuPort(args.n_qubits - 1, 50e3)
    rows = port.rows_for(150.0 / max(1, args.steps))
    for _ in range(args.warmup):
        port.seconds(0, min(rows, 1 << 22))
    secs = [port.seconds((s * rows) % (1 << port.n), rows) for s in range(args.steps)]
    step_s = [6.0 * dt * (1 << port.n) / rows for dt in secs]        # 6 H-applies per DP5 step attempt
    value = len(step_s) / sum(step_s)
    base = {"value": value, "unit": "ODE steps/s", "cores": port.cores, "kind": "port",
            "sample": port.sample_text(rows, secs), "extrapolated": rows != 1 << port.n, "rows_timed": rows,
            "seconds_per_sample": secs}
    line = {
        "impl": "reference", "metric": "ODE steps/s", "value": value, "unit": "ODE steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 / value,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": bench_config(args.n_qubits, args.method, world),
        "arm": "CPU port of the reference's CSR row product (oracle/c), all host threads; QuTiP itself cannot hold the "
               "1.2 TB CSR of this operator (SURVEY 6.3)",
        "timing": "wall clock of the sampled rows, scaled to 6 complete H-applies per step attempt; measured time "
                  f"{sum(secs):.1f} s, claimed (scaled) time {sum(step_s):.1f} s",
        "cpu_baseline": base,
        "e2e": {"value": value, "unit": "ODE steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_b200(args, rank, world, local_rank):
    import torch

    from quantumsimulations_b200 import compile_table, engine, simulate_rare_batch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the engine has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist

        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"   # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n = args.n_qubits
    n_sea = n - 1
    delta = 50e3 + 12.5e3 * rank
    params = sweep_params(n_sea, delta)
    table = compile_table(params)
    opts = engine.SolverOptions(atol=1e-10, rtol=1e-9, max_step=1e-5, nsteps=0, method=args.method)
    S = 16.0 * (1 << n)
    n_rhs = 6 if args.method == "dp5" else 12

    plan = engine.Plan(table, local_rank)
    plan.use_torch_workspace(args.method, local_rank)
    plan.session_begin(opts)
    with ClockSampler(local_rank) as clk:   # started before the warm-up: nvidia-smi needs ~1 s to begin streaming
        plan.session_step(args.warmup)
        barrier()
        clk.mark()
        t_now, st, ms = plan.session_step(args.steps)
        clk.mark()
    barrier()
    ms_max = allmax(ms)
    value = world * args.steps / (ms_max / 1e3)
    obs = plan.session_observe()
    t_end = t_now

    line = None
    if rank == 0:
        # ---- per-kernel shares (separate, serialised pass: never part of the timed number) ----------------
        engine.profile(True)
        plan.session_step(2)
        kms, kcnt, kbytes = engine.profile_read(with_bytes=True)
        engine.profile(False)
        tot = sum(kms.values()) or 1.0
        kernels = {}
        for k in kms:
            if kcnt[k]:
                avg = kms[k] / kcnt[k]
                kernels[k] = {"launches_per_step": kcnt[k] / 2.0, "avg_ms": avg, "share": kms[k] / tot,
                              "moved_bytes_per_launch": kbytes[k] / kcnt[k],
                              "moved_GBps": kbytes[k] / (kms[k] * 1e-3) / 1e9}
        moved_per_step = sum(kbytes.values()) / 2.0
        dominant = max(kernels, key=lambda k: kernels[k]["share"]) if kernels else None
        peak, how = measured_peak()
        alg_bytes_step = (n_rhs + 1) * 2.0 * S
        achieved = alg_bytes_step * args.steps / (ms / 1e3) / 1e9
        prof = ncu_traffic()
        roofline = {
            "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": (prof or {}).get("dram_bytes_per_step"),
            "traffic_source": ("committed ncu --set full capture (profiles/ncu_summary.json" +
                               (f", tag {prof.get('tag')}" if prof and prof.get("tag") else "") +
                               "), not re-measured in this run") if prof else None,
            "peak_source": how,
            "kernel": "one DP5 step attempt = 6 H-applies (pass_first[+stage combination]/mid/centre/mid/last; stages 1-4 "
                      "end in pass_last_first, which also begins the next stage), the embedded error estimate folded "
                      "into the last stage's pass_first / pass_last",
            "algorithmic_bytes_per_step": alg_bytes_step, "moved_bytes_per_step": moved_per_step,
            "moved_GBps": moved_per_step * args.steps / (ms / 1e3) / 1e9,
            "moved_frac_of_peak": moved_per_step * args.steps / (ms / 1e3) / 1e9 / peak,
            "dominant_kernel": dominant, "kernels": kernels,
        }
        # FP64 co-bound: what the three-diagonal-form H-apply actually executes.  Per amplitude and
        # apply: one complex add per butterfly stage (2 flops) over first 24 + mid 2 x 2 x |g1| + centre 2 x 2 x |g2| +
        # last 24 stages, three diagonal forms (~10 flops each) and the epilogue (~10).
        stages = 48 + 4 * (n - 12)
        flops_apply = (2.0 * stages + 40.0) * (1 << n)
        fp64_rate = n_rhs * flops_apply * args.steps / (ms / 1e3) / 1e12
        fp64_peak = engine.fp64_peak_tflops(local_rank)
        roofline["fp64"] = {"flops_per_step": n_rhs * flops_apply, "achieved_TFLOPs": fp64_rate, "peak_TFLOPs": fp64_peak,
                            "peak_source": "measured in this run (qsb_debug_fp64_peak: DFMA chains on all SMs, under the "
                                           "same power cap); nominal B200 FP64 is 40",
                            "frac": fp64_rate / fp64_peak,
                            "binding": "hbm" if roofline["moved_frac_of_peak"] > fp64_rate / fp64_peak else "fp64"}
    plan.session_end()
    plan.close()

    # ---- end to end through the public call: plan creation, table upload, evolve, observables back --------
    barrier()
    # four output times over 8x the span the timed session covered (~8 (W + K) step attempts): a real trajectory is
    # thousands of steps long, so plan creation (a 47 GB cudaMalloc among other things) must not dominate, but the leg
    # has to stay under ten seconds
    t_out = np.linspace(0.0, 8.0 * t_end, 4)
    t0 = time.perf_counter()
    with engine.Plan(table, local_rank) as p2:
        obs2, st2 = p2.evolve(t_out, opts)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    e2e_s = allmax(e2e_s)
    attempts2 = st2.n_accepted + st2.n_rejected
    bm = table.bit_major()
    h2d = bm["h"].nbytes + bm["J"].nbytes + 3 * 4096 * 8 + t_out.nbytes
    d2h = obs2.nbytes + 10 * 8 * attempts2
    e2e = {"value": world * attempts2 / e2e_s, "unit": "ODE steps/s", "h2d_bytes_per_step": h2d / attempts2,
           "d2h_bytes_per_step": d2h / attempts2, "steps": attempts2, "seconds": e2e_s,
           "call": "engine.Plan(table).evolve(t_out, opts) [what simulate_rare() runs], host in / host out"}

    # ---- C2 sweep (CI-sized window), trajectories sharded round-robin over ranks, no collective ----------
    dets = np.linspace(0.0, 3.0 * F1A, 13)
    plist = [sweep_params(6, d, v) for d in dets for v in ("center_off", "center_on", "shell_off")]
    mine = plist[rank::world]
    barrier()
    t0 = time.perf_counter()
    res, sst = simulate_rare_batch(mine, device=local_rank, method="dp5") if mine else ([], [])
    sweep_s = allmax(time.perf_counter() - t0)
    barrier()
    t0 = time.perf_counter()
    _, tst = simulate_rare_batch(mine, device=local_rank, method="taylor") if mine else ([], [])   # Taylor(20), tol x1e-3
    sweep_default_s = allmax(time.perf_counter() - t0)
    sweep_steps = float(sum(s.n_accepted + s.n_rejected for s in sst))
    sweep_kernel_ms = max([s.device_ms for s in sst], default=0.0)
    if dist is not None:
        tt = torch.tensor([sweep_steps], dtype=torch.float64, device="cuda")
        dist.all_reduce(tt)
        sweep_steps = float(tt.item())
        sweep_kernel_ms = allmax(sweep_kernel_ms)

    # ---- C2 at the reference's FULL size (SW:1201-1252: t_final = 30 s, 20000 outputs) on the propagator path ----------
    full = [sweep_params(6, d, v, t_final=30.0, steps=20000) for d in dets for v in ("center_off", "center_on", "shell_off")]
    fmine = full[rank::world]
    barrier()
    t0 = time.perf_counter()
    _, pci = simulate_rare_batch(mine, device=local_rank) if mine else ([], [])   # API default: "auto" -> propagator
    sweep_prop_ci_s = allmax(time.perf_counter() - t0)
    barrier()
    t0 = time.perf_counter()
    fres, fst = simulate_rare_batch(fmine, device=local_rank, raw_norm=True) if fmine else ([], [])
    sweep_full_s = allmax(time.perf_counter() - t0)
    full_kernel_ms = allmax(max([s.device_ms for s in fst], default=0.0))
    full_rhs = float(sum(s.n_rhs for s in fst))
    full_norm_dev = max([float(np.max(np.abs(r[1]["raw_norm"] - 1.0))) for r in fres], default=0.0)
    del fres

    # ---- C3 grid (CI-sized window): 19 f1A x 13 detunings x 3 variants, cost-balanced over ranks, no collective ----
    grid = None
    if not args.no_grid:
        from quantumsimulations_b200.dist import estimate_cost, partition

        f1as = np.arange(5.0e3, 50.0e3 + 1.0, 2.5e3)   # the 5..50 kHz step 2.5 kHz grid 2D_sweep_report.py is drawn on
        glist = [sweep_params(6, d, v, f1a=f) for f in f1as for d in np.linspace(0.0, 3.0 * f, 13)
                 for v in ("center_off", "center_on", "shell_off")]
        gmine = [glist[i] for i in partition([estimate_cost(p) for p in glist], world)[rank]]
        barrier()
        t0 = time.perf_counter()
        _, gst = simulate_rare_batch(gmine, device=local_rank, method="taylor") if gmine else ([], [])
        grid_s = allmax(time.perf_counter() - t0)
        barrier()
        t0 = time.perf_counter()
        _ = simulate_rare_batch(gmine, device=local_rank) if gmine else None
        grid_default_s = allmax(time.perf_counter() - t0)
        gsteps = float(sum(s.n_accepted + s.n_rejected for s in gst))
        grhs = float(sum(s.n_rhs for s in gst))
        if dist is not None:
            tt = torch.tensor([gsteps, grhs], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt)
            gsteps, grhs = float(tt[0].item()), float(tt[1].item())
        # the same grid at the reference's full trajectory size, propagator path
        gfull = [sweep_params(6, d, v, f1a=f, t_final=30.0, steps=20000) for f in f1as
                 for d in np.linspace(0.0, 3.0 * f, 13) for v in ("center_off", "center_on", "shell_off")]
        gfmine = [gfull[i] for i in partition([estimate_cost(p) for p in gfull], world)[rank]]
        barrier()
        t0 = time.perf_counter()
        _, gfst = simulate_rare_batch(gfmine, device=local_rank) if gfmine else ([], [])
        grid_full_s = allmax(time.perf_counter() - t0)
        grid_full_kernel_ms = allmax(max([s.device_ms for s in gfst], default=0.0))
        grid = {"full_size": {"workload": "C3 at the reference's trajectory size (t_final = 30 s, 20000 outputs), "
                                          "API default method (auto -> propagator)", "trajectories": len(gfull),
                              "wall_s": grid_full_s, "kernel_ms": grid_full_kernel_ms},
                "wall_s_default_method": grid_default_s,
                "workload": "C3 (CI-sized): 19 f1A (5..50 kHz) x 13 detunings (0..3 f1A) x 3 variants, N=7, t_final=0.03 s, "
                            "2000 outputs, one CTA per trajectory, stepping Taylor(20) integrator (method='taylor')",
                "trajectories": len(glist), "wall_s": grid_s, "trajectories_per_s": len(glist) / grid_s,
                "step_attempts": gsteps, "h_applies_per_s": grhs / grid_s,
                "parallelism": f"longest-first partition over {world} GPU(s), no collective"}

    sharded = None
    if not args.no_sharded:
        barrier()
        torch.cuda.empty_cache()
        sharded = sharded_leg(args, rank, world, local_rank, barrier, allmax)

    if rank == 0:
        # launches of one step attempt, counted in the profiled pass (2 attempts) + the partial-sum reduction kernel
        launches_per_step = sum(kcnt.values()) / 2.0 + 1
        cpu = cpu_port_step_rate(n_sea, 50e3) if world == 1 and not args.no_cpu else None
        zv = cpu_zvode_rate() if world == 1 and not args.no_cpu else None
        zv_all = cpu_zvode_sweep_all_cores() if world == 1 and not args.no_cpu else None
        c1 = c1_leg(local_rank, world == 1 and not args.no_cpu) if world == 1 else None
        line = {
            "metric": "ODE steps/s", "value": value, "unit": "ODE steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": bench_config(n, args.method, world),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches_per_step * args.steps),
            "clocks": clk.summary(),
            "accepted": st.n_accepted, "rejected": st.n_rejected, "rhs": st.n_rhs, "t_reached": t_end,
            "raw_norm": float(obs[7]),
            "sweep": {"workload": "C2 (CI-sized): 13 detunings x 3 variants, N=7, t_final=0.03 s, 2000 outputs, "
                                  "atol 1e-10 / rtol 1e-9, one CTA per trajectory",
                      "trajectories": len(plist), "wall_s": sweep_s, "kernel_ms": sweep_kernel_ms, "method": "dp5",
                      "steps_per_s": sweep_steps / sweep_s, "total_steps": sweep_steps,
                      "wall_s_taylor": sweep_default_s,
                      "taylor": "stepping Taylor order 20, atol/rtol x 1e-3 (method='taylor')",
                      "cpu_zvode": zv, "cpu_zvode_all_cores": zv_all,
                      "wall_s_default_method": sweep_prop_ci_s,
                      "default_method": "auto = Taylor(20) columns of exp(-iH dt/2^s), s on-chip squarings, one matvec per "
                                        "output (what simulate_rare runs when steps - 1 > 2^N)",
                      "full_size": {
                          "workload": "C2 at the reference's own size (SW:1201-1252): 13 detunings x 3 variants, N=7, "
                                      "t_final=30 s, 20000 outputs, atol 1e-10 / rtol 1e-9, max_step 1e-5; API default "
                                      "method (auto -> propagator: 128 Taylor(20) column trajectories over dt/2^s, s = "
                                      "8..10 on-chip squarings, then 19999 on-chip matvecs per trajectory)",
                          "trajectories": len(full), "wall_s": sweep_full_s, "kernel_ms": full_kernel_ms,
                          "column_h_applies": full_rhs, "max_raw_norm_deviation": full_norm_dev,
                          "cpu_zvode_extrapolated_s_per_trajectory_1core":
                              (30.0 / zv["window_s"] * zv["n_steps"] / zv["steps_per_s_1core"]) if zv else None}},
            "grid": grid,
            "c1": c1,
            "sharded": sharded,
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def _product_parity(plan, table, rank, seed, count=1 << 16):
    """max rel. error of `count` sampled rows of H psi (product state, one H-apply through the plan's kernels) against
    the closed form of oracle/product.py; rows are drawn from this rank's slice."""
    from oracle import product

    bm = table.bit_major()
    phi = product.random_phi(table.n, seed)
    rng = np.random.default_rng(1000 + rank)
    rows = rng.integers(0, plan.local_dim, size=count, dtype=np.uint64)
    rows[:2] = (0, plan.local_dim - 1)
    got = plan.apply_h_product(phi, rows)
    want = product.apply_h_rows(bm["h"], bm["J"], phi, rows + np.uint64(rank * plan.local_dim))
    return float(np.max(np.abs(got - want)) / np.max(np.abs(want)))


def _time_sharded(table, world, local_rank, opts, attempts, barrier, allmax, parity_seed=None, rank=0):
    from quantumsimulations_b200 import engine, shard

    plan = shard.sharded_plan_from_env(table, local_rank) if world > 1 else engine.Plan(table, local_rank)
    out = {"local_dim": plan.local_dim}
    plan.session_begin(opts)
    plan.session_step(1)
    barrier()
    _, st0, _ = plan.session_step(0)
    t_now, st, ms = plan.session_step(attempts)
    barrier()
    out["ms"] = allmax(ms)
    out["n_rhs"] = st.n_rhs - st0.n_rhs
    out["t_reached"] = t_now
    out["obs"] = plan.session_observe()
    engine.profile(True)
    plan.session_step(1)
    kms, kcnt = engine.profile_read()
    engine.profile(False)
    out["kernel_ms"] = {k: (kms[k] / kcnt[k] if kcnt[k] else None) for k in kms}
    n_apply = max(kcnt["pass_last"], 1)
    out["centre_ms"] = allmax(kms["pass_centre"] / n_apply)   # per H-apply (one launch, or one per branch when overlapped)
    out["kernel_ms"]["pass_centre_per_happly"] = kms["pass_centre"] / n_apply
    out["overlap"] = os.environ.get("QSB200_SHARD_OVERLAP", "1") != "0" and world > 1
    local_ms = sum(kms[k] for k in kms if k != "pass_centre") / max(kcnt["pass_last"], 1)
    out["hbm_passes_ms_per_happly"] = allmax(local_ms)
    plan.session_end()
    if parity_seed is not None:
        out["parity"] = allmax(_product_parity(plan, table, rank, parity_seed))
    plan.close()
    barrier()
    return out


def sharded_leg(args, rank, world, local_rank, barrier, allmax):
    """C5: ONE trajectory whose state vector is sharded over all ranks (top log2(world) index bits = rank), Taylor
    step attempts.  world == 1 gives the single-GPU anchor at the same per-GPU shard size.  Weak scaling at
    --shard-local-bits per GPU, a strong-scaling point at N = 30, and two parity fields: product-state rows of one
    H-apply against the closed form at the benchmarked size, and an N = 26 trajectory against the unsharded plan.
    Returns a dict on rank 0."""
    import torch

    from quantumsimulations_b200 import compile_table, engine, shard

    wl = world.bit_length() - 1
    n = args.shard_local_bits + wl
    table = compile_table(sweep_params(n - 1, 50e3))
    order = 8
    opts = engine.SolverOptions(atol=1e-10, rtol=1e-9, max_step=1e-5, nsteps=0, method="taylor", order=order)
    attempts = 2
    w = _time_sharded(table, world, local_rank, opts, attempts, barrier, allmax, parity_seed=n, rank=rank)
    torch.cuda.empty_cache()
    # strong-scaling point: N = 30 on every world size (the 1-GPU anchor is the weak leg when local bits = 30)
    n_strong = 30
    strong = None
    if world > 1 and n_strong - wl >= 12 and n_strong != n:
        st_table = compile_table(sweep_params(n_strong - 1, 50e3))
        sres = _time_sharded(st_table, world, local_rank, opts, attempts, barrier, allmax, parity_seed=n_strong, rank=rank)
        strong = {"n_qubits": n_strong, "ms_per_happly": sres["ms"] / sres["n_rhs"], "centre_ms": sres["centre_ms"],
                  "hbm_passes_ms_per_happly": sres["hbm_passes_ms_per_happly"], "parity_max_rel_err": sres["parity"]}
        torch.cuda.empty_cache()
    elif n_strong == n:
        strong = {"n_qubits": n_strong, "ms_per_happly": w["ms"] / w["n_rhs"], "note": "same run as the weak leg"}
    # sharded vs unsharded trajectory at N = 26 (every rank runs the sharded one; rank 0 repeats it on one plan)
    vs = None
    if world > 1:
        n_cmp = 26
        ctable = compile_table(sweep_params(n_cmp - 1, 50e3))
        t_out = np.linspace(0.0, 2.0e-6, 3)
        copts = engine.SolverOptions(atol=1e-10, rtol=1e-9, max_step=1e-5, nsteps=100000, method="taylor", order=order)
        plan = shard.sharded_plan_from_env(ctable, local_rank)
        sobs, sst = plan.evolve(t_out, copts)
        plan.close()
        barrier()
        if rank == 0:
            with engine.Plan(ctable, local_rank) as up:
                uobs, ust = up.evolve(t_out, copts)
            vs = {"n_qubits": n_cmp, "outputs": len(t_out), "max_abs_diff": float(np.max(np.abs(sobs - uobs))),
                  "accepted": [sst.n_accepted, ust.n_accepted], "rejected": [sst.n_rejected, ust.n_rejected],
                  "what": "all 8 observable columns of qsb_evolve, sharded over the ranks vs one unsharded plan on rank 0"}
        barrier()
        torch.cuda.empty_cache()
    if rank != 0:
        return None
    peak, how = measured_peak()
    s_loc = 16.0 * w["local_dim"]
    ms, n_rhs = w["ms"], w["n_rhs"]
    out = {
        "workload": f"C5: one trajectory, N={n} spins sharded over {world} GPU(s) ({args.shard_local_bits} local index bits, "
                    f"{s_loc / 1e9:.1f} GB per vector per GPU), Taylor({order}) step attempts",
        "n_qubits": n, "world": world, "method": f"taylor{order}", "attempts": attempts, "n_rhs": n_rhs,
        "steps_per_s": attempts / (ms / 1e3), "ms_per_happly": ms / n_rhs,
        "layouts": engine.tile_layouts(n, world),
        "hbm": {"algorithmic_GBps_per_gpu": (n_rhs + attempts) * 2.0 * s_loc / (ms / 1e3) / 1e9, "peak": peak, "peak_source": how},
        "kernel_ms": w["kernel_ms"], "hbm_passes_ms_per_happly": w["hbm_passes_ms_per_happly"],
        "overlap": ("centre pass of one branch on a low-priority stream under the HBM passes of the other, stream-ordered "
                    "NCCL barriers (kernel_ms are serialised per-kernel event times, ms_per_happly is the overlapped run)")
                   if w["overlap"] else "off (serialised passes, host-side barriers)",
        "raw_norm": float(w["obs"][7]), "Iz_sea": float(w["obs"][2]), "t_reached": w["t_reached"],
        "parity_max_rel_err": w["parity"],
        "parity": "rows of H psi on a product state (2^16 sampled rows per rank, one H-apply through the sharded kernels) "
                  "vs the closed form of oracle/product.py; max over ranks",
        "vs_unsharded": vs, "strong": strong,
    }
    out["hbm"]["frac"] = out["hbm"]["algorithmic_GBps_per_gpu"] / peak
    if world > 1:
        per_dir = 4.0 * (world - 1) / world * s_loc   # x and y branch, loads by me + stores by the peers (and vice versa)
        centre_ms = w["centre_ms"]
        out["nvlink"] = {"bytes_per_happly_per_gpu_per_direction": per_dir, "centre_ms": centre_ms,
                         "achieved_GBps_per_direction": per_dir / (centre_ms / 1e3) / 1e9, "peak": 770.0,
                         "peak_source": "B200_PROFILING.md measured peer copy, per direction",
                         "frac": per_dir / (centre_ms / 1e3) / 1e9 / 770.0,
                         "kernel": "pass_centre<sharded>: W_g D W_g on tiles gathered from / scattered to the peers' scratch (P2P ld/st)"}
    return out


def kernels_per_apply(n):
    rest = n - 12
    groups = (rest + 7) // 8 if rest > 0 else 1
    return 2 + 1 + 2 * (groups - 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n-qubits", type=int, default=28)
    ap.add_argument("--method", default="dp5", choices=["dp5", "taylor"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline legs")
    ap.add_argument("--no-sharded", action="store_true", help="skip the sharded-state (C5) leg")
    ap.add_argument("--no-grid", action="store_true", help="skip the 2D-grid (C3) leg")
    ap.add_argument("--shard-local-bits", type=int, default=30, help="index bits per GPU in the sharded leg (30: 17 GB vectors)")
    ap.add_argument("--zvode-worker", default=None, help=argparse.SUPPRESS)
    args = ap.parse_args()
    if args.zvode_worker:
        return zvode_worker(args.zvode_worker)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
