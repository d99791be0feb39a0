#!/bin/bash
mkdir -p gpurun_out
timeout -s KILL 900 python -m pytest tests/test_gpu_propagator.py -x -q -k "hbm" > gpurun_out/r2s_pytest_hbm.log 2>&1; tail -5 gpurun_out/r2s_pytest_hbm.log
timeout -s KILL 600 python scripts/time_propagator_hbm.py 13 > gpurun_out/r2s_time_hbm_n14.jsonl 2> gpurun_out/r2s_time_hbm_n14.err; cat gpurun_out/r2s_time_hbm_n14.jsonl; tail -3 gpurun_out/r2s_time_hbm_n14.err
python - <<'PY' > gpurun_out/r2s_sweep_n13.json 2> gpurun_out/r2s_sweep_n13.err
import json, os, sys, tempfile, time
sys.path.insert(0, os.getcwd())
import numpy as np
from quantumsimulations_b200.sweep import run_sweep_sea_detuning
G_SEA, G_RARE, B0_T, F1A = 8.1812e7, 6.976e7, 3.0, 50_000.0
with tempfile.TemporaryDirectory() as tmp:
    t0 = time.perf_counter()
    base = run_sweep_sea_detuning(f_Az=G_SEA * B0_T / (2 * np.pi), f1A=F1A, target_sea_detuning=F1A, gamma_sea=G_SEA, gamma_rare=G_RARE,
                                  sea_detunings_Hz=[0.0, 50e3, 100e3], n_sea=12, t_final=30.0, steps=20000, phi_sea=np.pi / 2, phi_rare=np.pi / 2,
                                  out_root=tmp, solver_atol=1e-10, solver_rtol=1e-9, solver_nsteps=10_000_000, solver_max_step=1e-5,
                                  coarse_window=100, verbose=False)
    wall = time.perf_counter() - t0
    rows = json.load(open(os.path.join(base, "summary.json")))["sweep_results"]
print(json.dumps({"what": "run_sweep_sea_detuning at its default n_sea=12 (N=13), 3 detunings x 3 variants, 30 s / 20000 outputs", "wall_s": wall,
                  "rows": [{k: r[k] for k in ("delta_Hz", "I_z_slope_off_center", "I_z_slope_on_center", "contrast_rare_center", "DeltaOmega_over_geff")} for r in rows]}))
PY
cat gpurun_out/r2s_sweep_n13.json; tail -3 gpurun_out/r2s_sweep_n13.err
