"""Timing of laserDTRM::calculate() on the device against the CPU oracle, outside bench.py
(whose one JSON line is the surface-force + MULES stage).  Lives under tests/ because it runs the oracle as the
checker and CPU baseline; not collected by pytest.  Usage on a B200:
    python tests/laser_timing.py [--n 200] [--nr 100] [--ntheta 200] [--reps 5] > gpurun_out/laser.json
The workload: BASELINE configs[3] (3-phase melt pool box, n^3 cells), a Gaussian beam of nr x ntheta rays pointing
down onto the first metal blob, FresnelLaser reflection at both metal surfaces, local-density absorption."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import oracle as orc, mesh_oracle as mo                      # noqa: E402  (the checker / CPU baseline)
from thermocapillaryinterfoam_b200 import abi, cases, mesh as hm         # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=200)
    ap.add_argument("--nr", type=int, default=100)
    ap.add_argument("--ntheta", type=int, default=200)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    spec = cases.case(4, a.n)
    m = mo.box_mesh(*spec.n, spec.lengths, spec.patch_types)
    xs, ys, zs = (hm.vertices(spec.n[d], spec.lengths[d]) for d in range(3))
    geom = cases.box_polymesh(m, xs, ys, zs)
    f = cases.synthetic_fields(spec, m)
    models = [dict(kind="FresnelLaser", **{"from": 2, "to": 0}, epsilon=0.25),
              dict(kind="FresnelLaser", **{"from": 2, "to": 1}, epsilon=0.4)]
    aC, eC, EC = (200.0, 150.0, 0.5), (0.0, 0.0, 0.0), (0.0, 0.0, 0.0)
    d = np.array([0.15, -1.0, 0.05]); d /= np.linalg.norm(d)
    r = np.cross(d, [0, 0, 1.0]); r /= np.linalg.norm(r)
    p0, p1, Ip, dA = cases.laser_seeds((0.45, 0.95, 0.5), d, r, a.nr, a.ntheta, 0.15, geom["boundsMag"], sigma=0.06, power=500.0)
    cell = cases.find_cells_box(spec.n, xs, ys, zs, p0)
    ctx = abi.Context(m, spec.nPhases)
    cases.load_case(ctx, spec, f)
    fc = np.asarray(m.faceCells, dtype=np.int64)
    ctx.set_cell_centres(m.C, np.asarray(m.Cfb) - np.asarray(m.C)[fc])
    t0 = time.perf_counter()
    ctx.laser_set_polymesh(m.Cf, m.Cfb, geom)
    t_setup = time.perf_counter() - t0
    ctx.laser_set_reflection(models)
    ctx.laser_set_local_density_absorption([0, 1, 2], aC, eC, EC)
    Q, nPart = ctx.laser_calculate(p0, p1, Ip, dA, cell)          # warm-up
    ctx.timers_enable(True)
    ctx.timers_reset()
    l0 = ctx.launch_count
    t0 = time.perf_counter()
    for _ in range(a.reps):
        Q, nPart = ctx.laser_calculate(p0, p1, Ip, dA, cell)
    wall = (time.perf_counter() - t0) / a.reps
    tm = ctx.timers_read()
    launches = (ctx.launch_count - l0) / a.reps
    t0 = time.perf_counter()
    for _ in range(a.reps):
        _, _ = ctx.laser_calculate(p0, p1, Ip, dA, cell, dense=False)
        cs, vs = ctx.laser_get_sparse(1 << 20)
    wall_sparse = (time.perf_counter() - t0) / a.reps
    Qs = np.zeros(m.nCells); Qs[cs] = vs
    out = dict(what="laserDTRM::calculate()", cells=int(m.nCells), seeds=int(len(Ip)), particles=int(nPart),
               generations=int(ctx.laser_generations), wall_ms_per_call=1e3 * wall,
               wall_ms_per_call_sparse_readback=1e3 * wall_sparse, cells_with_energy=int(len(cs)),
               sparse_vs_dense_max_dev_rel=float(np.abs(Qs - Q).max() / np.abs(Q).max()),
               device_ms_per_call={k: v["ms"] / a.reps for k, v in tm.items() if v["ms"] > 0},
               launches_per_call=launches, polymesh_setup_s=t_setup,
               absorbed_W=float((Q * m.V).sum()), beam_W=float((Ip * dA)[cell >= 0].sum()))
    if not a.no_cpu:
        al = f["alpha_i"]
        aa, ee, EE = (orc.local_density(al, c) for c in (aC, eC, EC))
        h, gh = orc.MeshHandle(m), orc.LaserGeomHandle(m, geom)
        t0 = time.perf_counter()
        ref = orc.laser_calculate(h, gh, models, al, f["alpha_b"], f["alphaBc"], aa, ee, EE, f["T_i"], p0, p1, Ip, dA, cell)
        out["cpu_oracle_ms_per_call"] = 1e3 * (time.perf_counter() - t0)
        out["cpu_cores"] = 1
        scale = np.abs(ref["Q"]).max()
        out["max_err_vs_oracle_rel"] = float(np.abs(Q - ref["Q"]).max() / scale)
        out["particles_oracle"] = int(ref["nParticles"])
    ctx.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
