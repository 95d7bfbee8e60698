#!/usr/bin/env python
"""bench.py -- surface-force + MULES Mcell-updates/s (BASELINE.json metric) on N B200s of one node.

One "step" = one stage: divCapillaryStress over all phase pairs + one solveAlphas
corrector (limit x P, limitSum, explicitSolve x P) on frozen synthetic fields of BASELINE.json
configs[4] (synthetic 3-D 3-phase box, 64 M cells; strong scaling: the SAME box split over N ranks
by the `simple` block decomposition, processor-patch halo exchange over NCCL).

  value : stages timed with CUDA events on the launching stream, inputs resident in HBM
  e2e   : the same stage through tcif_stage_host (host buffers in pinned memory; H2D of every
          input field and D2H of every output field inside the timed region)
  roofline / cpu_baseline : see DESIGN.md "Measurement"
  --impl reference : the CPU oracle (restatement of the reference path, oracle/) domain-decomposed
          on all host cores -- the reference's own OpenFOAM binary cannot be built here.
"""
from __future__ import annotations

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

METRIC = "surface-force+MULES Mcell-updates/s"
UNIT = "Mcell-updates/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=5)
    ap.add_argument("--n", "--edge", dest="n", type=int, default=0, help="cells per box edge (0 = BASELINE size); use --edge under torchrun, whose parser claims --n")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong: the SAME box split over N ranks (default, BASELINE configs[4]); weak: --n (default 200) cells "
                         "per edge PER RANK, the box grows with the simple (px py pz) layout")
    ap.add_argument("--cpu-n", type=int, default=0, help="edge of the CPU-baseline sample box (0 = auto)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


# ---- contract algorithmic bytes (SURVEY.md 8(d) / BASELINE.md section 3) ---------------------
def contract_bytes(nC, nIF, nBF, P, QT, L):
    BS = nC * (64 * P + 8 * QT + 40) + 80 * nIF + nBF * (56 + 40 * P + 8 * QT)
    BM = P * (nIF * (136 + 96 * L) + nC * (184 + 128 * L))
    return BS + BM


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []
        self.marks = {}

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append((time.time(), ln.strip()))

    def mark(self, name):
        """wall-clock marks: 'load' (work starts, warm-up included), 't0'/'t1' (timed region)"""
        self.marks[name] = time.time()

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        t0, t1 = self.marks.get("t0", 0.0), self.marks.get("t1", float("inf"))
        lines = [ln for (t, ln) in self.lines if t0 <= t <= t1 + 0.05]
        window = "timed region"
        if len(lines) < 2:      # timed region shorter than the sampling period: widen to the warm-up steps before it
            tl = self.marks.get("load", 0.0)
            lines = [ln for (t, ln) in self.lines if tl <= t <= t1 + 0.05]
            window = "warm-up + timed region (timed region shorter than the sampling period)"
        self.window = window
        for ln in lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if f[5 + k].lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "power_w_max": float(max(pw)), "samples": len(sm), "window": window}


# ---- CPU arm: the oracle, domain-decomposed on the host cores ----------------------------------
def cpu_layout(cores):
    nd = 1
    while nd * 2 <= min(cores, 64):
        nd *= 2
    procs = [1, 1, 1]
    d, k = 0, nd
    while k > 1:
        procs[d % 3] *= 2
        k //= 2
        d += 1
    return nd, tuple(procs)


def _mem_available_gb():
    try:
        with open("/proc/meminfo") as fh:
            for ln in fh:
                if ln.startswith("MemAvailable:"):
                    return int(ln.split()[1]) / 1e6
    except OSError:
        pass
    return 0.0


def cpu_stage_rate(config, n_edge, nStages, cores, warmup=0):
    """Mcell-updates/s of the CPU oracle on an n_edge^3 box of the workload, one thread per sub-domain (each worker
    first-touches its own sub-mesh and fields, oracle/tcif_oracle.c thr_main).  `warmup` untimed stages first."""
    from concurrent.futures import ThreadPoolExecutor
    from oracle import oracle as orc
    from thermocapillaryinterfoam_b200 import cases, mesh as hm
    spec = cases.case(config, n_edge)
    nd, procs = cpu_layout(cores)

    def block(r):
        m = hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, r)
        return m, cases.synthetic_fields(spec, m)
    with ThreadPoolExecutor(max_workers=min(nd, 8)) as ex:
        built = list(ex.map(block, range(nd)))
    meshes, fs = [b[0] for b in built], [b[1] for b in built]

    def run(n):
        return orc.stage_threads(meshes, orc.mules_cfg(3), [f["alpha_i"] for f in fs], [f["alpha_b"] for f in fs],
                                 [f["alphaBc"] for f in fs], spec.sigma, [f["T_i"] for f in fs], [f["T_b"] for f in fs],
                                 [f["phi_i"] for f in fs], [f["phi_b"] for f in fs], [f["phiAlpha_i"] for f in fs],
                                 [f["phiAlpha_b"] for f in fs], spec.deltaT, nStages=n)[0]
    if warmup > 0:
        run(warmup)
    secs = run(nStages)
    return spec.nCells * nStages / secs / 1e6, secs, nd, procs, spec


def cpu_one_core_rate(config, n_edge=96, nStages=2):
    """The serial oracle (one core, no decomposition) on a small box of the same workload."""
    from oracle import oracle as orc
    from thermocapillaryinterfoam_b200 import cases, mesh as hm
    spec = cases.case(config, n_edge)
    m = hm.box_block(spec.n, spec.lengths, spec.patch_types, (1, 1, 1), 0)
    f = cases.synthetic_fields(spec, m)
    t0 = time.perf_counter()
    for _ in range(nStages):
        orc.stage(m, orc.mules_cfg(3), f["alpha_i"], f["alpha_b"], f["alphaBc"], f["sigma"], f["T_i"], f["T_b"],
                  f["phi_i"], f["phi_b"], f["phiAlpha_i"], f["phiAlpha_b"], f["deltaT"])
    secs = time.perf_counter() - t0
    return spec.nCells * nStages / secs / 1e6, f"{spec.name} {n_edge}^3 = {spec.nCells} cells, {nStages} stages, serial, {secs:.1f} s"


def cpu_sample_edge(args, cores):
    if args.cpu_n:
        return args.cpu_n
    full = args.n or {1: 64, 2: 512, 3: 128, 4: 200, 5: 400}[args.config]
    return min(full, 160 if cores >= 8 else 96)      # ~10-20 s of CPU work at 16 stages on 16 cores


def reference_edge(args, cores):
    """--impl reference runs the workload it names: the full box when the host can hold it (configs[4] at 400^3 needs
    ~70 GB for 16 sub-meshes, their fields, the workers' first-touch copies and the oracle's temporaries), else the
    bounded sample -- and then the line's config.workload names the SAMPLE, not the full box."""
    full = args.n or {1: 64, 2: 512, 3: 128, 4: 200, 5: 400}[args.config]
    if args.cpu_n:
        return args.cpu_n, full
    need_gb = 1.1e-6 * full ** 3 + 4
    if _mem_available_gb() >= need_gb and cores >= 8:
        return full, full
    return cpu_sample_edge(args, cores), full


def run_reference(args, rank):
    if rank != 0:
        return
    from thermocapillaryinterfoam_b200 import cases
    cores = os.cpu_count() or 1
    edge, full = reference_edge(args, cores)
    # the run must end within a few minutes: one untimed stage on the chosen box first; if W + K stages of it would not
    # fit the budget (TCIF_REF_BUDGET_S, default 720 s: the driver allows the arm 1800 s) the bounded sample is run instead -- and named as such
    budget = float(os.environ.get("TCIF_REF_BUDGET_S", "720"))
    if edge == full and not args.cpu_n:
        t0 = time.time()
        _, s1, *_ = cpu_stage_rate(args.config, edge, 1, cores)
        setup = time.time() - t0 - s1
        if 2 * setup + s1 * (args.steps + args.warmup + 1) > budget:
            edge = cpu_sample_edge(args, cores)
    rate, secs, nd, procs, spec = cpu_stage_rate(args.config, edge, args.steps, cores, warmup=args.warmup)
    one_core, one_core_sample = cpu_one_core_rate(args.config)
    sample = (f"{spec.name} {edge}^3 = {spec.nCells} cells, {spec.nPhases} phases, {args.steps} stages after {args.warmup} warm-up "
              f"stages, {nd} sub-domains {procs[0]}x{procs[1]}x{procs[2]} (one thread each, first-touch local copies)")
    wl = (f"configs[{args.config - 1}] {spec.name} {spec.n[0]}x{spec.n[1]}x{spec.n[2]} = {spec.nCells} cells, "
          f"{spec.nPhases} phases")
    if edge != full:
        wl += f" (SAMPLE of the {full}^3 box: host memory or cores too small for the full one)"
    line = {"impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": wl, "sample": sample},
            "cpu_baseline": {"value": rate, "unit": UNIT, "cores": nd, "kind": "port", "sample": sample,
                             "host_cores": cores, "one_core": {"value": one_core, "unit": UNIT, "sample": one_core_sample}},
            "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "CPU oracle (restatement of the reference path; the OpenFOAM v2106 binary is unbuildable here)"}
    print(json.dumps(line), flush=True)


def workload_name(args):
    from thermocapillaryinterfoam_b200 import cases
    spec = cases.case(args.config, args.n or None)
    return f"configs[{args.config - 1}] {spec.name} {spec.n[0]}x{spec.n[1]}x{spec.n[2]} = {spec.nCells} cells, {spec.nPhases} phases"


# ---- GPU arm -----------------------------------------------------------------------------------
def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from thermocapillaryinterfoam_b200 import abi, cases, mesh as hm

    if not torch.cuda.is_available() or abi.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device -- the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    procs = hm.procs_for(world)
    if args.scaling == "weak":
        import dataclasses
        base = args.n or 200
        spec = cases.case(args.config, base)
        if spec.two_d:
            raise SystemExit("bench.py: --scaling weak is for the 3-D configs")
        spec = dataclasses.replace(spec, n=(base * procs[0], base * procs[1], base * procs[2]))
    else:
        spec = cases.case(args.config, args.n or None)
    t0 = time.time()
    m = hm.box_block(spec.n, spec.lengths, spec.patch_types, procs, rank)
    f = cases.synthetic_fields(spec, m)
    nC, nIF, nBF = m.nCells, m.nIF, m.nBF
    ctx = abi.Context(m, spec.nPhases, device=local_rank)
    m.close()
    del m
    if world > 1:
        ids = [abi.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init_nccl(rank, world, ids[0])
    stream = torch.cuda.Stream(device=dev)
    ctx.set_stream(stream.cuda_stream)
    cases.load_case(ctx, spec, f)
    dt = f["deltaT"]
    setup_s = time.time() - t0
    P = spec.nPhases
    QT = sum(1 for s in spec.sigma if s[0] != "constant")

    # ---- value: device-resident stages --------------------------------------------------------
    with torch.cuda.stream(stream):
        sampler = ClockSampler(local_rank)
        if rank == 0:
            sampler.start()
            time.sleep(0.3)            # let nvidia-smi come up before the GPU gets busy
        sampler.mark("load")
        for _ in range(args.warmup):
            ctx.stage(dt)
        launches0 = ctx.launch_count
        barrier()
        sampler.mark("t0")
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):                 # the timed region: per-kernel timers OFF (no event records inside)
            ctx.stage(dt)
        e1.record(stream)
        barrier()
        sampler.mark("t1")
        clocks = sampler.stop() if rank == 0 else None
        ms = max_over_ranks(e0.elapsed_time(e1))
        launches = ctx.launch_count - launches0
        # per-kernel breakdown (roofline of the dominant kernel): a SECOND loop with the CUDA-event timers on
        ksteps = max(1, min(args.steps, 5))
        ctx.timers_enable(True)
        ctx.timers_reset()
        for _ in range(ksteps):
            ctx.stage(dt)
        timers = ctx.timers_read()
        ctx.timers_enable(False)
        barrier()
    total_cells = spec.nCells
    value = total_cells * args.steps / (ms * 1e-3) / 1e6

    # ---- roofline: dominant kernel, CUDA-event time on the launching stream -------------------
    hbm, how = peaks()
    vol = {k: v for k, v in timers.items() if v["launches"] and v["bytes"] > 0 and k not in ("halo", "convert", "stats")}
    top = max(vol, key=lambda k: vol[k]["ms"])
    tk = vol[top]
    achieved = tk["bytes"] / (tk["ms"] * 1e-3) / 1e9
    step_ms_sum = sum(v["ms"] for v in timers.values())
    stage_bytes_compulsory = sum(v["bytes"] for k, v in timers.items() if k not in ("convert", "stats")) / ksteps
    contract = contract_bytes(nC, nIF, nBF, P, QT, 3)
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "ncu_traffic.json")      # written by profiles/summarise_ncu.py from the committed capture
    if os.path.exists(tp) and world == 1 and not args.n and args.config == 5 and args.scaling == "strong":
        with open(tp) as fh:
            tj = json.load(fh)
        traffic = tj.get("bytes_per_launch", {}).get(top)
        traffic_src = tj.get("source")
    roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": how, "launches": tk["launches"],
                "avg_launch_ms": tk["ms"] / tk["launches"], "kernel_share_of_step": tk["ms"] / step_ms_sum,
                "bytes_per_launch": tk["bytes"] / tk["launches"],
                "stage": {"compulsory_bytes_per_cell": stage_bytes_compulsory / nC,
                          "achieved_GBs_compulsory": stage_bytes_compulsory * args.steps / (ms * 1e-3) / 1e9,
                          "contract_bytes_per_cell": contract / nC,
                          "achieved_GBs_contract": contract * args.steps / (ms * 1e-3) / 1e9,
                          "frac_contract": contract * args.steps / (ms * 1e-3) / 1e9 / hbm},
                "per_kernel_ms_per_step": {k: v["ms"] / ksteps for k, v in timers.items() if v["launches"]},
                "per_kernel_note": f"second loop of {ksteps} stages with the per-kernel CUDA-event timers on; `value` is timed with them off"}

    # ---- e2e: host buffers through tcif_stage_host ------------------------------------------
    e2e = None
    extra = {}
    if not args.no_e2e:
        def pin(a):
            b = abi.pinned_empty(a.shape)
            b[...] = a
            return b
        ai0 = [np.array(a) for a in f["alpha_i"]]; ab0 = [np.array(a) for a in f["alpha_b"]]
        ai = [pin(a) for a in ai0]; ab = [pin(a) for a in ab0]
        pi0 = [pin(a) for a in f["phiAlpha_i"]]; pb0 = [pin(a) for a in f["phiAlpha_b"]]
        pi = [abi.pinned_empty(a.shape) for a in pi0]; pb = [abi.pinned_empty(a.shape) for a in pb0]
        Ti, Tb, phii, phib = pin(f["T_i"]), pin(f["T_b"]), pin(f["phi_i"]), pin(f["phi_b"])
        Fi, Fb = abi.pinned_empty((nC, 3)), abi.pinned_empty((nBF, 3))
        h2d = 8 * (P * (nC + nBF) + (nC + nBF) + (nIF + nBF) + P * (nIF + nBF))
        d2h = 8 * (3 * (nC + nBF) + P * (nC + nBF) + P * (nIF + nBF))

        def timed_host(fn, n):
            with torch.cuda.stream(stream):
                fn()
                barrier()
                tw = time.perf_counter()
                for _ in range(n):
                    fn()
                barrier()
                return max_over_ranks(time.perf_counter() - tw)

        # What a time-stepping caller really sends: alpha is only ever changed by the device (solve / the stage), so after
        # the first step its stamp matches and it is not uploaded again (tcif_host_alpha_stamps); F and alpha come back
        # as deltas against what the host arrays already hold (tcif_host_sparse_returns).  phi, T and the flux arrays
        # cross the link every step.  Bytes are the ones actually copied (tcif_host_bytes_moved), not a formula.
        ctx.host_sparse_returns(True)
        counter = [0]

        def stamped(call):
            def fn():
                k = counter[0]
                ctx.host_alpha_stamps([k] * P if k else None, [k + 1] * P)
                counter[0] = k + 1
                call()
            return fn

        # (1) exactly the stage the device-resident number times: every field it consumes comes from host memory and
        #     every field it produces goes back, pipelined over separate copy streams
        e2e_s = timed_host(stamped(lambda: ctx.stage_host(dt, ai, ab, Ti, Tb, phii, phib, pi0, pb0, pi, pb, Fi, Fb)), args.e2e_steps)
        moved = ctx.host_bytes_moved
        e2e = {"value": total_cells * args.e2e_steps / e2e_s / 1e6, "unit": UNIT,
               "h2d_bytes_per_step": int(sum_over_ranks(moved[0])), "d2h_bytes_per_step": int(sum_over_ranks(moved[1])),
               "dense_h2d_bytes_per_step": int(sum_over_ranks(h2d)), "dense_d2h_bytes_per_step": int(sum_over_ranks(d2h)),
               "steps": args.e2e_steps, "ms_per_step": 1e3 * e2e_s / args.e2e_steps,
               "api": "tcif_stage_host: H2D T,phi,phiAlphaCorr[P] (+alpha[P] in the first step only: stamps); D2H alphaPhi[P], F and alpha[P] as "
                      "deltas against what the same host arrays received last (dense when more than 1/4 changed); pinned host buffers, "
                      "uploads/compute/downloads pipelined; bytes = last step, as copied"}
        # (2) the reference's own call surface: fluid.solve() (flux assembly + implicit predictor + MULES + rhoPhi,
        #     all on the device) then fluid.divCapillaryStress(): more work per step, fewer bytes over PCIe
        ctx.set_cell_centres(np.asarray(f["C"]), np.asarray(f["deltab"]))
        for p in range(P):
            ctx.set_rho(p, (1000.0, 800.0, 1.2, 50.0)[p])
        ctl = abi.alpha_controls(1, 1, "vanLeer", "linear", True, 1e-10, 100)
        rpi, rpb = abi.pinned_empty((nIF,)), abi.pinned_empty((nBF,))

        for p in range(P):
            ai[p][...] = ai0[p]; ab[p][...] = ab0[p]
        counter[0] = 0                                   # the host rewrote alpha: unknown stamp, uploaded once more
        cs_s = timed_host(stamped(lambda: ctx.time_step_host(ctl, dt, ai, ab, Ti, Tb, phii, phib, rpi, rpb, Fi, Fb)), args.e2e_steps)
        moved = ctx.host_bytes_moved
        sweeps = [int(abi.lib().tcif_predictor_sweeps(ctx.h, p)) for p in range(P)]
        extra["e2e_call_surface"] = {
            "value": total_cells * args.e2e_steps / cs_s / 1e6, "unit": UNIT, "ms_per_step": 1e3 * cs_s / args.e2e_steps,
            "h2d_bytes_per_step": int(sum_over_ranks(moved[0])), "d2h_bytes_per_step": int(sum_over_ranks(moved[1])),
            "dense_h2d_bytes_per_step": int(sum_over_ranks(8 * ((P + 1) * (nC + nBF) + nIF + nBF))),
            "dense_d2h_bytes_per_step": int(sum_over_ranks(8 * ((P + 3) * (nC + nBF) + nIF + nBF))),
            "api": "tcif_time_step_host = multiphaseSystem::solve() (vanLeer + linear compression, implicit upwind "
                   "predictor, nAlphaCorr 1) + divCapillaryStress(); H2D T,phi (+alpha[P] first step only); D2H rhoPhi dense, "
                   "alpha[P] and F as deltas",
            "predictor_sweeps": sweeps}
        ctx.host_sparse_returns(False)
        # device-resident cost of that call surface
        with torch.cuda.stream(stream):
            ctx.solve_alphas(ctl, dt)
            barrier()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ctx.timers_enable(True)
            ctx.timers_reset()
            s0.record(stream)
            for _ in range(args.e2e_steps):
                ctx.solve_alphas(ctl, dt)
                ctx.div_capillary_stress()
            s1.record(stream)
            barrier()
            sm = max_over_ranks(s0.elapsed_time(s1))
            st = ctx.timers_read()
            ctx.timers_enable(False)
        extra["solve_plus_force_device"] = {"value": total_cells * args.e2e_steps / (sm * 1e-3) / 1e6, "unit": UNIT,
                                            "ms_per_step": sm / args.e2e_steps,
                                            "what": "tcif_solve_alphas + tcif_div_capillary_stress on resident fields",
                                            "per_kernel_ms_per_step": {k: v["ms"] / args.e2e_steps for k, v in st.items() if v["launches"]}}
        # what the host link of this box gives each rank when ALL ranks copy at once (the e2e numbers above are mostly
        # this): pinned 256 MB each way, CUDA events, slowest rank and the sum over ranks
        probe_n = 32 << 20
        hbuf = abi.pinned_empty((probe_n,)); hbuf[...] = 1.0
        dbuf = torch.empty(probe_n, dtype=torch.float64, device=dev)
        hten = torch.from_numpy(hbuf)
        gbs = []
        for direction in ("h2d", "d2h"):
            with torch.cuda.stream(stream):
                ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                for it in range(3):
                    if it == 1:
                        barrier(); ev0.record(stream)
                    if direction == "h2d":
                        dbuf.copy_(hten, non_blocking=True)
                    else:
                        hten.copy_(dbuf, non_blocking=True)
                ev1.record(stream)
                barrier()
                gbs.append(2 * 8 * probe_n / (ev0.elapsed_time(ev1) * 1e-3) / 1e9)
        extra["host_link_probe"] = {"what": "256 MB pinned copies, all ranks at once (torch copy_, CUDA events)",
                                    "h2d_gbs_slowest_rank": -max_over_ranks(-gbs[0]), "h2d_gbs_sum": sum_over_ranks(gbs[0]),
                                    "d2h_gbs_slowest_rank": -max_over_ranks(-gbs[1]), "d2h_gbs_sum": sum_over_ranks(gbs[1])}
        del hten, hbuf, dbuf
        del ai, ab, pi, pb, pi0, pb0, Ti, Tb, phii, phib, Fi, Fb, rpi, rpb

    stats = [ctx.alpha_stats(p) for p in range(P)]
    dev_bytes = ctx.device_bytes
    transport = ctx.comm_transport
    ctx.close()

    # ---- CPU baseline beside it (rank 0, N=1 only) --------------------------------------------
    cpu = None
    if world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        edge = cpu_sample_edge(args, cores)
        nst = 16
        rate, secs, nd, cprocs, cspec = cpu_stage_rate(args.config, edge, nst, cores)
        one_core, one_core_sample = cpu_one_core_rate(args.config)
        cpu = {"value": rate, "unit": UNIT, "cores": nd, "kind": "port", "host_cores": cores,
               "one_core": {"value": one_core, "unit": UNIT, "sample": one_core_sample},
               "sample": f"{cspec.name} {edge}^3 = {cspec.nCells} cells, {cspec.nPhases} phases, {nst} stages, "
                         f"{nd} sub-domains {cprocs[0]}x{cprocs[1]}x{cprocs[2]}, {secs:.1f} s"}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling,
                "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(args) if args.scaling == "strong" else
                           f"configs[{args.config - 1}] {spec.name} weak scaling {spec.n[0]}x{spec.n[1]}x{spec.n[2]} = {spec.nCells} cells, {spec.nPhases} phases",
                           "phases": P, "pairs": len(spec.sigma), "nLimiterIter": 3,
                           "decomposition": f"simple {procs[0]}x{procs[1]}x{procs[2]}", "cells_per_gpu": nC,
                           "halo_transport": transport,
                           "l2": "inputs larger than L2 (no flush needed)" if dev_bytes > 4 * 126e6 else
                                 "L2-resident case: latency-bound, no roofline claim",
                           "device_bytes_per_gpu": dev_bytes, "setup_s": round(setup_s, 1)},
                "clocks": clocks, "e2e": e2e, **extra, "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu,
                "alpha_stats": [{"sumValpha": s[0], "min": s[1], "max": s[2]} for s in stats]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}")
    if args.gpus > 1 and world == 1:
        raise SystemExit("bench.py: N>1 must be launched with torch.distributed.run (one rank per GPU)")
    if args.impl == "reference":
        run_reference(args, rank)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
