#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 bundle-adjustment hot path.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload C4] [--weak]

A "step" is one Levenberg-Marquardt solve of LM_ITERS fixed iterations (tolerances 0, as the
reference sets them: src/optim/bundle_adjustment.h:81-83) over one synthetic scene of a
BASELINE.json shape: residual + Jacobian evaluation over all observations (with the bilinear semantic
row where the config is a "semantic BA" one), point-block Schur elimination, reduced-system solve,
update, candidate cost -- the whole hot path.

Workload.  N = 1: C4 (1778 cams / 993 k pts / 5 M obs, ITERATIVE_SCHUR), the largest BASELINE config that
names one GPU; C2 and C3 ride along as `extra_workloads` (--extra).  N > 1: the SAME C4 problem split over
the ranks by point (strong scaling, the curve north_star asks for); --weak keeps the points per GPU fixed
instead; --workload C5 runs the 5 k-camera semantic scene.
`value` = observations x LM iterations per second with all inputs already resident in HBM
(sba_reset between steps); `e2e` = the same through the C ABI from HOST buffers (sba_create uploads,
sba_solve, sba_get_parameters reads back) with the copies inside the timed region -- page-locked caller
arrays in `e2e`, the pageable std::vector case a COLMAP caller has in `e2e_pageable`.
--impl reference: the CPU restatement of the Ceres path (oracle/) on all host cores, same scene, same options.
"""
import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
LM_ITERS = 10
# Initial guess: scene.py's perturbation (t 0.01, rotation 0.002 rad, X 0.01) times this.  At x1 Gauss-Newton is
# at the noise floor after 6 iterations and the rest of a 10-iteration solve measures rounding; at x30 (t 0.3, 0.06 rad,
# X 0.3 on a unit-cube scene seen from 10 units) every one of the 10 iterations moves the cost and the trust region
# really works (rejected steps, radius changes), on the device and on the CPU alike.
PERTURB_SCALE = 30.0
METRIC = "residual+Jacobian observations/s through the LM solve (observations x LM iterations / s)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C4", help="C1..C5 (BASELINE.json configs); C4 is the headline")
    ap.add_argument("--extra", default="C2,C3", help="smaller workloads appended at N = 1 (solve rate, e2e, stage table); '' = none")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--weak", action="store_true", help="N > 1: keep the points per GPU fixed (default: strong, the named config is the TOTAL problem)")
    ap.add_argument("--strong", action="store_true", help="(default at N > 1; accepted for compatibility)")
    ap.add_argument("--cpu-threads", type=int, default=0, help="threads of the CPU arm (default: every core this process may use)")
    return ap.parse_args()


def host_cores(args):
    """Cores the CPU arm uses: passed to the oracle explicitly, so torchrun's OMP_NUM_THREADS=1 cannot shrink it."""
    if args.cpu_threads > 0:
        return args.cpu_threads
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region (B200_PROFILING.md): NVML in-process every 5 ms
    (the timed region of the C2 headline is tens of milliseconds -- too short for an `nvidia-smi -lms` child)."""
    REASONS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self.err = [], set(), None, None
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self._sample()
            self.th = threading.Thread(target=self._run, daemon=True)
            self.th.start()
        except Exception as e:  # noqa: BLE001 -- the bench line says why there are no clocks
            self.nv, self.err = None, repr(e)

    def _sample(self):
        self.samples.append(float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
        mask = int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
        for name, bit in self.REASONS.items():
            if mask & bit:
                self.reasons.add(name)

    def _run(self):
        while not self._stop.wait(0.005):
            try:
                self._sample()
            except Exception:  # noqa: BLE001
                return

    def stop(self):
        if self.nv is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: %s" % self.err], "samples": 0}
        try:
            self._sample()
        except Exception:  # noqa: BLE001
            pass
        self._stop.set()
        self.th.join(timeout=1)
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def dist_env():
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


def make_workload(m, name, n_gpus, weak=False):
    """Strong scaling (default): the named config is the whole problem.  Weak: points (and with them observations) per
    GPU fixed, cameras fixed."""
    cfg = dict(m.scene.CONFIGS[name])
    cfg["n_pt"] = cfg["n_pt"] * (n_gpus if weak else 1)
    prob, _ = m.scene.make_scene(seed=0, perturb_scale=PERTURB_SCALE, **cfg)  # C2 / C3 / C5 carry their distance maps
    return prob, cfg


def solver_name(m, n_img):
    """BundleAdjuster::Solve's policy (src/optim/bundle_adjustment.cc:276-286)."""
    return ("DENSE_SCHUR device Cholesky" if n_img <= 50 else
            "SPARSE_SCHUR sizes -> device dense Cholesky" if n_img <= 1000 else "ITERATIVE_SCHUR PCG + SCHUR_JACOBI")


def config_dict(m, name, world, prob, weak):
    """The SAME dictionary in both arms (the driver compares them)."""
    sem = prob.sem_maps is not None
    return {"workload": "%s x%d: %d cams / %d pts / %d obs, %s, %d LM iterations (tolerances 0), %s" % (
                name, world, len(prob.image_camera), len(prob.points), prob.num_observations,
                "3 residual rows (reprojection + bilinear semantic row, %dx%d float32 map per image, weight %g)" % (
                    prob.sem_maps.shape[2], prob.sem_maps.shape[1], prob.sem_weight) if sem else "2 residual rows (reprojection)",
                LM_ITERS, solver_name(m, len(prob.image_camera))),
            "initial_guess": "scene.py perturbation x%g" % PERTURB_SCALE,
            "scaling": "weak (points per GPU fixed)" if weak and world > 1 else "strong (one problem split by point)" if world > 1 else "one GPU",
            "l2": "the working set is larger than L2 at C3-C5; stage timings flush L2 (256 MB memset) between launches",
            "sharding": "points (with their observations) split across ranks, cameras replicated"}


def set_options(o, n_img, max_iters=LM_ITERS):
    o.max_num_iterations = max_iters
    o.num_config_images = n_img
    return o


def solver_options(m, n_img, max_iters=LM_ITERS):
    return set_options(m.capi.default_solver_options(), n_img, max_iters)


def oracle_options(m, n_img, max_iters=LM_ITERS):
    """The same defaults restated by the oracle (oracle_default_solver_options): the CPU arm never opens libsba_b200.so."""
    from oracle import loader
    o = m.capi.SolverOptions()
    loader.load(m.capi).oracle_default_solver_options(C.byref(o))
    return set_options(o, n_img, max_iters)


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def cpu_solve(m, prob, opts, threads, cap=64):
    from oracle import loader
    ora = loader.load(m.capi)
    s = m.capi.Summary()
    tr = (m.capi.IterationSummary * cap)()
    n = C.c_int32()
    d = prob.desc()
    t = time.perf_counter()
    ora.oracle_ba_solve(C.byref(d), C.byref(opts), threads, C.byref(s), tr, cap, C.byref(n), None, None, None, None)
    return time.perf_counter() - t, s, [tr[i] for i in range(n.value)]


def accept_string(trace):
    """One character per LM iteration: A accepted, r rejected, i invalid step."""
    return "".join("A" if t.step_is_successful else ("r" if t.step_is_valid else "i") for t in trace[1:])


def run_reference(args, m):
    """--impl reference: the reference's CPU implementation of the path on the host cores.  Ceres is not
    installable here (no Eigen/Ceres/glog), so this arm times the oracle port of the Ceres path."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    weak = args.weak and args.gpus > 1
    prob, cfg = make_workload(m, args.workload, args.gpus, weak)
    o = oracle_options(m, len(prob.image_camera))
    threads = host_cores(args)
    # bounded: the whole arm ends within a few minutes whatever K, W and the scene size are -- warm-ups, then timed
    # steps, are dropped once the projected total would pass the budget (at least one timed solve always runs)
    budget_s, t_begin, last = 150.0, time.perf_counter(), 0.0
    for i in range(args.warmup):
        if (time.perf_counter() - t_begin) + last * 2 > budget_s / 2 or (i == 0 and prob.num_observations > 2000000):
            break  # (a multi-million-observation solve takes tens of seconds: its first run is the timed one)
        last = cpu_solve(m, prob, o, threads)[0]
    ts = []
    for _ in range(args.steps):
        if ts and (time.perf_counter() - t_begin) + ts[-1] > budget_s:
            break
        dt, s, tr = cpu_solve(m, prob, o, threads)
        ts.append(dt)
    n_iter = s.num_successful_steps + s.num_unsuccessful_steps
    sec = float(np.sum(ts)) / len(ts)
    val = prob.num_observations * n_iter / sec
    line = {
        "impl": "reference", "metric": METRIC,
        "value": val, "unit": "observations/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak" if weak else "strong" if args.gpus > 1 else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "lm_iterations_per_s": n_iter / sec,
        "config": config_dict(m, args.workload, args.gpus, prob, weak),
        "lm_iterations_ran": n_iter, "accept_reject": accept_string(tr),
        "cpu_baseline": {"value": val, "unit": "observations/s", "cores": threads, "kind": "port",
                         "sample": "%d full solves of %d LM iterations on the whole workload (%d asked; 150 s budget)" % (
                             len(ts), LM_ITERS, args.steps)},
        "steps_run": len(ts),
        "e2e": {"value": val, "unit": "observations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "final_cost": s.final_cost,
    }
    print(json.dumps(line), flush=True)


def pba_second_opinion(m):
    """The reference's own lib/PBA CPU solver (oracle/_ref/libpba_ref.so, compiled from /root/reference/lib/PBA where the
    sources lie) beside this library on the C1 shape -- the one BASELINE config PBA accepts (SIMPLE_RADIAL, unshared
    intrinsics; bundle_adjustment.cc:629-651).  A different LM, reported as a second opinion only."""
    ref = os.path.join(ROOT, "oracle", "_ref", "libpba_ref.so")
    if not os.path.exists(ref):
        return {"unavailable": "oracle/_ref/libpba_ref.so not built"}
    lib = C.CDLL(ref)
    lib.pba_ref_solve.argtypes = [C.POINTER(m.capi.ProblemDesc), C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double),
                                  C.POINTER(C.c_double)]
    prob, _ = m.scene.make_config("C1")
    d = prob.desc()
    out = np.zeros(5)
    threads = max(1, len(os.sched_getaffinity(0)))
    # lib/PBA prints a banner from C: keep stdout to the one JSON line (the banner goes to stderr)
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    try:
        rc = lib.pba_ref_solve(C.byref(d), LM_ITERS, threads, 1, out.ctypes.data_as(C.POINTER(C.c_double)), None)
        C.CDLL(None).fflush(None)
    finally:
        os.dup2(saved, 1)
        os.close(saved)
    if rc != 0:
        return {"unavailable": "PBA rejected the problem"}
    s = m.Solver(prob)
    o = solver_options(m, len(prob.image_camera))
    s.solve(o); s.reset()
    summ = s.solve(o)
    s.close()
    it_gpu = summ.num_successful_steps + summ.num_unsuccessful_steps
    return {"workload": "C1: %d cams / %d pts / %d obs, SIMPLE_RADIAL" % (len(prob.image_camera), len(prob.points), prob.num_observations),
            "pba_cpu_double": {"observations_per_s": prob.num_observations * max(out[2], 1) / out[3], "lm_iterations": int(out[2]),
                               "seconds": out[3], "threads": threads, "final_cost": out[1]},
            "b200": {"observations_per_s": prob.num_observations * it_gpu / summ.device_time_s, "lm_iterations": it_gpu,
                     "seconds": summ.device_time_s, "final_cost": summ.final_cost}}


def stage_table(solver, stages, repeat, hbm_peak):
    out = {}
    for st in stages:
        try:
            ms, nbytes = solver.bench_stage(st, repeat)
        except Exception as e:  # a stage that does not apply to this solver regime
            out[st] = {"error": str(e)}
            continue
        gbs = nbytes / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        out[st] = {"ms": ms, "algorithmic_bytes": nbytes, "GBps": gbs, "frac_of_hbm_peak": gbs / hbm_peak}
    return out


def resident_rate(s, o, n_obs, steps, warmup, barrier=lambda: None):
    """`steps` re-solves of one handle (sba_reset restores the creation parameters): device time, wall time, last summary."""
    for _ in range(warmup):
        s.reset(); summ = s.solve(o)
    barrier()
    dev_s, wall_s = 0.0, 0.0
    for _ in range(steps):
        s.reset()
        barrier()
        t = time.perf_counter()
        summ = s.solve(o)  # ends with a stream synchronize
        wall_s += time.perf_counter() - t
        dev_s += summ.device_time_s
    barrier()
    return dev_s, wall_s, summ


def e2e_rate(new_solver, o, steps, warmup, barrier=lambda: None):
    """create (H2D) -> solve -> read back (D2H) -> destroy, a fresh handle every step, wall clock."""
    total, phases, out, each = 0.0, np.zeros(3), None, []
    for i in range(warmup + steps):
        barrier()
        t = time.perf_counter()
        s2 = new_solver()
        t1 = time.perf_counter()
        s2.solve(o)
        t2 = time.perf_counter()
        out = s2.parameters()
        s2.close()
        dt = time.perf_counter() - t
        if i >= warmup:
            total += dt
            phases += np.array([t1 - t, t2 - t1, t + dt - t2])
            each.append(round(dt * 1e3, 2))
    ph = dict(zip(("create_h2d", "solve", "read_back_destroy"), (phases / steps * 1e3).round(3).tolist()))
    ph["each_step_total"] = each  # (a single slow step -- a host hiccup -- shows here instead of hiding in the mean)
    return total, ph, out


def host_flatten(prob):
    """Host side of the drop-in, measured beside the device path: the C++ shim's Flatten (shim/sba_bundle_adjuster.h) over a mock
    Reconstruction of the workload's size -- what a COLMAP caller pays before sba_create sees an array.  Built by build();
    None when the helper is missing or fails (it is a report, not part of any timed region)."""
    exe = os.path.join(ROOT, "tools", "shim_flatten_bench.bin")
    try:
        out = subprocess.run([exe, str(len(prob.image_camera)), str(len(prob.points)), "--json"], capture_output=True, text=True, timeout=180)
        d = json.loads(out.stdout.strip().splitlines()[-1])
        d["what"] = ("BundleAdjusterT::Flatten over a mock Reconstruction of this many images / points (mean track 5.03), one host "
                     "thread, median / min / max of 3; not included in e2e")
        return d
    except Exception as e:  # noqa: BLE001
        return {"unavailable": str(e)[:200]}


def io_bytes(prob, out):
    h2d = int(prob.image_qvec.nbytes + prob.image_tvec.nbytes + prob.camera_params.nbytes + prob.points.nbytes +
              prob.obs_xy.nbytes + prob.obs_image.nbytes + prob.obs_point.nbytes +
              (prob.sem_maps.nbytes if prob.sem_maps is not None else 0))
    return h2d, int(sum(a.nbytes for a in out))


def stages_for(m, summ):
    dense = summ.linear_solver_type_used != m.capi.ITERATIVE_SCHUR
    return ["eval", "cost", "point_blocks", "point_refresh", "rhs", "backsub"] + (
        ["schur_dense", "cholesky"] if dense else ["schur_jacobi", "matvec", "schur_blocks", "matvec_blocks"])


def run_b200(args, m):
    rank, world, local = dist_env()
    lib = m.capi.load()
    if lib.sba_device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: the product has no CPU fallback")
    m.capi.check(lib.sba_set_device(local))
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    weak = args.weak and world > 1
    prob, cfg = make_workload(m, args.workload, world, weak)
    n_obs = prob.num_observations
    o = solver_options(m, len(prob.image_camera))

    def barrier():
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    uid = None
    if world > 1:
        obj = [m.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(obj, src=0)
        uid = obj[0]

    def new_solver():
        s = m.Solver(prob, shard_rank=rank, shard_count=world)
        if world > 1:
            s.init_comm(world, rank, uid)
        return s

    # ---- resident: inputs already in HBM when the timed region starts ----
    s = new_solver()
    s.solve(o)  # first solve of the handle: pair lists, block structure, first-touch allocations (untimed, before warm-up)
    clocks = None
    l0 = [0]

    def start_clocks():
        nonlocal clocks
        clocks = ClockSampler(local)
        l0[0] = s.launches()
    for _ in range(args.warmup):
        s.reset(); s.solve(o)
    start_clocks()
    dev_s, wall_s, summ = resident_rate(s, o, n_obs, args.steps, 0, barrier)
    clk = clocks.stop()
    launches = s.launches() - l0[0]
    dev_s, wall_s = max_over_ranks(dev_s), max_over_ranks(wall_s)
    n_iter = summ.num_successful_steps + summ.num_unsuccessful_steps
    value = n_obs * n_iter * args.steps / dev_s
    trace = s.trace()

    # ---- end to end through the C ABI from host buffers: pageable arrays first (what a COLMAP caller's std::vectors
    #      are), then page-locked ones (sba_host_alloc: the copies inside sba_create are DMA from the caller's buffers) ----
    e2e_steps = max(1, min(args.steps, 10))
    pg_s, pg_phases, out = e2e_rate(new_solver, o, e2e_steps, min(args.warmup, 2), barrier)
    pg_s = max_over_ranks(pg_s)
    prob.pin()
    e2e_s, e2e_phases, out = e2e_rate(new_solver, o, e2e_steps, min(args.warmup, 2), barrier)
    e2e_s = max_over_ranks(e2e_s)
    h2d, d2h = io_bytes(prob, out)
    # workloads with label maps: the maps uploaded ONCE (sba_maps_create, outside the timed region, as a mapper that
    # adjusts the same images again and again would keep them), everything else from host buffers every step
    e2e_resident_maps = None
    if prob.sem_maps is not None and world == 1:
        maps = m.ResidentMaps(prob.sem_maps)
        rm_s, rm_phases, out_rm = e2e_rate(lambda: m.Solver(prob, maps=maps), o, e2e_steps, min(args.warmup, 2), barrier)
        maps.close()
        e2e_resident_maps = {"value": n_obs * n_iter * e2e_steps / rm_s, "unit": "observations/s",
                             "ms_per_step": rm_s / e2e_steps * 1e3, "steps": e2e_steps,
                             "h2d_bytes_per_step": h2d - int(prob.sem_maps.nbytes), "d2h_bytes_per_step": d2h,
                             "maps": "resident on the device (sba_maps_create, %d bytes uploaded once outside the timed region)" % prob.sem_maps.nbytes,
                             "phases_ms": rm_phases,
                             "same_result": bool(all(np.array_equal(a, b) for a, b in zip(out, out_rm)))}

    # ---- stage times (L2 flushed between launches) and the roofline of the dominant HBM-bound stage ----
    hbm_peak, peak_kind = peaks()
    table = stage_table(s, stages_for(m, summ), 20 if n_obs < 1000000 else 8, hbm_peak)
    hbm_stages = {k: v for k, v in table.items() if "ms" in v and k not in ("cholesky", "matvec", "schur_jacobi", "matvec_blocks")}
    top = max(hbm_stages, key=lambda k: hbm_stages[k]["ms"])
    traffic = None
    try:  # per-launch DRAM bytes of the same stage from the committed ncu --set full capture (profiles/traffic.json)
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get("%s/%s" % (args.workload, top), {}).get("bytes")
    except Exception:  # noqa: BLE001
        pass
    roof = {"bound": "hbm", "kernel": top, "achieved": table[top]["GBps"], "peak": hbm_peak, "unit": "GB/s",
            "frac": table[top]["GBps"] / hbm_peak, "traffic": traffic, "peak_source": peak_kind,
            "eval_frac": table["eval"]["GBps"] / hbm_peak, "eval_frac_of_8TBps": table["eval"]["GBps"] / 8000.0,
            "note": "stage = all kernels of the named step of ONE LM iteration, L2 flushed (256 MB memset) before each "
                    "launch, CUDA events on the solver's stream; achieved = algorithmic bytes (DESIGN.md section 4) / time"}
    line = {
        "metric": METRIC,
        "value": value, "unit": "observations/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_s / args.steps * 1e3, "wall_ms_per_step": wall_s / args.steps * 1e3,
        "higher_is_better": True, "scaling": "weak" if (weak or world == 1) else "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "lm_iterations_per_s": n_iter * args.steps / dev_s, "lm_iterations_ran": n_iter, "accept_reject": accept_string(trace),
        "linear_solver_iterations": int(summ.num_linear_solver_iterations),
        "eval_observations_per_s": n_obs / (table["eval"]["ms"] * 1e-3) if world == 1 else None,
        "config": config_dict(m, args.workload, world, prob, weak),
        "clocks": clk, "gpu_launches": int(launches),
        "e2e": {"value": n_obs * n_iter * e2e_steps / e2e_s, "unit": "observations/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_s / e2e_steps * 1e3,
                "steps": e2e_steps, "host_memory": "page-locked (sba_host_alloc)", "phases_ms": e2e_phases},
        "e2e_pageable": {"value": n_obs * n_iter * e2e_steps / pg_s, "unit": "observations/s",
                         "ms_per_step": pg_s / e2e_steps * 1e3, "steps": e2e_steps,
                         "host_memory": "pageable numpy arrays (a COLMAP caller's std::vector)", "phases_ms": pg_phases},
        "e2e_maps_resident": e2e_resident_maps,
        "host_flatten": host_flatten(prob) if (rank == 0 and world == 1) else None,
        "roofline": roof, "stages": table, "final_cost": summ.final_cost,
        # per-stage device time of the LAST timed solve (CUDA events between the stages; all zero when the iteration
        # replays as a CUDA graph -- run with SBA_NO_GRAPHS=1 to fill them)
        "solve_stage_ms": {"evaluate": summ.time_evaluate_s * 1e3, "eliminate": summ.time_eliminate_s * 1e3,
                           "linear_solve": summ.time_linear_solve_s * 1e3, "update": summ.time_update_s * 1e3,
                           "device_total": summ.device_time_s * 1e3},
    }
    if world > 1:
        # per-rank timeline of the last timed solve (filled when SBA_NO_GRAPHS=1 keeps plain launches with stage events)
        per_rank = [None] * world
        dist.all_gather_object(per_rank, dict(line["solve_stage_ms"], rank=rank))
        line["solve_stage_ms_per_rank"] = per_rank
        # correctness gate of the sharded solve: rank 0 solves the SAME problem unsharded on its own GPU
        s.close()
        rel = None
        if rank == 0:
            s1 = m.Solver(prob)
            one = s1.solve(o)
            s1.close()
            rel = abs(one.final_cost - summ.final_cost) / one.final_cost
            line["parity_vs_1gpu"] = {"final_cost_1gpu": one.final_cost, "final_cost": summ.final_cost, "rel_diff": rel,
                                      "lm_iterations_1gpu": one.num_successful_steps + one.num_unsuccessful_steps,
                                      "ok": bool(rel <= 1e-6)}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # bounded sample: the same scene, the first CPU_ITERS LM iterations (the metric is per observation x iteration)
        cpu_iters = LM_ITERS if n_obs < 6000000 else 2  # C4: the whole solve, ~10 s on 16 threads; C5: 2 iterations, ~9 s
        threads = host_cores(args)
        dt, so, tro = cpu_solve(m, prob, oracle_options(m, len(prob.image_camera), cpu_iters), threads)
        it_o = so.num_successful_steps + so.num_unsuccessful_steps
        k = min(len(tro), len(trace)) - 1
        line["cpu_baseline"] = {"value": n_obs * it_o / dt, "unit": "observations/s", "cores": threads, "kind": "port",
                                "sample": "1 solve of the first %d of the %d LM iterations on the whole workload (%.1f s)" % (it_o, LM_ITERS, dt),
                                "cost_after_sample": tro[k].cost, "device_cost_at_same_iteration": trace[k].cost,
                                "cost_rel_diff_at_same_iteration": abs(tro[k].cost - trace[k].cost) / tro[k].cost,
                                "accept_reject": accept_string(tro)}
    if rank == 0 and world == 1 and args.extra:
        s.close()
        line["extra_workloads"] = {}
        for name in [x for x in args.extra.split(",") if x and x != args.workload]:
            line["extra_workloads"][name] = run_extra(args, m, name, hbm_peak)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["pba_reference"] = pba_second_opinion(m)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def run_extra(args, m, name, hbm_peak):
    """A smaller BASELINE config beside the headline: resident solve rate (median of three re-solves of one handle, the
    steady state of the headline's timed steps), end to end from pinned host buffers, stage table, CPU arm (C2 only)."""
    big, _ = make_workload(m, name, 1)
    n_obs = big.num_observations
    ob = solver_options(m, len(big.image_camera))
    sb = m.Solver(big)
    sb.solve(ob)   # first solve of the handle: pair lists, block structure, first-touch allocations
    sb.reset(); sb.solve(ob)   # second: the iteration's CUDA graphs are captured and instantiated here
    runs = []
    for _ in range(3):
        sb.reset()
        runs.append(sb.solve(ob))
    sm = sorted(runs, key=lambda r: r.device_time_s)[1]
    itb = sm.num_successful_steps + sm.num_unsuccessful_steps
    tr = sb.trace()
    out = {"config": config_dict(m, name, 1, big, False), "n_obs": n_obs, "ms_per_solve": sm.device_time_s * 1e3,
           "lm_iterations_per_s": itb / sm.device_time_s, "observations_per_s": n_obs * itb / sm.device_time_s,
           "lm_iterations_ran": itb, "accept_reject": accept_string(tr),
           "linear_solver_iterations": int(sm.num_linear_solver_iterations), "final_cost": sm.final_cost,
           "stages": stage_table(sb, stages_for(m, sm), 10, hbm_peak)}
    sb.close()
    big.pin()
    e_s, e_ph, o3 = e2e_rate(lambda: m.Solver(big), ob, 3, 1)
    h2d, d2h = io_bytes(big, o3)
    out["e2e"] = {"value": n_obs * itb * 3 / e_s, "unit": "observations/s", "ms_per_step": e_s / 3 * 1e3,
                  "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "phases_ms": e_ph}
    if not args.no_cpu_baseline:
        cpu_iters = LM_ITERS if n_obs < 400000 else 2
        dt, so, tro = cpu_solve(m, big, oracle_options(m, len(big.image_camera), cpu_iters), host_cores(args))
        it_o = so.num_successful_steps + so.num_unsuccessful_steps
        k = min(len(tro), len(tr)) - 1
        out["cpu_baseline"] = {"value": n_obs * it_o / dt, "unit": "observations/s", "cores": host_cores(args), "kind": "port",
                               "sample": "1 solve of %d LM iterations (%.1f s)" % (it_o, dt), "accept_reject": accept_string(tro),
                               "cost_rel_diff_at_same_iteration": abs(tro[k].cost - tr[k].cost) / tro[k].cost}
    return out


def main():
    args = parse()
    m = importlib.import_module("semantic-bundle-adjustment-colmap_b200")
    if args.impl == "reference":
        run_reference(args, m)
    else:
        run_b200(args, m)


if __name__ == "__main__":
    main()
