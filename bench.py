#!/usr/bin/env python
"""bench.py -- headline benchmark of the B200 bundle-adjustment hot path.

    python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload C2]

A "step" is one Levenberg-Marquardt solve of LM_ITERS fixed iterations (tolerances 0, as the
reference sets them: src/optim/bundle_adjustment.h:81-83) over one synthetic scene of the
BASELINE.json shape: residual + Jacobian evaluation over all observations, point-block Schur
elimination, reduced-system solve, update, candidate cost -- the whole hot path.
`value` = observations x LM iterations per second with all inputs already resident in HBM
(sba_reset between steps); `e2e` = the same through the C-ABI from HOST buffers (sba_create uploads,
sba_solve, sba_get_parameters reads back) with the copies inside the timed region.
N > 1: one process per GPU (torchrun), observations sharded by point, NCCL allreduce of the
camera-space quantities; weak scaling (points per GPU fixed).
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C2", help="C1..C5 (BASELINE.json configs); C2 is the headline")
    ap.add_argument("--extra", default="C4", help="larger workload whose stage rooflines are appended ('' = none)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--strong", action="store_true",
                    help="strong scaling: the named config is the TOTAL problem, split over the ranks (default: weak, per-GPU work fixed)")
    return ap.parse_args()


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


def make_workload(m, name, n_gpus, strong=False):
    """Weak scaling (default): points (and with them observations) per GPU fixed, cameras fixed."""
    cfg = dict(m.scene.CONFIGS[name])
    cfg["n_pt"] = cfg["n_pt"] * (1 if strong else n_gpus)
    prob, _ = m.scene.make_scene(seed=0, **cfg)  # C2 / C3 / C5 carry their distance maps (scene.CONFIGS)
    return prob, cfg


def solver_options(m, n_img):
    o = m.capi.default_solver_options()
    o.max_num_iterations = LM_ITERS
    o.num_config_images = n_img
    return o


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def cpu_solve(m, prob, opts, threads):
    from oracle import loader
    ora = loader.load(m.capi)
    s = m.capi.Summary()
    d = prob.desc()
    t = time.perf_counter()
    ora.oracle_ba_solve(C.byref(d), C.byref(opts), threads, C.byref(s), None, 0, None, None, None, None, None)
    return time.perf_counter() - t, s, ora.oracle_num_threads() if threads <= 0 else threads


def run_reference(args, m):
    """--impl reference: the reference's CPU implementation of the path on the host cores.  Ceres is not
    installable here (no Eigen/Ceres/glog), so this arm times the oracle port of the Ceres path."""
    rank, world, _ = dist_env()
    if rank != 0:
        return
    prob, cfg = make_workload(m, args.workload, args.gpus, args.strong)
    o = solver_options(m, len(prob.image_camera))
    # bounded: the whole arm ends within a few minutes whatever K, W and the scene size are -- warm-ups, then timed
    # steps, are dropped once the projected total would pass the budget (at least one timed solve always runs)
    budget_s, t_begin, last = 150.0, time.perf_counter(), 0.0
    for _ in range(args.warmup):
        if last > 0.0 and (time.perf_counter() - t_begin) + last * (1 + args.steps) > budget_s:
            break
        last = cpu_solve(m, prob, o, 0)[0]
    ts = []
    for _ in range(args.steps):
        if ts and (time.perf_counter() - t_begin) + ts[-1] > budget_s:
            break
        dt, s, threads = cpu_solve(m, prob, o, 0)
        ts.append(dt)
    n_iter = s.num_successful_steps + s.num_unsuccessful_steps
    sec = float(np.sum(ts)) / len(ts)
    val = prob.num_observations * n_iter / sec
    line = {
        "impl": "reference", "metric": "residual+Jacobian observations/s through the LM solve (observations x LM iterations / s)",
        "value": val, "unit": "observations/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "lm_iterations_per_s": n_iter / sec,
        "config": {"workload": "%s x%d: %d cams / %d pts / %d obs, %d LM iterations, reference solver policy" % (
            args.workload, args.gpus, len(prob.image_camera), len(prob.points), prob.num_observations, LM_ITERS)},
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
    threads = os.cpu_count() or 1
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
    prob, cfg = make_workload(m, args.workload, args.gpus, args.strong)
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

    s = new_solver()
    for _ in range(args.warmup):
        s.reset(); summ = s.solve(o)
    barrier()
    clocks = ClockSampler(local)
    l0 = s.launches()
    dev_s, wall_s = 0.0, 0.0
    for _ in range(args.steps):
        s.reset()
        barrier()
        t = time.perf_counter()
        summ = s.solve(o)  # ends with a stream synchronize
        wall_s += time.perf_counter() - t
        dev_s += summ.device_time_s
    barrier()
    clk = clocks.stop()
    launches = s.launches() - l0
    dev_s, wall_s = max_over_ranks(dev_s), max_over_ranks(wall_s)
    n_iter = summ.num_successful_steps + summ.num_unsuccessful_steps
    value = n_obs * n_iter * args.steps / dev_s

    # ---- end to end through the C ABI from host buffers (the caller's arrays are page-locked, as a production
    #      caller that hands the same reconstruction to repeated solves would keep them) ----
    prob.pin()
    e2e_s, e2e_phases = 0.0, np.zeros(3)
    for i in range(args.warmup + args.steps):
        barrier()
        t = time.perf_counter()
        s2 = new_solver()
        t1 = time.perf_counter()
        s2.solve(o)
        t2 = time.perf_counter()
        q, tt, cam, X = s2.parameters()
        s2.close()
        dt = time.perf_counter() - t
        if i >= args.warmup:
            e2e_s += dt
            e2e_phases += np.array([t1 - t, t2 - t1, t + dt - t2])
    e2e_s = max_over_ranks(e2e_s)
    h2d = int(prob.image_qvec.nbytes + prob.image_tvec.nbytes + prob.camera_params.nbytes + prob.points.nbytes +
              prob.obs_xy.nbytes + prob.obs_image.nbytes + prob.obs_point.nbytes +
              (prob.sem_maps.nbytes if prob.sem_maps is not None else 0))
    d2h = int(q.nbytes + tt.nbytes + cam.nbytes + X.nbytes)

    # ---- stage times (L2 flushed between launches) and the roofline of the dominant HBM-bound stage ----
    hbm_peak, peak_kind = peaks()
    dense = summ.linear_solver_type_used != m.capi.ITERATIVE_SCHUR
    stages = ["eval", "cost", "point_blocks", "rhs", "backsub"] + (["schur_dense", "cholesky"] if dense else ["schur_jacobi", "matvec", "schur_blocks", "matvec_blocks"])
    table = stage_table(s, stages, 20, hbm_peak)
    hbm_stages = {k: v for k, v in table.items() if "ms" in v and k != "cholesky"}
    top = max(hbm_stages, key=lambda k: hbm_stages[k]["ms"])
    traffic = None
    try:  # per-launch DRAM bytes of the same stage from the committed ncu --set full capture (profiles/traffic.json)
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            traffic = json.load(f).get("%s/%s" % (args.workload, top), {}).get("bytes")
    except Exception:  # noqa: BLE001
        pass
    roof = {"bound": "hbm", "kernel": top, "achieved": table[top]["GBps"], "peak": hbm_peak, "unit": "GB/s",
            "frac": table[top]["GBps"] / hbm_peak, "traffic": traffic, "peak_source": peak_kind,
            "eval_frac": table["eval"]["GBps"] / hbm_peak,
            "note": "stage = all kernels of the named step, L2 flushed before each launch; at C2 the whole working set "
                    "(28 MB) is smaller than L2 and every stage is launch/latency-bound -- the C4 stage table in "
                    "extra_workloads is the HBM-bound regime"}
    line = {
        "metric": "residual+Jacobian observations/s through the LM solve (observations x LM iterations / s)",
        "value": value, "unit": "observations/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_s / args.steps * 1e3, "wall_ms_per_step": wall_s / args.steps * 1e3,
        "higher_is_better": True, "scaling": "strong" if args.strong else "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "lm_iterations_per_s": n_iter * args.steps / dev_s,
        "eval_observations_per_s": n_obs / world / (table["eval"]["ms"] * 1e-3) * world,
        "config": {"workload": "%s x%d: %d cams / %d pts / %d obs, max %d LM iterations (tolerances 0; ran %d), %s" % (
            args.workload, world, len(prob.image_camera), len(prob.points), n_obs, LM_ITERS, n_iter,
            "DENSE_SCHUR device Cholesky" if dense else "ITERATIVE_SCHUR PCG + SCHUR_JACOBI"),
            "l2": "stage timings flush L2 (256 MB memset) between launches; the LM step re-reads its own Jacobian",
            "sharding": "points (with their observations) split across ranks, cameras replicated"},
        "clocks": clk, "gpu_launches": int(launches),
        "e2e": {"value": n_obs * n_iter * args.steps / e2e_s, "unit": "observations/s",
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": e2e_s / args.steps * 1e3,
                "phases_ms": dict(zip(("create_h2d", "solve", "read_back_destroy"), (e2e_phases / args.steps * 1e3).round(3).tolist()))},
        "roofline": roof, "stages": table, "final_cost": summ.final_cost,
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        dt, so, threads = cpu_solve(m, prob, o, 0)
        it_o = so.num_successful_steps + so.num_unsuccessful_steps
        line["cpu_baseline"] = {"value": n_obs * it_o / dt, "unit": "observations/s", "cores": threads, "kind": "port",
                                "sample": "1 full solve of %d LM iterations on the whole workload (%.2f s)" % (LM_ITERS, dt),
                                "final_cost": so.final_cost,
                                "final_cost_rel_diff": abs(so.final_cost - summ.final_cost) / so.final_cost}
    if rank == 0 and world == 1 and args.extra and args.extra != args.workload:
        s.close()
        big, _ = make_workload(m, args.extra, 1)
        sb = m.Solver(big)
        ob = solver_options(m, len(big.image_camera))
        sb.solve(ob)   # first solve of the handle: pair lists, block structure, first-touch allocations
        sb.reset()
        sb.solve(ob)   # second: the iteration's CUDA graphs are captured and instantiated here
        runs = []
        for _ in range(3):  # steady state of a re-solved handle, like the headline's timed steps: median of three
            sb.reset()
            runs.append(sb.solve(ob))
        sm = sorted(runs, key=lambda r: r.device_time_s)[1]
        itb = sm.num_successful_steps + sm.num_unsuccessful_steps
        it_stages = ["eval", "cost", "point_blocks", "rhs", "backsub"] + (
            ["schur_jacobi", "matvec", "schur_blocks", "matvec_blocks"] if sm.linear_solver_type_used == m.capi.ITERATIVE_SCHUR else ["schur_dense", "cholesky"])
        line["extra_workloads"] = {args.extra: {
            "n_obs": big.num_observations, "lm_iterations_per_s": itb / sm.device_time_s,
            "observations_per_s": big.num_observations * itb / sm.device_time_s,
            "linear_solver_iterations": int(sm.num_linear_solver_iterations), "final_cost": sm.final_cost,
            "stages": stage_table(sb, it_stages, 10, hbm_peak)}}
        sb.close()
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line["pba_reference"] = pba_second_opinion(m)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse()
    m = importlib.import_module("semantic-bundle-adjustment-colmap_b200")
    if args.impl == "reference":
        run_reference(args, m)
    else:
        run_b200(args, m)


if __name__ == "__main__":
    main()
