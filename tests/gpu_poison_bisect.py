"""Diagnostics (not a test): find the device allocation whose initial contents a solve depends on.
Runs the scene of tests/test_gpu_bsr.py::test_block_sparse_schur_matches_oracle in subprocesses with SBA_POISON ranges."""
import os
import subprocess
import sys

CHILD = r'''
import importlib, sys, os
import numpy as np
sys.path.insert(0, os.getcwd()); sys.path.insert(0, "tests")
m = importlib.import_module("semantic-bundle-adjustment-colmap_b200")
import test_gpu_bsr as T
prob = T._scene(m, "radial", seed=8)
os.environ["SBA_IMPLICIT_SCHUR"] = "0"
o = m.capi.default_solver_options(); o.max_num_iterations, o.linear_solver_type = 6, 2
s = m.Solver(prob); summ = s.solve(o)
print("RESULT %.17g %d %.17g" % (summ.final_cost, summ.num_linear_solver_iterations, float(np.abs(s.parameters()[3]).sum())))
'''


def run(lo, hi, log=False):
    env = dict(os.environ, SBA_POISON="1", SBA_POISON_FROM=str(lo), SBA_POISON_TO=str(hi))
    if log:
        env["SBA_POISON_LOG"] = "1"
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, capture_output=True, text=True)
    res = [l for l in r.stdout.splitlines() if l.startswith("RESULT")]
    return (res[0] if res else "FAILED " + r.stderr[-300:]), r.stderr


base, err = run(0, 0, log=True)
n = sum(1 for l in err.splitlines() if l.startswith("[sba poison] alloc"))
print("baseline (all zero):", base, "allocations:", n)
full, _ = run(0, 1 << 30)
print("all poisoned:", full)
lo, hi = 0, n
while hi - lo > 1 and full != base:
    mid = (lo + hi) // 2
    r, _ = run(lo, mid)
    if r != base:
        hi = mid
    else:
        r2, _ = run(mid, hi)
        if r2 != base:
            lo = mid
        else:
            print("needs both halves", lo, mid, hi); break
print("culprit ordinal range:", lo, hi)
for l in err.splitlines():
    if l.startswith("[sba poison] alloc %d:" % lo) or l.startswith("[sba poison] alloc %d:" % (lo - 1)) or l.startswith("[sba poison] alloc %d:" % (lo + 1)):
        print(l)
print("\n".join(err.splitlines()[:n + 5]))
