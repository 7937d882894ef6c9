"""Diagnostics (not a test): does a solve depend on what the process did before it (block cache, statics)?"""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.getcwd()); sys.path.insert(0, "tests")
m = importlib.import_module("semantic-bundle-adjustment-colmap_b200")
import test_gpu_bsr as T


def solve(prob, solver=2, iters=6):
    o = m.capi.default_solver_options(); o.max_num_iterations, o.linear_solver_type = iters, solver
    s = m.Solver(prob); summ = s.solve(o)
    out = (summ.final_cost, summ.num_linear_solver_iterations, float(np.abs(s.parameters()[3]).sum()),
           [t.linear_solver_iterations for t in s.trace()])
    s.close()
    return out


os.environ["SBA_IMPLICIT_SCHUR"] = "0"
target = T._scene(m, "radial", seed=8)
print("fresh      ", solve(target))
print("again      ", solve(target))
hist = {
    "pinhole": lambda: solve(T._scene(m, "pinhole_const", seed=3)),
    "dc10": lambda: solve(T._scene(m, "all_intrinsics_free", seed=3)),
    "dense": lambda: solve(m.scene.make_scene(n_cam=12, n_pt=1500, mean_track=5.0, seed=21)[0], solver=0),
    "dup": lambda: solve(T._scene(m, "duplicates", seed=3)),
    "sem": lambda: solve(m.scene.make_scene(n_cam=12, n_pt=1500, mean_track=5.0, model="SIMPLE_PINHOLE", refine_intr=False,
                                            width=320, height=240, semantic="distance", seed=8)[0]),
    "radial13": lambda: solve(T._scene(m, "radial", seed=13), iters=7),
}
for name, fn in hist.items():
    fn()
    print("after %-8s" % name, solve(target))
