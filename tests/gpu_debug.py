"""Ad-hoc GPU diagnostics (not a test): prints the GPU and oracle LM tables side by side."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np
from util import oracle_evaluate, oracle_solve
m = importlib.import_module("semantic-bundle-adjustment-colmap_b200")
from oracle import loader
ora = loader.load(m.capi)
name = sys.argv[1] if len(sys.argv) > 1 else "C1"
solver = int(sys.argv[2]) if len(sys.argv) > 2 else -1
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 10
prob, _ = m.scene.make_config(name)
print(name, "obs", prob.num_observations)
t = time.time(); s = m.Solver(prob); print("create %.1f ms" % ((time.time() - t) * 1e3))
t = time.time(); s2 = m.Solver(prob); print("create again %.1f ms" % ((time.time() - t) * 1e3)); s2.close()
ev = s.evaluate()
ref = oracle_evaluate(ora, m.capi, prob)
for k in ("r", "pose", "intr", "point"):
    print(k, "max abs diff", np.max(np.abs(ev[k] - ref[k])), "ref max", np.max(np.abs(ref[k])))
print("cost", ev["cost"], ref["cost"])
o = m.capi.default_solver_options(); o.max_num_iterations = iters; o.linear_solver_type = solver
t = time.time(); sg = s.solve(o); tg = time.time() - t
t = time.time(); rc, so, tro, po = oracle_solve(ora, m.capi, prob, o); to = time.time() - t
print("gpu %.3fs oracle %.3fs" % (tg, to), "final", sg.final_cost, so.final_cost, "rel", abs(sg.final_cost - so.final_cost) / so.final_cost)
for a, b in zip(s.trace(), tro):
    print("%3d %d%d cost %.10e / %.10e  g %.3e / %.3e  step %.3e / %.3e  rad %.2e / %.2e  lin %d / %d" % (
        a.iteration, a.step_is_successful, b.step_is_successful, a.cost, b.cost, a.gradient_max_norm, b.gradient_max_norm,
        a.step_norm, b.step_norm, a.trust_region_radius, b.trust_region_radius, a.linear_solver_iterations, b.linear_solver_iterations))
print({k: v for k, v in sg.as_dict().items() if k.startswith("time")}, "launches", s.launches())
