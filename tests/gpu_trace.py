"""Prints the LM iteration table of one BASELINE-shaped solve (diagnostics)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
m = importlib.import_module("semantic-bundle-adjustment-colmap_b200")
name = sys.argv[1] if len(sys.argv) > 1 else "C2"
prob, _ = m.scene.make_config(name)
o = m.capi.default_solver_options(); o.max_num_iterations = int(sys.argv[2]) if len(sys.argv) > 2 else 10
s = m.Solver(prob)
summ = s.solve(o)
print("stage ms: evaluate %.2f eliminate %.2f linear_solve %.2f update %.2f ; linear iterations %d" % (
    summ.time_evaluate_s * 1e3, summ.time_eliminate_s * 1e3, summ.time_linear_solve_s * 1e3, summ.time_update_s * 1e3,
    summ.num_linear_solver_iterations))
print("termination", summ.termination_type, "iterations", summ.num_iterations, "device ms", summ.device_time_s * 1e3)
for t in s.trace():
    print("%3d ok=%d valid=%d cost %.12e change %.3e |g| %.3e step %.3e rho %.3e radius %.3e lin %d" % (
        t.iteration, t.step_is_successful, t.step_is_valid, t.cost, t.cost_change, t.gradient_max_norm, t.step_norm,
        t.relative_decrease, t.trust_region_radius, t.linear_solver_iterations))
