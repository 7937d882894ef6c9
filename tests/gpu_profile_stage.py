"""Runs one stage of the hot path a few times on a BASELINE-shaped scene (target of ncu captures)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
m = importlib.import_module("semantic-bundle-adjustment-colmap_b200")
name, stage, repeat = sys.argv[1], sys.argv[2], int(sys.argv[3])
prob, _ = m.scene.make_config(name)
s = m.Solver(prob)
ms, nbytes = s.bench_stage(stage, repeat)
print("%s %s: %.4f ms, %.1f GB/s algorithmic (%d obs)" % (name, stage, ms, nbytes / ms / 1e6, prob.num_observations))
