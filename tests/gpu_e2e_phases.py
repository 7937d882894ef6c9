"""Host-side phase times of one end-to-end C-ABI round trip (create / solve / read back / destroy)."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
m = importlib.import_module("semantic-bundle-adjustment-colmap_b200")
name = sys.argv[1] if len(sys.argv) > 1 else "C2"
maps = (len(sys.argv) > 2 and sys.argv[2] == "maps")
prob, _ = m.scene.make_config(name, label_maps=maps)
o = m.capi.default_solver_options(); o.max_num_iterations = 10
for it in range(int(os.environ.get('E2E_ROUNDS', '4'))):
    t0 = time.perf_counter(); s = m.Solver(prob)
    t1 = time.perf_counter(); summ = s.solve(o)
    t2 = time.perf_counter(); s.parameters()
    t3 = time.perf_counter(); s.close()
    t4 = time.perf_counter()
    print("iter %d: create %.2f solve %.2f (device %.2f) read %.2f destroy %.2f ms" % (
        it, (t1 - t0) * 1e3, (t2 - t1) * 1e3, summ.device_time_s * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3), flush=True)
