"""Committed fixtures of tests/golden/ (written by tests/golden/make_golden.py).

CPU : the oracle reproduces the reference's transcribed known answers (reference_known_answers.json) and its own
      recorded outputs (oracle_scenes.npz) -- the latter pins the checker against accidental change.
GPU : the CUDA path through the C ABI against the same recorded outputs, with the tolerances of SURVEY.md section 8:
      indexing bit-exact, r / J <= 1e-10 relative, final cost <= 1e-6 relative after the same iteration count."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from util import assert_eval_parity, oracle_evaluate, oracle_solve

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SCENES = ["radial_dense", "pinhole_const_intr_iterative", "radial_cauchy_dense"]
dp = C.POINTER(C.c_double)


@pytest.fixture(scope="module")
def known():
    with open(os.path.join(GOLD, "reference_known_answers.json")) as f:
        return json.load(f)


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLD, "oracle_scenes.npz"))


def _scene(sba, gold, name):
    kw = json.loads(bytes(gold[name + "/kwargs"]).decode())
    prob, _ = sba.scene.make_scene(**kw)
    loss, scale, solver = gold[name + "/solver"]
    # a7: the generator's residual order is part of the fixture -- bit-exact
    assert np.array_equal(prob.obs_image, gold[name + "/obs_image"]) and np.array_equal(prob.obs_point, gold[name + "/obs_point"])
    return prob, int(loss), float(scale), int(solver)


def _gold_eval(gold, name):
    ref = {k: gold[name + "/" + k] for k in ("r", "pose", "intr", "point")}
    ref["cost"] = float(gold[name + "/cost"][0])
    return ref


def test_known_answers_of_the_reference_functor_tests(oracle, known):
    kc = known["cost_functions"]
    f = lambda v: np.ascontiguousarray(v, dtype=np.float64)
    for case in kc["cases"]:
        a = [f(kc["qvec"]), f(kc["tvec"]), f(case["X"]), f(case["params"]), f(kc["obs"])]
        r = np.zeros(2)
        oracle.oracle_eval_observation(0, *[v.ctypes.data_as(dp) for v in a], r.ctypes.data_as(dp), None, None, None, None)
        assert r.tolist() == case["residual"]
        g = kc["rig"]
        a = [f(g["rig_qvec"]), f(g["rig_tvec"]), f(g["rel_qvec"]), f(g["rel_tvec"]), f(case["X"]), f(case["params"]), f(kc["obs"])]
        oracle.oracle_eval_rig_observation(0, *[v.ctypes.data_as(dp) for v in a], r.ctypes.data_as(dp))
        assert r.tolist() == case["residual"]


def test_known_answers_bilinear(oracle, known):
    kb = known["bilinear"]
    img = np.zeros((kb["height"], kb["width"]), np.float32)
    for x, y, v in kb["set"]:
        img[y, x] = v
    ok = C.c_int32()
    for x, y, v in kb["cases"]:
        got = oracle.oracle_bilinear(img.ctypes.data_as(C.POINTER(C.c_float)), kb["height"], kb["width"], x, y, C.byref(ok))
        assert ok.value == 1 and got == v


def test_known_round_trips(oracle, known):
    km = known["camera_models"]
    for model, prm in km["params"]:
        p = np.zeros(12); p[:len(prm)] = prm
        for u in np.arange(-0.5, 0.51, 0.1):
            for v in np.arange(-0.5, 0.51, 0.1):
                x, y, u2, v2 = C.c_double(), C.c_double(), C.c_double(), C.c_double()
                oracle.oracle_world_to_image(model, p.ctypes.data_as(dp), u, v, C.byref(x), C.byref(y))
                oracle.oracle_image_to_world(model, p.ctypes.data_as(dp), x.value, y.value, C.byref(u2), C.byref(v2))
                assert abs(u - u2.value) < km["round_trip_tolerance"] and abs(v - v2.value) < km["round_trip_tolerance"]


@pytest.mark.parametrize("name", SCENES)
def test_oracle_reproduces_its_recorded_outputs(sba, oracle, gold, name):
    prob, loss, scale, solver = _scene(sba, gold, name)
    ev = oracle_evaluate(oracle, sba.capi, prob, loss_type=loss, loss_scale=scale, threads=1)
    assert_eval_parity(ev, _gold_eval(gold, name), prob, tol=1e-13)
    o = sba.capi.default_solver_options()
    o.max_num_iterations, o.linear_solver_type, o.loss_function_type, o.loss_function_scale = 10, solver, loss, scale
    rc, s, trace, _ = oracle_solve(oracle, sba.capi, prob, o, threads=1)
    assert rc == 0 and s.final_cost == pytest.approx(float(gold[name + "/cost"][2]), rel=1e-10)
    # accept / reject sequence: identical while the steps still move the cost; at the noise floor (|cost change| below
    # 1e-9 of the cost) the sign of a 1e-16 difference decides and is not a property of the algorithm
    gc, gok = gold[name + "/trace_cost"], gold[name + "/trace_ok"].tolist()
    for k, t in enumerate(trace[:len(gc)]):
        if k and abs(gc[k] - gc[k - 1]) > 1e-9 * gc[k]:
            assert t.step_is_successful == gok[k] and t.cost == pytest.approx(gc[k], rel=1e-10)


@pytest.mark.gpu
@pytest.mark.parametrize("name", SCENES)
def test_cuda_matches_recorded_outputs(sba, gold, name):
    prob, loss, scale, solver = _scene(sba, gold, name)
    s = sba.Solver(prob)
    ev = s.evaluate(loss_type=loss, loss_scale=scale)
    assert_eval_parity(ev, _gold_eval(gold, name), prob, tol=1e-10)
    o = sba.capi.default_solver_options()
    o.max_num_iterations, o.linear_solver_type, o.loss_function_type, o.loss_function_scale = 10, solver, loss, scale
    summ = s.solve(o)
    _, c0, c1 = gold[name + "/cost"]
    assert summ.initial_cost == pytest.approx(c0, rel=1e-12)
    assert abs(summ.final_cost - c1) <= 1e-6 * c1
    tr = s.trace()
    gc, gok, glin = gold[name + "/trace_cost"], gold[name + "/trace_ok"], gold[name + "/trace_lin_it"]
    for k, t in enumerate(tr[:len(gc)]):
        if k and abs(gc[k] - gc[k - 1]) > 1e-9 * gc[k]:
            assert t.step_is_successful == gok[k] and t.cost == pytest.approx(gc[k], rel=1e-9)
            assert t.linear_solver_iterations == glin[k]
    assert np.allclose(s.parameters()[3], gold[name + "/points"], rtol=0, atol=1e-5)
    s.close()
