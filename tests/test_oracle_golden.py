"""Pins the CPU oracle against every known answer the reference's own tests hold for this path.

  (1) src/base/cost_functions_test.cc:41-132   exact residuals of the three reprojection functors
  (2) src/base/camera_models_test.cc:41-218    parameter vectors + WorldToImage/ImageToWorld round trips
  (3) src/util/bitmap_test.cc:277-288          bilinear interpolation (extension X1's sampling rule)
plus self-consistency of the oracle's autodiff Jacobians (central differences) -- the reference pins no
Jacobian values anywhere.
"""
import ctypes as C

import numpy as np
import pytest

dp = C.POINTER(C.c_double)


def _p(a):
    return a.ctypes.data_as(dp)


def _eval(oracle, model, q, t, X, prm, xy, jac=False):
    q, t, X, prm, xy = (np.ascontiguousarray(v, dtype=np.float64) for v in (q, t, X, prm, xy))
    r = np.zeros(2)
    if not jac:
        oracle.oracle_eval_observation(model, _p(q), _p(t), _p(X), _p(prm), _p(xy), _p(r), None, None, None, None)
        return r
    P = oracle.oracle_model_num_params(model)
    Jq, Jt, JX, Jk = np.zeros((2, 4)), np.zeros((2, 3)), np.zeros((2, 3)), np.zeros((2, P))
    oracle.oracle_eval_observation(model, _p(q), _p(t), _p(X), _p(prm), _p(xy), _p(r), _p(Jq), _p(Jt), _p(JX), _p(Jk))
    return r, Jq, Jt, JX, Jk


# (1) -----------------------------------------------------------------------------------------
KNOWN = [  # (point3D, focal) -> residual; cost_functions_test.cc:49-68
    ((0, 0, 1), 1, (0, 0)), ((0, 1, 1), 1, (0, 1)), ((0, 1, 1), 2, (0, 2)), ((-1, 1, 1), 2, (-2, 2))]


@pytest.mark.parametrize("X,f,expect", KNOWN)
def test_bundle_adjustment_cost_function_known_answers(oracle, X, f, expect):
    # BundleAdjustmentCostFunction and ...ConstantPoseCostFunction share the arithmetic (cost_functions.h:57-81,115-140)
    r = _eval(oracle, 0, (1, 0, 0, 0), (0, 0, 0), X, (f, 0, 0), (0, 0))
    assert r[0] == expect[0] and r[1] == expect[1]


@pytest.mark.parametrize("X,f,expect", KNOWN)
def test_rig_cost_function_known_answers(oracle, X, f, expect):
    # cost_functions_test.cc:101-132: rig t = (0,0,-1), rel t = (0,0,1) cancel
    a = [np.ascontiguousarray(v, dtype=np.float64) for v in
         ((1, 0, 0, 0), (0, 0, -1), (1, 0, 0, 0), (0, 0, 1), X, (f, 0, 0), (0, 0))]
    r = np.zeros(2)
    oracle.oracle_eval_rig_observation(0, *[_p(v) for v in a], _p(r))
    assert r[0] == expect[0] and r[1] == expect[1]


# (2) -----------------------------------------------------------------------------------------
MODEL_PARAMS = [  # camera_models_test.cc:133-218
    (0, [655.123, 386.123, 511.123]),
    (1, [651.123, 655.123, 386.123, 511.123]),
    (2, [651.123, 386.123, 511.123, 0]), (2, [651.123, 386.123, 511.123, 0.1]),
    (3, [651.123, 386.123, 511.123, 0, 0]), (3, [651.123, 386.123, 511.123, 0.1, 0]),
    (3, [651.123, 386.123, 511.123, 0.05, 0]), (3, [651.123, 386.123, 511.123, 0.05, 0.03]),
    (4, [651.123, 655.123, 386.123, 511.123, -0.471, 0.223, -0.001, 0.001]),
    (5, [651.123, 655.123, 386.123, 511.123, -0.471, 0.223, -0.001, 0.001]),
    (6, [651.123, 655.123, 386.123, 511.123, -0.471, 0.223, -0.001, 0.001, 0.001, 0.02, -0.02, 0.001]),
    (7, [651.123, 655.123, 386.123, 511.123, 0.9]), (7, [651.123, 655.123, 386.123, 511.123, 0]),
    (7, [651.123, 655.123, 386.123, 511.123, 1e-6]), (7, [651.123, 655.123, 386.123, 511.123, 1e-2]),
    (8, [651.123, 386.123, 511.123, 0]), (8, [651.123, 386.123, 511.123, 0.1]),
    (9, [651.123, 386.123, 511.123, 0, 0]), (9, [651.123, 386.123, 511.123, 0.1, 0]),
    (9, [651.123, 386.123, 511.123, 0.05, 0]), (9, [651.123, 386.123, 511.123, 0.05, 0.03]),
    (10, [651.123, 655.123, 386.123, 511.123, -0.471, 0.223, -0.001, 0.001, 0.001, 0.02, -0.02, 0.001]),
]


@pytest.mark.parametrize("model,params", MODEL_PARAMS)
def test_camera_model_round_trips(oracle, model, params):
    assert oracle.oracle_model_num_params(model) == len(params)
    prm = np.zeros(12); prm[:len(params)] = params
    x, y, u, v = C.c_double(), C.c_double(), C.c_double(), C.c_double()
    for u0 in np.arange(-0.5, 0.5 + 1e-9, 0.1):        # camera_models_test.cc:110-114
        for v0 in np.arange(-0.5, 0.5 + 1e-9, 0.1):
            oracle.oracle_world_to_image(model, _p(prm), u0, v0, C.byref(x), C.byref(y))
            oracle.oracle_image_to_world(model, _p(prm), x.value, y.value, C.byref(u), C.byref(v))
            assert abs(u.value - u0) < 1e-6 and abs(v.value - v0) < 1e-6
    for x0 in range(0, 801, 50):                         # :116-120
        for y0 in range(0, 801, 50):
            oracle.oracle_image_to_world(model, _p(prm), float(x0), float(y0), C.byref(u), C.byref(v))
            oracle.oracle_world_to_image(model, _p(prm), u.value, v.value, C.byref(x), C.byref(y))
            assert abs(x.value - x0) < 1e-6 and abs(y.value - y0) < 1e-6


def test_simple_pinhole_and_radial_closed_forms(oracle):
    # camera_models.h:614-626, 746-757 evaluated by hand
    x, y = C.c_double(), C.c_double()
    prm = np.zeros(12); prm[:4] = [1200.0, 500.0, 400.0, 0.1]
    oracle.oracle_world_to_image(2, _p(prm), 0.25, -0.5, C.byref(x), C.byref(y))
    r2 = 0.25 ** 2 + 0.5 ** 2
    assert x.value == pytest.approx(1200 * (0.25 + 0.25 * 0.1 * r2) + 500, rel=1e-15)
    assert y.value == pytest.approx(1200 * (-0.5 - 0.5 * 0.1 * r2) + 400, rel=1e-15)
    oracle.oracle_world_to_image(0, _p(prm), 0.25, -0.5, C.byref(x), C.byref(y))
    assert (x.value, y.value) == (1200 * 0.25 + 500, 1200 * -0.5 + 400)


# Jacobians: autodiff-equivalent blocks vs central differences -------------------------------------
@pytest.mark.parametrize("model,params", MODEL_PARAMS)
def test_oracle_jacobians_match_central_differences(oracle, model, params):
    rng = np.random.default_rng(model * 7 + len(params))
    q = rng.normal(size=4); q /= np.linalg.norm(q)
    t = rng.normal(size=3) * 0.2 + np.array([0, 0, 4.0])
    X = rng.uniform(-1, 1, 3)
    prm = np.zeros(12); prm[:len(params)] = params
    xy = np.array([400.0, 500.0])
    r, Jq, Jt, JX, Jk = _eval(oracle, model, q, t, X, prm, xy, jac=True)
    blocks = [(q, Jq), (t, Jt), (X, JX), (prm, Jk)]
    for vec, J in blocks:
        for j in range(J.shape[1]):
            if model == 7 and vec is prm and j == 4 and abs(prm[4] ** 2 - 1e-4) < 1e-8:
                # omega = 1e-2 sits exactly on the reference's branch border (omega^2 < 1e-4,
                # camera_models.h:1147) where its order-3 Taylor branch is discontinuous with the
                # closed form; a central difference would straddle the two branches
                continue
            h = 1e-6 * max(1.0, abs(vec[j]))
            old = vec[j]
            vec[j] = old + h; rp = _eval(oracle, model, q, t, X, prm, xy)
            vec[j] = old - h; rm = _eval(oracle, model, q, t, X, prm, xy)
            vec[j] = old
            fd = (rp - rm) / (2 * h)
            assert np.allclose(J[:, j], fd, rtol=2e-5, atol=2e-5 * max(1.0, np.abs(J).max())), (model, j, J[:, j], fd)


# (3) -----------------------------------------------------------------------------------------
def test_bilinear_known_answers(oracle):
    # bitmap_test.cc:277-288 (red channel): 11x11 zeros, pixel (5,5) = 1
    m = np.zeros((11, 11), np.float32); m[5, 5] = 1.0
    ok = C.c_int32()
    fp = m.ctypes.data_as(C.POINTER(C.c_float))
    assert oracle.oracle_bilinear(fp, 11, 11, 5.0, 5.0, C.byref(ok)) == 1.0 and ok.value == 1
    assert oracle.oracle_bilinear(fp, 11, 11, 5.5, 5.0, C.byref(ok)) == 0.5
    assert oracle.oracle_bilinear(fp, 11, 11, 5.5, 5.5, C.byref(ok)) == 0.25
    oracle.oracle_bilinear(fp, 11, 11, 10.5, 5.0, C.byref(ok))
    assert ok.value == 0  # x1 >= width is rejected, bitmap.cc:249-251


# Ceres Solver's own documentation prints the iteration table of one solve: Powell's singular function
# (docs/source/nnls_tutorial.rst, section "Powell's Function": f1 = x1 + 10 x2, f2 = sqrt(5) (x3 - x4), f3 = (x2 - 2 x3)^2,
# f4 = sqrt(10) (x1 - x4)^2 from x = (3, -1, 0, 1); Solver::Options defaults, DENSE_QR, minimizer_progress_to_stdout).
# Columns transcribed from that page: iter, cost, cost_change, |gradient| (max norm), |step|, tr_ratio, tr_radius.
# No Ceres ran here -- the numbers are the published ones -- but 15 rows of 6 printed quantities leave no room for a
# different trust-region loop: radius update, acceptance test, LM diagonal and its clamp, Jacobi scaling, termination.
POWELL_TABLE = """
   0  1.075000e+02    0.00e+00    1.55e+02   0.00e+00   0.00e+00  1.00e+04
   1  5.036190e+00    1.02e+02    2.00e+01   2.16e+00   9.53e-01  3.00e+04
   2  3.148168e-01    4.72e+00    2.50e+00   6.23e-01   9.37e-01  9.00e+04
   3  1.967760e-02    2.95e-01    3.13e-01   3.08e-01   9.37e-01  2.70e+05
   4  1.229900e-03    1.84e-02    3.91e-02   1.54e-01   9.37e-01  8.10e+05
   5  7.687123e-05    1.15e-03    4.89e-03   7.69e-02   9.37e-01  2.43e+06
   6  4.804625e-06    7.21e-05    6.11e-04   3.85e-02   9.37e-01  7.29e+06
   7  3.003028e-07    4.50e-06    7.64e-05   1.92e-02   9.37e-01  2.19e+07
   8  1.877006e-08    2.82e-07    9.54e-06   9.62e-03   9.37e-01  6.56e+07
   9  1.173223e-09    1.76e-08    1.19e-06   4.81e-03   9.37e-01  1.97e+08
  10  7.333425e-11    1.10e-09    1.49e-07   2.40e-03   9.37e-01  5.90e+08
  11  4.584044e-12    6.88e-11    1.86e-08   1.20e-03   9.37e-01  1.77e+09
  12  2.865573e-13    4.30e-12    2.33e-09   6.02e-04   9.37e-01  5.31e+09
  13  1.791438e-14    2.69e-13    2.91e-10   3.01e-04   9.37e-01  1.59e+10
  14  1.120029e-15    1.68e-14    3.64e-11   1.51e-04   9.37e-01  4.78e+10
"""


def _powell(sba, oracle, max_iterations):
    o = sba.capi.default_solver_options()  # Solver::Options defaults of Ceres 2.1
    o.max_num_iterations = max_iterations
    o.function_tolerance, o.gradient_tolerance, o.parameter_tolerance = 1e-6, 1e-10, 1e-8
    x = (C.c_double * 4)(3.0, -1.0, 0.0, 1.0)
    s = sba.capi.Summary()
    tr = (sba.capi.IterationSummary * 64)()
    n = C.c_int32()
    assert oracle.oracle_powell_solve(C.byref(o), x, C.byref(s), tr, 64, C.byref(n)) == 0
    return s, [tr[i] for i in range(n.value)], list(x)


def test_lm_loop_reproduces_the_iteration_table_ceres_publishes(sba, oracle):
    """The trust-region loop of the oracle (oracle/dense_lm_oracle.cpp SolveDense -- the loop behind the semantic,
    cylinder and rig solves, and equal to the Schur-eliminated loop of ba_oracle.cpp on bundle-adjustment problems,
    tests/test_rig_solve.py) against the one LM trajectory Ceres documents, digit for digit as printed."""
    s, trace, x = _powell(sba, oracle, 100)
    want = [line.split() for line in POWELL_TABLE.strip().splitlines()]
    assert len(trace) == len(want) == 15
    for row, w in zip(trace, want):
        got = ["%d" % row.iteration, "%.6e" % row.cost, "%.2e" % row.cost_change, "%.2e" % row.gradient_max_norm,
               "%.2e" % row.step_norm, "%.2e" % row.relative_decrease, "%.2e" % row.trust_region_radius]
        assert got == w, (got, w)
    # "Gradient tolerance reached. Gradient max norm: 3.642190e-11 <= 1.000000e-10" and the final point of the report
    assert s.termination_type == sba.capi.CONVERGENCE and "%.6e" % trace[-1].gradient_max_norm == "3.642190e-11"
    assert ["%g" % v for v in x] == ["0.000146222", "-1.46222e-05", "2.40957e-05", "2.40957e-05"]
    assert "%.6e" % s.initial_cost == "1.075000e+02" and "%.6e" % s.final_cost == "1.120029e-15"
    assert (s.num_successful_steps, s.num_unsuccessful_steps) == (14, 0)
    # the same page in the documentation of older releases stops one iteration earlier and prints that point
    s13, trace13, x13 = _powell(sba, oracle, 13)
    assert s13.termination_type == sba.capi.NO_CONVERGENCE and len(trace13) == 14
    assert ["%g" % v for v in x13] == ["0.000292189", "-2.92189e-05", "4.79511e-05", "4.79511e-05"]


HELLO_TABLE = """
   0  4.512500e+01    0.00e+00    9.50e+00   0.00e+00   0.00e+00  1.00e+04
   1  4.511598e-07    4.51e+01    9.50e-04   9.50e+00   1.00e+00  3.00e+04
   2  5.012552e-16    4.51e-07    3.17e-08   9.50e-04   1.00e+00  9.00e+04
"""


def test_lm_loop_reproduces_the_hello_world_table(sba, oracle):
    """The other iteration table of the same page (section "Hello World!": r = 10 - x from x = 0.5; the report prints
    "x : 0.5 -> 10", final cost 5.012552e-16, CONVERGENCE): a full step with ratio 1 and the function-tolerance exit."""
    o = sba.capi.default_solver_options()
    o.max_num_iterations = 100
    o.function_tolerance, o.gradient_tolerance, o.parameter_tolerance = 1e-6, 1e-10, 1e-8
    x = (C.c_double * 1)(0.5)
    s = sba.capi.Summary()
    tr = (sba.capi.IterationSummary * 16)()
    n = C.c_int32()
    assert oracle.oracle_helloworld_solve(C.byref(o), x, C.byref(s), tr, 16, C.byref(n)) == 0
    want = [line.split() for line in HELLO_TABLE.strip().splitlines()]
    assert n.value == len(want)
    for row, w in zip([tr[i] for i in range(n.value)], want):
        got = ["%d" % row.iteration, "%.6e" % row.cost, "%.2e" % row.cost_change, "%.2e" % row.gradient_max_norm,
               "%.2e" % row.step_norm, "%.2e" % row.relative_decrease, "%.2e" % row.trust_region_radius]
        assert got == w, (got, w)
    assert s.termination_type == sba.capi.CONVERGENCE and "%g" % x[0] == "10" and "%.6e" % s.final_cost == "5.012552e-16"


@pytest.mark.parametrize("loss,scale", [(1, 1.0), (1, 0.7), (2, 1.0), (2, 2.5)])
def test_loss_and_corrector_against_the_published_definitions(sba, oracle, loss, scale):
    """Ceres' documentation defines the losses this path uses (nnls_modeling, "Instances": SoftLOneLoss rho(s) = 2 (sqrt(1 + s) - 1),
    CauchyLoss rho(s) = log(1 + s), each scaled as a^2 rho(s / a^2)) and the corrector (corrector.h: residual and Jacobian
    are scaled so that the Gauss-Newton model of 1/2 rho(|r|^2) has gradient rho' J^T r; with rho'' <= 0 -- both losses
    are concave -- the correction is the plain sqrt(rho') scaling).  In 50-digit arithmetic from the raw residuals:
    cost = 1/2 sum rho(s), corrected rows = sqrt(rho'(s)) x raw rows, and J_c^T r_c = rho' J^T r per block."""
    from mpmath import mp, mpf, sqrt as msqrt, log as mlog
    from util import oracle_evaluate
    mp.dps = 50
    prob, _ = sba.scene.make_scene(n_cam=4, n_pt=60, mean_track=3.0, model="SIMPLE_RADIAL", refine_intr=True, seed=6)
    raw = oracle_evaluate(oracle, sba.capi, prob, loss, scale, apply_loss=False)
    cor = oracle_evaluate(oracle, sba.capi, prob, loss, scale, apply_loss=True)
    a2 = mpf(scale) ** 2
    total = mpf(0)
    for k in range(prob.num_observations):
        s = mpf(float(raw["r"][k, 0])) ** 2 + mpf(float(raw["r"][k, 1])) ** 2
        if loss == 1:
            rho, d1 = 2 * a2 * (msqrt(1 + s / a2) - 1), 1 / msqrt(1 + s / a2)
        else:
            rho, d1 = a2 * mlog(1 + s / a2), 1 / (1 + s / a2)
        total += rho / 2
        w = float(msqrt(d1))
        assert np.allclose(cor["r"][k], w * raw["r"][k], rtol=1e-13, atol=0)
        for blk in ("pose", "intr", "point"):
            assert np.allclose(cor[blk][k], w * raw[blk][k], rtol=1e-13, atol=1e-300)
            assert np.allclose(cor[blk][k].T @ cor["r"][k], float(d1) * (raw[blk][k].T @ raw["r"][k]), rtol=1e-12, atol=1e-12)
    assert abs(cor["cost"] - float(total)) <= 1e-13 * float(total) and abs(raw["cost"] - float(total)) <= 1e-13 * float(total)
    assert float(total) < 0.5 * float(np.sum(raw["r"] ** 2)) or scale > 2  # the loss does damp these residuals
