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
