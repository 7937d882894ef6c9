"""High-precision cross-check of the oracle (last row): the residual of
BundleAdjustmentCostFunction (/root/reference/src/base/cost_functions.h:57-81) for all 11 camera models
(/root/reference/src/base/camera_models.h:614-1480), written a third time -- in mpmath at 50 digits, straight from the
reference's formulas -- and differentiated by central differences at that precision (step 1e-25: truncation and rounding
both below 1e-24).  The oracle's double-precision Jets (oracle/jet.h + projection.h) must agree with it to double
rounding, for the reference's own per-model parameter vectors (camera_models_test.cc:133-218).  The device kernels are
held to the Jets at 1e-10 by the GPU suite, so this closes the chain formula -> Jet -> kernel."""
import ctypes as C

import numpy as np
import pytest
from mpmath import mp, mpf, sqrt, atan, tan

from test_oracle_golden import MODEL_PARAMS, _eval

mp.dps = 50
EPS = mpf(2) ** -52  # std::numeric_limits<double>::epsilon()


def rotate(q, X):  # ceres::UnitQuaternionRotatePoint: the polynomial, no normalisation (Ceres 2.1 rotation.h)
    q0, q1, q2, q3 = q
    X0, X1, X2 = X
    return (X0 + 2 * ((-q2 * q2 - q3 * q3) * X0 + (q1 * q2 - q0 * q3) * X1 + (q0 * q2 + q1 * q3) * X2),
            X1 + 2 * ((q0 * q3 + q1 * q2) * X0 + (-q1 * q1 - q3 * q3) * X1 + (q2 * q3 - q0 * q1) * X2),
            X2 + 2 * ((q1 * q3 - q0 * q2) * X0 + (q0 * q1 + q2 * q3) * X1 + (-q1 * q1 - q2 * q2) * X2))


def fisheye(u, v, ks):  # theta_d / r scaling of the fisheye models (camera_models.h:963-990, 1272-1290, 1348-1368)
    r = sqrt(u * u + v * v)
    if not r > EPS:
        return mpf(0), mpf(0)
    th = atan(r)
    poly = mpf(1)
    for n, k in enumerate(ks):
        poly += k * th ** (2 * (n + 1))
    return u * th * poly / r - u, v * th * poly / r - v


def world_to_image(model, p, u, v):
    if model == 0:  # SIMPLE_PINHOLE f, cx, cy (:615-625)
        return p[0] * u + p[1], p[0] * v + p[2]
    if model == 1:  # PINHOLE (:664-675)
        return p[0] * u + p[2], p[1] * v + p[3]
    if model in (2, 3):  # SIMPLE_RADIAL k / RADIAL k1, k2 (:715-760, :784-830)
        r2 = u * u + v * v
        radial = p[3] * r2 + (p[4] * r2 * r2 if model == 3 else 0)
        return p[0] * (u + u * radial) + p[1], p[0] * (v + v * radial) + p[2]
    if model == 4:  # OPENCV k1, k2, p1, p2 (:854-905)
        k1, k2, p1, p2 = p[4:8]
        u2, uv, v2 = u * u, u * v, v * v
        r2 = u2 + v2
        radial = k1 * r2 + k2 * r2 * r2
        du = u * radial + 2 * p1 * uv + p2 * (r2 + 2 * u2)
        dv = v * radial + 2 * p2 * uv + p1 * (r2 + 2 * v2)
        return p[0] * (u + du) + p[2], p[1] * (v + dv) + p[3]
    if model == 5:  # OPENCV_FISHEYE k1..k4 (:929-990)
        du, dv = fisheye(u, v, p[4:8])
        return p[0] * (u + du) + p[2], p[1] * (v + dv) + p[3]
    if model == 6:  # FULL_OPENCV k1, k2, p1, p2, k3, k4, k5, k6 (:1024-1082)
        k1, k2, p1, p2, k3, k4, k5, k6 = p[4:12]
        u2, uv, v2 = u * u, u * v, v * v
        r2 = u2 + v2
        r4 = r2 * r2
        r6 = r4 * r2
        radial = (1 + k1 * r2 + k2 * r4 + k3 * r6) / (1 + k4 * r2 + k5 * r4 + k6 * r6)
        du = u * radial + 2 * p1 * uv + p2 * (r2 + 2 * u2) - u
        dv = v * radial + 2 * p2 * uv + p1 * (r2 + 2 * v2) - v
        return p[0] * (u + du) + p[2], p[1] * (v + dv) + p[3]
    if model == 7:  # FOV omega (:1105-1170): x = f * u * factor + c (the distortion is multiplicative)
        omega = p[4]
        radius2, omega2 = u * u + v * v, omega * omega
        if omega2 < mpf("1e-4"):
            factor = (omega2 * radius2) / 3 - omega2 / 12 + 1
        elif radius2 < mpf("1e-4"):
            th = tan(omega / 2)
            factor = (-2 * th * (4 * radius2 * th * th - 3)) / (3 * omega)
        else:
            radius = sqrt(radius2)
            factor = atan(radius * 2 * tan(omega / 2)) / (radius * omega)
        return p[0] * u * factor + p[2], p[1] * v * factor + p[3]
    if model in (8, 9):  # SIMPLE_RADIAL_FISHEYE k / RADIAL_FISHEYE k1, k2 (:1240-1290, :1316-1368)
        du, dv = fisheye(u, v, p[3:4] if model == 8 else p[3:5])
        return p[0] * (u + du) + p[1], p[0] * (v + dv) + p[2]
    if model == 10:  # THIN_PRISM_FISHEYE k1, k2, p1, p2, k3, k4, sx1, sy1 (:1405-1480)
        r = sqrt(u * u + v * v)
        if r > EPS:
            th = atan(r)
            uu, vv = th * u / r, th * v / r
        else:
            uu, vv = u, v
        k1, k2, p1, p2, k3, k4, sx1, sy1 = p[4:12]
        u2, uv, v2 = uu * uu, uu * vv, vv * vv
        r2 = u2 + v2
        r4 = r2 * r2
        r6 = r4 * r2
        r8 = r6 * r2
        radial = k1 * r2 + k2 * r4 + k3 * r6 + k4 * r8
        du = uu * radial + 2 * p1 * uv + p2 * (r2 + 2 * u2) + sx1 * r2
        dv = vv * radial + 2 * p2 * uv + p1 * (r2 + 2 * v2) + sy1 * r2
        return p[0] * (uu + du) + p[2], p[1] * (vv + dv) + p[3]
    raise AssertionError(model)


def residual(model, z, n_params, xy):
    """z = q(4) | t(3) | X(3) | params(P): cost_functions.h:57-81."""
    q, t, X, p = z[0:4], z[4:7], z[7:10], z[10:10 + n_params]
    pr = rotate(q, X)
    px, py, pz = pr[0] + t[0], pr[1] + t[1], pr[2] + t[2]
    x, y = world_to_image(model, p, px / pz, py / pz)
    return x - xy[0], y - xy[1]


def check_point(oracle, model, params, q, t, X, rtol):
    P = len(params)
    xy = (mpf(400.25), mpf(500.75))
    z = [mpf(float(v)) for v in list(q) + list(t) + list(X) + list(params)]
    r_hp = residual(model, z, P, xy)
    h = mpf("1e-25")
    J_hp = np.zeros((2, 10 + P))
    for j in range(10 + P):
        zp, zm = list(z), list(z)
        zp[j] += h
        zm[j] -= h
        a, b = residual(model, zp, P, xy), residual(model, zm, P, xy)
        J_hp[0, j], J_hp[1, j] = float((a[0] - b[0]) / (2 * h)), float((a[1] - b[1]) / (2 * h))
    r, Jq, Jt, JX, Jk = _eval(oracle, model, q, t, X, params, (400.25, 500.75), jac=True)
    J = np.hstack([Jq, Jt, JX, Jk])
    scale = max(1.0, float(np.abs(J_hp).max()))
    assert abs(r[0] - float(r_hp[0])) <= rtol * max(1.0, abs(float(r_hp[0]))) and abs(r[1] - float(r_hp[1])) <= rtol * max(1.0, abs(float(r_hp[1])))
    assert np.allclose(J, J_hp, rtol=rtol, atol=rtol * scale), (model, np.abs(J - J_hp).max(), scale)


@pytest.mark.parametrize("model,params", MODEL_PARAMS)
def test_jets_against_fifty_digit_arithmetic(oracle, model, params):
    rng = np.random.default_rng(100 + model)
    for _ in range(6):
        q = rng.normal(size=4)
        q /= np.linalg.norm(q)
        t = rng.uniform(-0.5, 0.5, 3) + [0, 0, 4.0]
        X = rng.uniform(-1.0, 1.0, 3)
        check_point(oracle, model, params, q, t, X, rtol=2e-13)
    # a non-unit quaternion: the functor rotates with the polynomial, it does not normalise (cost_functions.h:66)
    check_point(oracle, model, params, 1.1 * q, t, X, rtol=2e-13)


@pytest.mark.parametrize("model,params", [(m, p) for m, p in MODEL_PARAMS if m in (5, 7, 8, 9, 10)])
def test_branch_points_against_fifty_digit_arithmetic(oracle, model, params):
    """The models with a case split: on the optical axis (r = 0: the fisheye models take their identity branch; FOV its
    small-radius series) and just off it."""
    q, t = np.array([1.0, 0, 0, 0]), np.array([0.0, 0, 0])
    for X in ([0.0, 0.0, 2.0], [1e-3, -2e-3, 2.0], [1e-7, 0.0, 2.0]):
        r, *_ = _eval(oracle, model, q, t, X, params, (400.25, 500.75), jac=True)
        z = [mpf(float(v)) for v in list(q) + list(t) + list(X) + list(params)]
        r_hp = residual(model, z, len(params), (mpf(400.25), mpf(500.75)))
        assert abs(r[0] - float(r_hp[0])) <= 1e-12 * max(1.0, abs(float(r_hp[0])))
        assert abs(r[1] - float(r_hp[1])) <= 1e-12 * max(1.0, abs(float(r_hp[1])))


def quat_product(z, w):  # ceres::QuaternionProduct(z, w): Hamilton product, w-first storage
    return (z[0] * w[0] - z[1] * w[1] - z[2] * w[2] - z[3] * w[3], z[0] * w[1] + z[1] * w[0] + z[2] * w[3] - z[3] * w[2],
            z[0] * w[2] - z[1] * w[3] + z[2] * w[0] + z[3] * w[1], z[0] * w[3] + z[1] * w[2] - z[2] * w[1] + z[3] * w[0])


def rig_residual(model, z, n_params, xy):
    """z = rig_q(4) | rig_t(3) | rel_q(4) | rel_t(3) | X(3) | params(P): RigBundleAdjustmentCostFunction,
    cost_functions.h:173-209."""
    rig_q, rig_t, rel_q, rel_t, X, p = z[0:4], z[4:7], z[7:11], z[11:14], z[14:17], z[17:17 + n_params]
    q = quat_product(rel_q, rig_q)
    tr = rotate(rel_q, rig_t)
    t = (tr[0] + rel_t[0], tr[1] + rel_t[1], tr[2] + rel_t[2])
    pr = rotate(q, X)
    px, py, pz = pr[0] + t[0], pr[1] + t[1], pr[2] + t[2]
    x, y = world_to_image(model, p, px / pz, py / pz)
    return x - xy[0], y - xy[1]


@pytest.mark.parametrize("model,params", [MODEL_PARAMS[i] for i in (0, 3, 7, 8, 10, 11, 20, 21)])
def test_rig_jets_against_fifty_digit_arithmetic(oracle, model, params):
    """The rig functor's 2 x 29 autodiff Jacobian (rig pose, relative pose, point, intrinsics padded to 12): the layout of
    sba_rig_evaluate, which the device's analytic rig kernel is held to."""
    rng = np.random.default_rng(200 + model)
    P = len(params)
    dp_ = C.POINTER(C.c_double)
    for _ in range(4):
        qs = rng.normal(size=(2, 4))
        qs /= np.linalg.norm(qs, axis=1, keepdims=True)
        rig_t, rel_t = rng.uniform(-0.5, 0.5, 3) + [0, 0, 4.0], rng.uniform(-0.2, 0.2, 3)
        X = rng.uniform(-1.0, 1.0, 3)
        prm = np.zeros(12)
        prm[:P] = params
        xy = np.array([400.25, 500.75])
        r, J = np.zeros(2), np.zeros((2, 29))
        args = [np.ascontiguousarray(v, dtype=np.float64) for v in (qs[0], rig_t, qs[1], rel_t, X, prm, xy)]
        oracle.oracle_eval_rig_observation_jac(model, *[v.ctypes.data_as(dp_) for v in args], r.ctypes.data_as(dp_), J.ctypes.data_as(dp_))
        z = [mpf(float(v)) for v in list(qs[0]) + list(rig_t) + list(qs[1]) + list(rel_t) + list(X) + list(params)]
        xy_hp = (mpf(400.25), mpf(500.75))
        r_hp = rig_residual(model, z, P, xy_hp)
        h = mpf("1e-25")
        J_hp = np.zeros((2, 17 + P))
        for j in range(17 + P):
            zp, zm = list(z), list(z)
            zp[j] += h
            zm[j] -= h
            a, b = rig_residual(model, zp, P, xy_hp), rig_residual(model, zm, P, xy_hp)
            J_hp[0, j], J_hp[1, j] = float((a[0] - b[0]) / (2 * h)), float((a[1] - b[1]) / (2 * h))
        scale = max(1.0, float(np.abs(J_hp).max()))
        assert abs(r[0] - float(r_hp[0])) <= 2e-13 * max(1.0, abs(float(r_hp[0])))
        assert np.allclose(J[:, :17 + P], J_hp, rtol=2e-13, atol=2e-13 * scale), (model, np.abs(J[:, :17 + P] - J_hp).max())
        assert not J[:, 17 + P:].any()


def test_pose_tangent_jacobian_is_the_derivative_along_the_quaternion_manifold(sba, oracle):
    """ceres::QuaternionManifold (manifold.h): Plus(x, d) = [cos|d|, sin|d| / |d| d] (x) x.  The pose
    block the oracle hands to the solver is 2 x 6 in the tangent layout (rotation increment, translation): its first three
    columns must be d r(Plus(q, d)) / d d at d = 0, taken here at 50 digits through the third writing of the functor."""
    from mpmath import cos as mcos, sin as msin
    from util import oracle_evaluate
    prob, _ = sba.scene.make_scene(n_cam=3, n_pt=12, mean_track=3.0, model="SIMPLE_RADIAL", refine_intr=True, seed=11)
    ev = oracle_evaluate(oracle, sba.capi, prob, 0, 1.0, apply_loss=False, qvec=True)
    P = 4

    def plus(q, d):
        n = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2])
        if n == 0:
            return q
        s = msin(n) / n
        return quat_product((mcos(n), s * d[0], s * d[1], s * d[2]), q)
    h = mpf("1e-25")
    checked = 0
    for k in range(0, prob.num_observations, 5):
        i, pt = int(prob.obs_image[k]), int(prob.obs_point[k])
        c = int(prob.image_camera[i])
        if prob.image_pose_constant[i]:
            continue
        checked += 1
        q = [mpf(float(v)) for v in prob.image_qvec[i]]
        rest = [mpf(float(v)) for v in list(prob.image_tvec[i]) + list(prob.points[pt]) + list(prob.camera_params[c][:P])]
        xy = (mpf(float(prob.obs_xy[k, 0])), mpf(float(prob.obs_xy[k, 1])))
        J_hp = np.zeros((2, 3))
        for j in range(3):
            dplus, dminus = [mpf(0)] * 3, [mpf(0)] * 3
            dplus[j], dminus[j] = h, -h
            a = residual(2, list(plus(q, dplus)) + rest, P, xy)
            b = residual(2, list(plus(q, dminus)) + rest, P, xy)
            J_hp[0, j], J_hp[1, j] = float((a[0] - b[0]) / (2 * h)), float((a[1] - b[1]) / (2 * h))
        scale = max(1.0, float(np.abs(J_hp).max()))
        assert np.allclose(ev["pose"][k][:, :3], J_hp, rtol=1e-12, atol=1e-12 * scale), (k, ev["pose"][k][:, :3], J_hp)
    assert checked >= 3
