"""RefineAbsolutePose (/root/reference/src/estimators/pose.cc:198-322) through the device
solve: the one-image problem with constant points, Cauchy loss and the reference's camera parameterisation, against
the CPU oracle's LM on the same flattened problem.  "Parity unpinned" like every LM trajectory:
the reference has no test that asserts a refined pose."""
import numpy as np
import pytest

from util import oracle_solve


def correspondences(sba, n=200, seed=0, outliers=20, model="SIMPLE_RADIAL"):
    rng = np.random.default_rng(seed)
    params = {"SIMPLE_RADIAL": [800.0, 500.0, 500.0, 0.02], "SIMPLE_PINHOLE": [800.0, 500.0, 500.0],
              "OPENCV": [800.0, 820.0, 500.0, 500.0, 0.02, -0.01, 0.001, 0.002]}[model]
    cam = sba.Camera(1, model, params, 1000, 1000)
    q = np.array([0.98, 0.05, -0.12, 0.08]); q /= np.linalg.norm(q)
    t = np.array([0.2, -0.1, 0.5])
    R = sba.scene.rotmat_from_quat(q[None])[0]
    Xc = np.column_stack([rng.uniform(-2, 2, n), rng.uniform(-2, 2, n), rng.uniform(4, 9, n)])
    X = (Xc - t) @ R            # world points: R X + t = Xc
    # exact projections through the device's own per-observation pass are not needed: a pinhole + radial model by hand
    u, v = Xc[:, 0] / Xc[:, 2], Xc[:, 1] / Xc[:, 2]
    if model == "SIMPLE_PINHOLE":
        x, y = params[0] * u + params[1], params[0] * v + params[2]
    elif model == "SIMPLE_RADIAL":
        r2 = u * u + v * v
        x, y = params[0] * u * (1 + params[3] * r2) + params[1], params[0] * v * (1 + params[3] * r2) + params[2]
    else:
        k1, k2, p1, p2 = params[4:]
        r2 = u * u + v * v
        rad = k1 * r2 + k2 * r2 * r2
        du = u * rad + 2 * p1 * u * v + p2 * (r2 + 2 * u * u)
        dv = v * rad + 2 * p2 * u * v + p1 * (r2 + 2 * v * v)
        x, y = params[0] * (u + du) + params[2], params[1] * (v + dv) + params[3]
    xy = np.column_stack([x, y]) + rng.normal(0, 0.7, (n, 2))
    mask = np.ones(n, bool)
    bad = rng.choice(n, outliers, replace=False)
    xy[bad] += rng.normal(0, 60, (outliers, 2))
    mask[bad[: outliers // 2]] = False       # half the gross outliers are flagged, half are left to the Cauchy loss
    q0 = q + rng.normal(0, 0.01, 4)
    t0 = t + rng.normal(0, 0.05, 3)
    cam.params[0] *= 1.03
    return mask, xy, X, q0, t0, cam, (q, t)


def test_flatten_structure(sba):
    est = sba.estimators
    mask, xy, X, q0, t0, cam, _ = correspondences(sba)
    n_in = int(mask.sum())
    for rf, re, eff in ((True, True, 6 + 2), (True, False, 6 + 1), (False, True, 6 + 1), (False, False, 6)):
        prob = est.flatten(est.AbsolutePoseRefinementOptions(refine_focal_length=rf, refine_extra_params=re), mask, xy, X, q0, t0, cam)
        c = prob.counts()
        assert prob.num_observations == n_in and c.num_residuals_reduced == 2 * n_in
        assert c.num_effective_parameters_reduced == eff, (rf, re, c.num_effective_parameters_reduced)
        assert prob.point_constant.all() and np.array_equal(prob.obs_xy, xy[mask])
    with pytest.raises(ValueError):
        est.flatten(est.AbsolutePoseRefinementOptions(), mask[:-1], xy, X, q0, t0, cam)
    with pytest.raises(ValueError):
        est.AbsolutePoseRefinementOptions(loss_function_scale=-1.0).Check()
    # no inlier: nothing to solve, the (empty) solution is usable
    assert est.RefineAbsolutePose(est.AbsolutePoseRefinementOptions(), np.zeros(len(mask), bool), xy, X, q0.copy(), t0.copy(), cam) is True


@pytest.mark.gpu
@pytest.mark.parametrize("model,rf,re", [("SIMPLE_RADIAL", True, True), ("SIMPLE_RADIAL", False, False), ("SIMPLE_PINHOLE", True, False),
                                         ("OPENCV", True, True)])
def test_refine_absolute_pose_matches_oracle(sba, oracle, model, rf, re):
    est, capi = sba.estimators, sba.capi
    mask, xy, X, q0, t0, cam, (q_true, t_true) = correspondences(sba, model=model, seed=3)
    opt = est.AbsolutePoseRefinementOptions(refine_focal_length=rf, refine_extra_params=re, print_summary=False)
    prob = est.flatten(opt, mask, xy, X, q0, t0, cam)
    rc, s_ref, tr_ref, (q_ref, t_ref, cam_ref, _) = oracle_solve(oracle, capi, prob, est.solver_options(opt))
    assert rc == 0
    q, t = q0.copy(), t0.copy()
    params0 = cam.params.copy()
    out = []
    assert est.RefineAbsolutePose(opt, mask, xy, X, q, t, cam, summary_out=out) is True
    s = out[0]
    assert s.termination_type == s_ref.termination_type == capi.CONVERGENCE
    assert s.num_iterations == s_ref.num_iterations and s.num_successful_steps == s_ref.num_successful_steps
    assert abs(s.final_cost - s_ref.final_cost) <= 1e-9 * s_ref.final_cost and s.final_cost < s.initial_cost
    assert np.allclose(q, q_ref[0], rtol=0, atol=1e-9) and np.allclose(t, t_ref[0], rtol=0, atol=1e-8)
    assert np.allclose(cam.params, cam_ref[0, :cam.NumParams()], rtol=1e-9, atol=1e-10)
    # the principal point never moves; focal length / extra parameters only when asked (pose.cc:249-281)
    pp = list(cam.PrincipalPointIdxs())
    assert np.array_equal(cam.params[pp], params0[pp])
    if not rf:
        assert np.array_equal(cam.params[list(cam.FocalLengthIdxs())], params0[list(cam.FocalLengthIdxs())])
    elif model != "OPENCV":
        assert abs(cam.params[0] - 800.0) < abs(params0[0] - 800.0)
    if not re and len(cam.ExtraParamsIdxs()):
        assert np.array_equal(cam.params[list(cam.ExtraParamsIdxs())], params0[list(cam.ExtraParamsIdxs())])
    if not rf:
        return  # with the focal length held 3 % off, the pose compensates: no claim about the distance to the truth
    # the refined pose is closer to the truth than the start
    qn = q / np.linalg.norm(q)
    assert min(np.linalg.norm(qn - q_true), np.linalg.norm(qn + q_true)) < 0.5 * np.linalg.norm(q0 / np.linalg.norm(q0) - q_true)
    assert np.linalg.norm(t - t_true) < np.linalg.norm(t0 - t_true)
