"""GPU parity over the 11 camera models (reference parameter vectors of
src/base/camera_models_test.cc:133-218), mixed-model problems, shared intrinsics, constant
blocks, duplicate observations and robust losses -- CUDA path vs the CPU oracle."""
import numpy as np
import pytest

from test_oracle_golden import MODEL_PARAMS
from util import assert_eval_parity, oracle_evaluate, oracle_solve

pytestmark = pytest.mark.gpu


def scene_with_model(sba, model, params, n_cam=6, n_pt=300, seed=0, free_all=True):
    prob, _ = sba.scene.make_scene(n_cam=n_cam, n_pt=n_pt, mean_track=4.0, model="SIMPLE_PINHOLE", width=772, height=1022,
                                   refine_intr=True, seed=seed)
    P = len(params)
    prm = np.zeros((n_cam, 12)); prm[:, :P] = params
    rng = np.random.default_rng(seed + 100)
    prm[:, :P] *= 1.0 + 1e-3 * rng.normal(size=(n_cam, P))      # distinct intrinsics per camera
    mask = np.zeros(n_cam, np.uint16) if free_all else np.full(n_cam, sum(1 << j for j in sba.capi.MODEL_PP_IDXS[model]), np.uint16)
    return sba.FlatProblem(prob.image_qvec, prob.image_tvec, prob.image_camera, np.full(n_cam, model), prm, prob.points,
                           prob.obs_image, prob.obs_point, prob.obs_xy, prob.image_pose_constant,
                           prob.image_tvec_constant_mask, mask, prob.point_constant)


@pytest.mark.parametrize("model,params", MODEL_PARAMS)
def test_eval_parity_every_model(sba, oracle, model, params):
    prob = scene_with_model(sba, model, params, seed=model)
    s = sba.Solver(prob)
    ev = s.evaluate(want=("r", "pose", "intr", "point", "qvec"))
    ref = oracle_evaluate(oracle, sba.capi, prob, qvec=True)
    assert_eval_parity(ev, ref, prob)
    assert np.linalg.norm(ev["qvec"] - ref["qvec"]) <= 1e-10 * np.linalg.norm(ref["qvec"])
    s.close()


def test_fisheye_center_branch(sba, oracle):
    """r <= DBL_EPSILON takes the undistorted branch (camera_models.h:970-983): a point on the optical axis."""
    for model, params in [(5, MODEL_PARAMS[9][1]), (8, MODEL_PARAMS[16][1]), (9, MODEL_PARAMS[20][1]), (10, MODEL_PARAMS[21][1]),
                          (7, MODEL_PARAMS[11][1])]:
        q = np.array([[1.0, 0, 0, 0]]); t = np.array([[0.0, 0, 0]])
        X = np.array([[0.0, 0.0, 3.0], [0.3, -0.2, 4.0]])
        prm = np.zeros((1, 12)); prm[0, :len(params)] = params
        prob = sba.FlatProblem(q, t, [0], [model], prm, X, [0, 0], [0, 1], np.array([[380.0, 500.0], [400.0, 520.0]]))
        s = sba.Solver(prob)
        ev = s.evaluate()
        ref = oracle_evaluate(oracle, sba.capi, prob)
        assert_eval_parity(ev, ref, prob)
        s.close()


def test_mixed_models_in_one_problem(sba, oracle):
    base = scene_with_model(sba, 0, MODEL_PARAMS[0][1], n_cam=11, n_pt=600, seed=3)
    models, prm = [], np.zeros((11, 12))
    pick = {0: 0, 1: 1, 2: 3, 3: 7, 4: 8, 5: 9, 6: 10, 7: 11, 8: 16, 9: 20, 10: 21}
    for m in range(11):
        mm, pp = MODEL_PARAMS[pick[m]]
        assert mm == m
        models.append(m); prm[m, :len(pp)] = pp
    prob = sba.FlatProblem(base.image_qvec, base.image_tvec, base.image_camera, models, prm, base.points, base.obs_image,
                           base.obs_point, base.obs_xy, base.image_pose_constant, base.image_tvec_constant_mask,
                           np.zeros(11, np.uint16), base.point_constant)
    s = sba.Solver(prob)
    assert_eval_parity(s.evaluate(), oracle_evaluate(oracle, sba.capi, prob), prob)
    o = sba.capi.default_solver_options(); o.max_num_iterations = 4
    rc, so, tro, _ = oracle_solve(oracle, sba.capi, prob, o)
    sg = s.solve(o)
    assert abs(sg.final_cost - so.final_cost) <= 1e-6 * so.final_cost
    s.close()


def variant(sba, **kw):
    prob, _ = sba.scene.make_scene(n_cam=kw.pop("n_cam", 8), n_pt=kw.pop("n_pt", 600), mean_track=5.0, seed=kw.pop("seed", 4))
    return prob


@pytest.mark.parametrize("solver", [0, 2])
def test_shared_intrinsics(sba, oracle, solver):
    """All images share one camera with free f, k: the camera block of S has many writers."""
    p = variant(sba)
    prob = sba.FlatProblem(p.image_qvec, p.image_tvec, np.zeros(8, np.int32), [2], p.camera_params[:1], p.points, p.obs_image,
                           p.obs_point, p.obs_xy, p.image_pose_constant, p.image_tvec_constant_mask,
                           p.camera_param_constant_mask[:1], p.point_constant)
    o = sba.capi.default_solver_options(); o.max_num_iterations = 6; o.linear_solver_type = solver
    rc, so, tro, po = oracle_solve(oracle, sba.capi, prob, o)
    s = sba.Solver(prob)
    assert_eval_parity(s.evaluate(), oracle_evaluate(oracle, sba.capi, prob), prob)
    sg = s.solve(o)
    assert sg.num_effective_parameters_reduced == so.num_effective_parameters_reduced
    assert abs(sg.final_cost - so.final_cost) <= 1e-6 * so.final_cost
    trg = s.trace()
    for a, b in zip(trg[:4], tro[:4]):
        assert a.cost == pytest.approx(b.cost, rel=1e-8)
    s.close()


@pytest.mark.parametrize("solver", [0, 2])
def test_constant_blocks_and_duplicates(sba, oracle, solver):
    p = variant(sba, seed=9)
    rng = np.random.default_rng(1)
    pt_const = (rng.random(len(p.points)) < 0.2).astype(np.uint8)
    pose_const = p.image_pose_constant.copy(); pose_const[3] = 1
    tmask = p.image_tvec_constant_mask.copy(); tmask[2] = 0b101; tmask[4] = 0b111
    cmask = p.camera_param_constant_mask.copy(); cmask[5] = 0b1111; cmask[6] = 0b0000
    # duplicate observations: a point seen twice by the same image
    dup = np.arange(0, 40)
    obs_image = np.concatenate([p.obs_image, p.obs_image[dup]])
    obs_point = np.concatenate([p.obs_point, p.obs_point[dup]])
    obs_xy = np.concatenate([p.obs_xy, p.obs_xy[dup] + 0.5])
    prob = sba.FlatProblem(p.image_qvec, p.image_tvec, p.image_camera, p.camera_model, p.camera_params, p.points, obs_image,
                           obs_point, obs_xy, pose_const, tmask, cmask, pt_const)
    o = sba.capi.default_solver_options(); o.max_num_iterations = 6; o.linear_solver_type = solver
    rc, so, tro, po = oracle_solve(oracle, sba.capi, prob, o)
    s = sba.Solver(prob)
    assert_eval_parity(s.evaluate(), oracle_evaluate(oracle, sba.capi, prob), prob)
    sg = s.solve(o)
    assert (sg.num_residuals_reduced, sg.num_effective_parameters_reduced) == (so.num_residuals_reduced, so.num_effective_parameters_reduced)
    assert abs(sg.final_cost - so.final_cost) <= 1e-6 * so.final_cost
    q, t, cam, X = s.parameters()
    assert np.array_equal(X[pt_const == 1], prob.points[pt_const == 1])        # constant blocks untouched, bit for bit
    assert np.array_equal(q[3], prob.image_qvec[3]) and np.array_equal(t[4], prob.image_tvec[4])
    assert t[2][0] == prob.image_tvec[2][0] and t[2][2] == prob.image_tvec[2][2] and t[2][1] != prob.image_tvec[2][1]
    assert np.array_equal(cam[5], prob.camera_params[5]) and cam[6][1] != prob.camera_params[6][1]
    s.close()


@pytest.mark.parametrize("loss,solver", [(1, 0), (2, 0), (1, 2)])
def test_solve_parity_robust_loss(sba, oracle, loss, solver):
    prob = variant(sba, seed=12)
    prob.obs_xy[::17] += 40.0  # outliers
    o = sba.capi.default_solver_options()
    o.max_num_iterations = 8; o.loss_function_type = loss; o.loss_function_scale = 2.0; o.linear_solver_type = solver
    rc, so, tro, po = oracle_solve(oracle, sba.capi, prob, o)
    s = sba.Solver(prob)
    sg = s.solve(o)
    assert sg.initial_cost == pytest.approx(so.initial_cost, rel=1e-12)
    assert abs(sg.final_cost - so.final_cost) <= 1e-6 * so.final_cost
    s.close()


def test_no_jacobi_scaling_and_radius_options(sba, oracle):
    prob = variant(sba, seed=13)
    o = sba.capi.default_solver_options()
    o.max_num_iterations = 5; o.jacobi_scaling = 0; o.initial_trust_region_radius = 1e2
    rc, so, tro, po = oracle_solve(oracle, sba.capi, prob, o)
    s = sba.Solver(prob)
    sg = s.solve(o)
    assert abs(sg.final_cost - so.final_cost) <= 1e-6 * so.final_cost
    assert [r.trust_region_radius for r in s.trace()] == pytest.approx([r.trust_region_radius for r in tro], rel=1e-6)
    s.close()


def test_handle_is_one_shot_and_reset(sba):
    prob = variant(sba, seed=14)
    s = sba.Solver(prob)
    o = sba.capi.default_solver_options(); o.max_num_iterations = 2
    a = s.solve(o)
    with pytest.raises(sba.capi.SbaError) as e:
        s.solve(o)
    assert e.value.code == sba.capi.ERR_ALREADY_SOLVED
    s.reset()
    b = s.solve(o)
    assert a.final_cost == b.final_cost  # bit-reproducible run to run
    s.close()
