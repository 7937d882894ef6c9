"""Extension X1: the bilinear semantic residual row inside every residual block --
evaluation (residual + its three Jacobian blocks) and the LM solve, CUDA path vs the CPU oracle.

The reference has no implementation of this row (its only hint is the TODO at
/root/reference/src/base/semantic_cost_functions.h:153); the interpolation convention is the reference's
Bitmap::InterpolateBilinear (/root/reference/src/util/bitmap.cc:238-269, known answers of bitmap_test.cc:277-288 in
tests/test_oracle_golden.py).  The oracle's Jacobian is the Jet of the interpolant; the CPU tests below pin it against
central differences of the oracle's own residual."""
import ctypes as C

import numpy as np
import pytest

from util import assert_eval_parity, assert_sem_parity, oracle_evaluate_rows, oracle_solve


def _scene(sba, model="SIMPLE_PINHOLE", refine=False, n_cam=8, n_pt=600, seed=5, weight=0.25, **kw):
    return sba.scene.make_scene(n_cam=n_cam, n_pt=n_pt, mean_track=5.0, model=model, refine_intr=refine, seed=seed,
                                width=320, height=240, semantic="distance", sem_weight=weight, **kw)[0]


# ------------------------------------------------------------------ CPU: the oracle row itself
def test_oracle_counts_three_rows(sba, oracle):
    prob = _scene(sba)
    s = sba.capi.Summary()
    d = prob.desc()
    assert oracle.oracle_ba_counts(C.byref(d), C.byref(s)) == 0
    n = prob.num_observations
    assert s.num_residual_blocks == n and s.num_residuals == 3 * n and s.num_residuals_reduced == 3 * n
    c = prob.counts()  # the product's host-side count (no device needed)
    assert (c.num_residuals, c.num_residuals_reduced) == (3 * n, 3 * n)


def test_oracle_semantic_jacobian_matches_central_differences(sba, oracle):
    """The Jet row against central differences of the oracle's raw third residual, per ambient parameter."""
    prob = _scene(sba, model="SIMPLE_RADIAL", refine=True, n_cam=4, n_pt=80, seed=9)
    prob.image_pose_constant[:] = 0
    prob.image_tvec_constant_mask[:] = 0
    prob.camera_param_constant_mask[:] = 0
    capi = sba.capi
    ref = oracle_evaluate_rows(oracle, capi, prob, apply_loss=False)
    assert np.count_nonzero(ref["sem"]) > 0.8 * len(ref["sem"])
    assert np.count_nonzero(ref["sem_point"]) > 0

    def sem_rows():
        out = np.zeros(prob.num_observations)
        d = prob.desc()
        assert oracle.oracle_ba_sem_rows(C.byref(d), capi.ptr(out, capi._dp)) == 0
        return out
    # raw row == the residual-only restatement (the Jet quotient rounds u = p_x / p_z once differently: ulps)
    assert np.allclose(sem_rows(), ref["sem"], rtol=1e-12, atol=1e-12)
    h = 1e-6
    # points
    for a in range(3):
        base = prob.points.copy()
        prob.points[:, a] = base[:, a] + h; fp = sem_rows()
        prob.points[:, a] = base[:, a] - h; fm = sem_rows()
        prob.points[:, a] = base[:, a]
        num = (fp - fm) / (2 * h)
        ana = ref["sem_point"][:, a]
        ok = np.abs(num - ana) <= 1e-5 * np.maximum(np.abs(ana), 1.0)
        assert ok.mean() > 0.98  # a perturbed sample may cross into the next cell, where the slope changes
    # tvec (tangent == ambient)
    for a in range(3):
        base = prob.image_tvec.copy()
        prob.image_tvec[:, a] = base[:, a] + h; fp = sem_rows()
        prob.image_tvec[:, a] = base[:, a] - h; fm = sem_rows()
        prob.image_tvec[:, a] = base[:, a]
        num = (fp - fm) / (2 * h)
        ana = ref["sem_pose"][:, 3 + a]
        assert (np.abs(num - ana) <= 1e-5 * np.maximum(np.abs(ana), 1.0)).mean() > 0.98
    # focal length (ambient parameter 0)
    base = prob.camera_params.copy()
    hf = 1e-4
    prob.camera_params[:, 0] = base[:, 0] + hf; fp = sem_rows()
    prob.camera_params[:, 0] = base[:, 0] - hf; fm = sem_rows()
    prob.camera_params[:, 0] = base[:, 0]
    num = (fp - fm) / (2 * hf)
    ana = ref["sem_intr"][:, 0]
    assert (np.abs(num - ana) <= 1e-5 * np.maximum(np.abs(ana), 1e-3)).mean() > 0.98


def test_oracle_solve_with_semantic_row_decreases_cost(sba, oracle):
    prob = _scene(sba, n_cam=6, n_pt=300)
    o = sba.capi.default_solver_options()
    o.max_num_iterations = 8
    rc, s, tr, _ = oracle_solve(oracle, sba.capi, prob, o)
    assert rc == 0 and s.final_cost < 0.9 * s.initial_cost and s.num_successful_steps >= 3
    # the semantic term is really in the objective: weight 0 gives another (smaller) initial cost
    prob.sem_weight = 0.0
    rc, s0, _, _ = oracle_solve(oracle, sba.capi, prob, o)
    assert s0.initial_cost < s.initial_cost


# ------------------------------------------------------------------ GPU: CUDA path vs oracle
def _eval_both(sba, oracle, prob, loss=0, scale=1.0, apply_loss=True):
    s = sba.Solver(prob)
    ev = s.evaluate(loss_type=loss, loss_scale=scale, apply_loss=apply_loss,
                    want=("r", "pose", "intr", "point", "sem", "sem_jac"))
    s.close()
    ref = oracle_evaluate_rows(oracle, sba.capi, prob, loss_type=loss, loss_scale=scale, apply_loss=apply_loss)
    return ev, ref


@pytest.mark.gpu
@pytest.mark.parametrize("model,refine", [("SIMPLE_PINHOLE", False), ("SIMPLE_PINHOLE", True), ("SIMPLE_RADIAL", True)])
def test_semantic_row_eval_parity(sba, oracle, model, refine):
    """r_sem and its pose / intrinsics / point Jacobian blocks <= 1e-10 vs the Jet oracle (dc = 6, 10, 8)."""
    prob = _scene(sba, model=model, refine=refine)
    ev, ref = _eval_both(sba, oracle, prob)
    assert np.count_nonzero(ref["sem"]) > 0.5 * len(ref["sem"]) and np.count_nonzero(ref["sem_pose"]) > 0
    assert_eval_parity(ev, ref, prob)
    assert_sem_parity(ev, ref)
    # raw functor outputs (CostFunction::Evaluate level)
    ev, ref = _eval_both(sba, oracle, prob, apply_loss=False)
    assert_eval_parity(ev, ref, prob)
    assert_sem_parity(ev, ref)


@pytest.mark.gpu
@pytest.mark.parametrize("loss", [1, 2])
def test_semantic_row_robust_loss(sba, oracle, loss):
    """The corrector acts on the squared norm of all three rows."""
    prob = _scene(sba, model="SIMPLE_RADIAL", refine=True, seed=6, weight=0.5)
    ev, ref = _eval_both(sba, oracle, prob, loss=loss, scale=1.5)
    assert_eval_parity(ev, ref, prob)
    assert_sem_parity(ev, ref)


@pytest.mark.gpu
def test_semantic_row_all_models_and_constant_blocks(sba, oracle):
    """Every camera model (dc tiers 8..18), samples that fall outside the map (zero row), constant poses / points."""
    capi = sba.capi
    for model in capi.CAMERA_MODELS:
        base = _scene(sba, model="SIMPLE_PINHOLE", n_cam=5, n_pt=200, seed=7)
        mid = capi.CAMERA_MODELS[model]
        P = capi.MODEL_NUM_PARAMS[mid]
        prm = np.zeros((5, 12))
        f = 1.2 * 320
        two = model not in ("SIMPLE_PINHOLE", "SIMPLE_RADIAL", "RADIAL", "SIMPLE_RADIAL_FISHEYE", "RADIAL_FISHEYE")
        prm[:, :4 if two else 3] = [f, f, 160.0, 120.0] if two else [f, 160.0, 120.0]
        prm[:, (4 if two else 3):P] = 1e-3 * (1 + np.arange(P - (4 if two else 3)))
        if model == "FOV":
            prm[:, 4] = 0.3
        prob = sba.FlatProblem(base.image_qvec, base.image_tvec, base.image_camera, np.full(5, mid), prm, base.points,
                               base.obs_image, base.obs_point, base.obs_xy, image_pose_constant=base.image_pose_constant,
                               image_tvec_constant_mask=base.image_tvec_constant_mask, sem_maps=base.sem_maps,
                               sem_weight=0.3)
        prob.point_constant[::7] = 1
        prob.image_pose_constant[3] = 1
        ev, ref = _eval_both(sba, oracle, prob)
        assert_eval_parity(ev, ref, prob)
        assert_sem_parity(ev, ref)
    # a map smaller than the image: many projections are rejected -> exactly zero residual and row on both sides
    small = base.sem_maps[:, :100, :140].copy()
    prob = sba.FlatProblem(base.image_qvec, base.image_tvec, base.image_camera, base.camera_model, base.camera_params,
                           base.points, base.obs_image, base.obs_point, base.obs_xy, sem_maps=small, sem_weight=1.0)
    ev, ref = _eval_both(sba, oracle, prob)
    assert 0.1 < (ref["sem"] == 0).mean() < 0.95
    assert_sem_parity(ev, ref)
    assert ((ev["sem"] == 0) == (ref["sem"] == 0)).all()


def _assert_solve_parity(sba, oracle, prob, o, atol_pts=1e-5):
    rc, so, tro, po = oracle_solve(oracle, sba.capi, prob, o)
    s = sba.Solver(prob)
    sg = s.solve(o)
    trg = s.trace()
    assert rc == 0
    assert sg.initial_cost == pytest.approx(so.initial_cost, rel=1e-12)
    assert abs(sg.final_cost - so.final_cost) <= 1e-6 * so.final_cost
    assert sg.num_residuals == so.num_residuals and sg.num_residuals_reduced == so.num_residuals_reduced
    assert so.num_successful_steps >= 3
    for a, b in zip(trg, tro):
        if abs(b.cost_change) > 1e-9 * b.cost:
            assert a.step_is_successful == b.step_is_successful
            assert a.cost == pytest.approx(b.cost, rel=1e-9)
    X = s.parameters()[3]
    s.close()
    assert np.allclose(X, po[3], rtol=0, atol=atol_pts)
    return sg, so


@pytest.mark.gpu
@pytest.mark.parametrize("solver", [0, 2])
@pytest.mark.parametrize("model,refine", [("SIMPLE_PINHOLE", False), ("SIMPLE_RADIAL", True)])
def test_semantic_row_solve_parity(sba, oracle, solver, model, refine):
    """LM over three-row blocks: same accept / reject sequence, final cost <= 1e-6 (DENSE_SCHUR and ITERATIVE_SCHUR)."""
    prob = _scene(sba, model=model, refine=refine, n_cam=12, n_pt=1500, seed=8)
    o = sba.capi.default_solver_options()
    o.max_num_iterations, o.linear_solver_type = 10, solver
    sg, so = _assert_solve_parity(sba, oracle, prob, o)
    # and the term matters: the same scene without maps ends somewhere else
    bare = sba.FlatProblem(prob.image_qvec, prob.image_tvec, prob.image_camera, prob.camera_model, prob.camera_params,
                           prob.points, prob.obs_image, prob.obs_point, prob.obs_xy,
                           image_pose_constant=prob.image_pose_constant,
                           image_tvec_constant_mask=prob.image_tvec_constant_mask,
                           camera_param_constant_mask=prob.camera_param_constant_mask)
    s = sba.Solver(bare)
    s0 = s.solve(o)
    s.close()
    assert s0.num_residuals * 3 == sg.num_residuals * 2 and s0.final_cost < sg.final_cost


@pytest.mark.gpu
def test_semantic_row_solve_parity_c2(sba, oracle):
    """BASELINE config C2 as named: semantic BA, 50 cams / 20 k pts / ~150 k obs + 640x480 maps, DENSE_SCHUR."""
    prob, _ = sba.scene.make_config("C2")
    assert prob.sem_maps is not None and prob.sem_weight > 0
    o = sba.capi.default_solver_options()
    o.max_num_iterations = 10
    _assert_solve_parity(sba, oracle, prob, o)


@pytest.mark.gpu
def test_semantic_row_solve_parity_c2_bench_scene(sba, oracle):
    """The scene bench.py solves (initial guess perturbed 30x harder): every one of the 10 iterations moves the cost and
    the trust region rejects steps -- the device and the oracle must take the SAME accept / reject sequence."""
    prob, _ = sba.scene.make_config("C2", perturb_scale=30.0)
    o = sba.capi.default_solver_options()
    o.max_num_iterations = 10
    rc, so, tro, _ = oracle_solve(oracle, sba.capi, prob, o)
    s = sba.Solver(prob)
    sg = s.solve(o)
    trg = s.trace()
    s.close()
    assert len(trg) == len(tro) == 11
    assert [t.step_is_successful for t in trg] == [t.step_is_successful for t in tro]
    assert so.num_unsuccessful_steps >= 1 and all(abs(t.cost_change) > 1e-9 * t.cost for t in tro[1:])
    for a, b in zip(trg, tro):
        assert a.cost == pytest.approx(b.cost, rel=1e-8) and a.trust_region_radius == pytest.approx(b.trust_region_radius, rel=1e-6)
    assert abs(sg.final_cost - so.final_cost) <= 1e-6 * so.final_cost


@pytest.mark.gpu
def test_semantic_row_implicit_products_and_robust_solve(sba, oracle, monkeypatch):
    """Implicit Schur products (matvec_point / cam_reduce with three rows) and a Cauchy-loss solve."""
    prob = _scene(sba, model="SIMPLE_RADIAL", refine=True, n_cam=12, n_pt=1500, seed=10)
    o = sba.capi.default_solver_options()
    o.max_num_iterations, o.linear_solver_type = 8, 2
    o.loss_function_type, o.loss_function_scale = 2, 2.0
    monkeypatch.setenv("SBA_IMPLICIT_SCHUR", "1")
    _assert_solve_parity(sba, oracle, prob, o)
    monkeypatch.delenv("SBA_IMPLICIT_SCHUR")
    s = sba.Solver(prob)
    n = s.schur_product()
    x = np.random.default_rng(0).normal(size=n)
    y_imp, y_blk = s.schur_product(x, use_blocks=False), s.schur_product(x, use_blocks=True)
    s.close()
    assert np.linalg.norm(y_imp - y_blk) <= 1e-11 * np.linalg.norm(y_blk)


@pytest.mark.gpu
def test_resident_maps_solve_is_bit_identical(sba):
    """sba_maps_create + sba_create_with_maps (maps uploaded once, problems built over them) against sba_create with the
    maps in the descriptor: the same bits, for two problems over one set, also after the caller's reference is gone."""
    capi = sba.capi
    prob = _scene(sba, model="SIMPLE_RADIAL", refine=True, n_cam=12, n_pt=1500, seed=8)
    o = capi.default_solver_options()
    o.max_num_iterations = 6
    s = sba.Solver(prob)
    ref = s.solve(o)
    ref_params = s.parameters()
    ref_eval = s.evaluate(want=("r", "sem"))
    s.close()
    maps = sba.ResidentMaps(prob.sem_maps)
    a, b = sba.Solver(prob, maps=maps), sba.Solver(prob, maps=maps)
    maps.close()  # the set lives on until the last handle over it is destroyed
    for t in (a, b):
        got = t.solve(o)
        assert (got.final_cost, got.num_residuals, got.num_successful_steps) == (ref.final_cost, ref.num_residuals, ref.num_successful_steps)
        assert all(np.array_equal(x, y) for x, y in zip(t.parameters(), ref_params))
        ev = t.evaluate(want=("r", "sem"))
        assert np.array_equal(ev["sem"], ref_eval["sem"]) and np.array_equal(ev["r"], ref_eval["r"])
        t.close()
    with pytest.raises(ValueError):
        sba.Solver(prob, maps=maps)  # closed
    # error behaviour of the entry point: a second set in the descriptor, a set of another size, a null set
    maps = sba.ResidentMaps(prob.sem_maps)
    h = C.c_void_p()
    d = prob.desc()  # carries sem_maps
    assert capi.load().sba_create_with_maps(C.byref(d), maps._m, C.byref(h)) == capi.ERR_INVALID_ARGUMENT and not h
    assert b"sem_maps must be null" in capi.load().sba_last_error()
    small = sba.ResidentMaps(prob.sem_maps[:5])
    d = prob.desc(with_maps=False)
    assert capi.load().sba_create_with_maps(C.byref(d), small._m, C.byref(h)) == capi.ERR_INVALID_ARGUMENT and not h
    assert capi.load().sba_create_with_maps(C.byref(d), None, C.byref(h)) == capi.ERR_INVALID_ARGUMENT and not h
    m0 = C.c_void_p()
    assert capi.load().sba_maps_create(0, 4, 4, None, C.byref(m0)) == capi.ERR_INVALID_ARGUMENT and not m0
    small.close(); maps.close()


@pytest.mark.gpu
def test_resident_maps_under_point_shards(sba):
    """The multi-GPU layout over a resident map set (every rank keeps the maps of all images):
    each shard's three rows are the unsharded handle's, bit for bit, and the shards partition the observations."""
    prob = _scene(sba, model="SIMPLE_RADIAL", refine=True, n_cam=9, n_pt=900, seed=12)
    s = sba.Solver(prob)
    full = s.evaluate(want=("r", "sem"))
    s.close()
    maps = sba.ResidentMaps(prob.sem_maps)
    covered = np.zeros(prob.num_observations, bool)
    for rank in range(3):
        _, obs_keep = prob.shard_masks(3, rank)
        own = obs_keep.astype(bool)
        sh = sba.Solver(prob, shard_rank=rank, shard_count=3, maps=maps)
        ev = sh.evaluate(want=("r", "sem"))
        sh.close()
        assert np.array_equal(ev["r"][own], full["r"][own]) and np.array_equal(ev["sem"][own], full["sem"][own])
        assert not ev["r"][~own].any() and not ev["sem"][~own].any()
        covered |= own
    maps.close()
    assert covered.all() and np.count_nonzero(full["sem"]) > 0
