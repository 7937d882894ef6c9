"""Size-independent properties at BASELINE.json's full sizes (C4: 1778 cameras / 993 k points / 5 M observations, the
headline workload; C3 and C5 -- the north-star configuration -- with the three-row residual).  The oracle needs ~10 s per LM iteration at this size, so parity
here is through properties the domain offers: the cost is a checksum of the residuals, two evaluation kernels and the
allocation-free one must agree on it, the reduced camera matrix is linear and symmetric and its two representations
(implicit products over J, explicit block-sparse S) are the same operator, the order of the caller's observation list
does not matter, and a solve is reproducible bit for bit and never accepts a step that raises the cost."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def c4(sba):
    prob, _ = sba.scene.make_config("C4", perturb_scale=30.0)
    return prob


@pytest.mark.gpu
def test_c4_cost_is_the_checksum_of_the_residuals(sba, c4):
    s = sba.Solver(c4)
    try:
        ev = s.evaluate(want=("r",))
        r = ev["r"]
        assert r.shape == (c4.num_observations, 2) and np.isfinite(r).all()
        checksum = 0.5 * float(np.sum(r.astype(np.longdouble) ** 2))
        assert abs(ev["cost"] - checksum) <= 1e-12 * checksum
        assert abs(s.cost_no_alloc() - checksum) <= 1e-12 * checksum          # the residual-only kernel, static scratch
        # robust loss: rho(s) = a^2 log(1 + s / a^2) per block (Cauchy, ceres loss_function.cc)
        a2 = 4.0
        sq = np.sum(r.astype(np.longdouble) ** 2, axis=1)
        cauchy = 0.5 * float(np.sum(a2 * np.log1p(sq / a2)))
        got = s.evaluate(loss_type=sba.capi.LOSS_CAUCHY, loss_scale=2.0, apply_loss=False, want=())["cost"]
        assert abs(got - cauchy) <= 1e-12 * cauchy
    finally:
        s.close()


@pytest.mark.gpu
def test_c4_observation_order_does_not_matter(sba, c4):
    """The device sorts the caller's list itself: a shuffled list is the same problem (same cost to the last bits of a
    fixed-order sum over the SAME point-sorted order, since the sort is stable only within a point: rounding only)."""
    rng = np.random.default_rng(5)
    order = rng.permutation(c4.num_observations)
    shuffled = sba.FlatProblem(c4.image_qvec, c4.image_tvec, c4.image_camera, c4.camera_model, c4.camera_params, c4.points,
                               c4.obs_image[order], c4.obs_point[order], c4.obs_xy[order],
                               camera_param_constant_mask=c4.camera_param_constant_mask, point_constant=c4.point_constant,
                               image_pose_constant=c4.image_pose_constant, image_tvec_constant_mask=c4.image_tvec_constant_mask)
    a, b = sba.Solver(c4), sba.Solver(shuffled)
    try:
        ea, eb = a.evaluate(want=("r",)), b.evaluate(want=("r",))
        assert abs(ea["cost"] - eb["cost"]) <= 1e-13 * ea["cost"]
        assert np.array_equal(eb["r"], ea["r"][order])                          # residuals come back in the caller's order
    finally:
        a.close(); b.close()


@pytest.mark.gpu
def test_c4_reduced_camera_matrix_is_one_linear_symmetric_operator(sba, c4):
    s = sba.Solver(c4)
    try:
        n = s.schur_product()
        assert n > 10000
        rng = np.random.default_rng(2)
        x, y = rng.normal(size=n), rng.normal(size=n)
        for blocks in (False, True):
            Sx, Sy = s.schur_product(x, use_blocks=blocks), s.schur_product(y, use_blocks=blocks)
            Sxy = s.schur_product(2.0 * x - 3.0 * y, use_blocks=blocks)
            assert np.linalg.norm(Sxy - (2.0 * Sx - 3.0 * Sy)) <= 1e-12 * np.linalg.norm(Sxy)      # linear
            assert abs(y @ Sx - x @ Sy) <= 1e-11 * abs(y @ Sx)                                        # symmetric
            assert x @ Sx > 0.0                                                                       # positive definite with D^2
        assert np.linalg.norm(s.schur_product(x, use_blocks=True) - s.schur_product(x, use_blocks=False)) <= \
            1e-12 * np.linalg.norm(s.schur_product(x, use_blocks=False))                             # the same operator
    finally:
        s.close()


@pytest.mark.gpu
def test_c4_solve_is_reproducible_and_monotone(sba, c4):
    o = sba.capi.default_solver_options()
    o.max_num_iterations = 10
    o.function_tolerance = o.gradient_tolerance = o.parameter_tolerance = 0.0
    s = sba.Solver(c4)
    try:
        first = s.solve(o)
        trace = s.trace()
        q1, t1, cam1, X1 = s.parameters()
        costs = [r.cost for r in trace]
        assert all(b <= a for a, b in zip(costs, costs[1:]))                   # a rejected step keeps the cost, an accepted one lowers it
        assert sum(r.step_is_successful for r in trace) >= 3 and first.final_cost < 0.01 * first.initial_cost
        s.reset()
        again = s.solve(o)                                                       # CUDA-graph replay of the iteration
        q2, t2, cam2, X2 = s.parameters()
        assert again.final_cost == first.final_cost
        assert np.array_equal(X1, X2) and np.array_equal(q1, q2) and np.array_equal(cam1, cam2)
    finally:
        s.close()


@pytest.mark.gpu
def test_c3_three_row_cost_is_the_checksum_of_all_rows(sba):
    prob, _ = sba.scene.make_config("C3")
    s = sba.Solver(prob)
    try:
        ev = s.evaluate(want=("r", "sem"))
        checksum = 0.5 * (float(np.sum(ev["r"].astype(np.longdouble) ** 2)) + float(np.sum(ev["sem"].astype(np.longdouble) ** 2)))
        assert np.count_nonzero(ev["sem"]) > 0.5 * len(ev["sem"])
        assert abs(ev["cost"] - checksum) <= 1e-12 * checksum
        assert abs(s.cost_no_alloc() - checksum) <= 1e-12 * checksum
    finally:
        s.close()


@pytest.mark.gpu
def test_c5_three_rows_and_resident_maps_at_full_size(sba):
    """The north-star configuration (5000 cameras / 2 M points / 15 M observations, 6.1 GB of label maps): the cost is the
    checksum of all three residual rows, two LM iterations never raise it, and a problem built over resident maps
    (sba_maps_create) ends on the same bits as one that carries the maps in its descriptor."""
    prob, _ = sba.scene.make_config("C5")
    o = sba.capi.default_solver_options()
    o.max_num_iterations = 2
    o.function_tolerance = o.gradient_tolerance = o.parameter_tolerance = 0.0
    maps = sba.ResidentMaps(prob.sem_maps)
    s = sba.Solver(prob, maps=maps)
    try:
        ev = s.evaluate(want=("r", "sem"))
        assert ev["r"].shape == (prob.num_observations, 2) and np.isfinite(ev["r"]).all() and np.isfinite(ev["sem"]).all()
        checksum = 0.5 * (float(np.sum(ev["r"].astype(np.longdouble) ** 2)) + float(np.sum(ev["sem"].astype(np.longdouble) ** 2)))
        assert np.count_nonzero(ev["sem"]) > 0.5 * len(ev["sem"])
        assert abs(ev["cost"] - checksum) <= 1e-12 * checksum
        del ev
        a = s.solve(o)
        costs = [r.cost for r in s.trace()]
        assert abs(costs[0] - checksum) <= 1e-12 * checksum
        assert all(y <= x for x, y in zip(costs, costs[1:])) and a.final_cost < a.initial_cost
        assert a.num_residuals == 3 * prob.num_observations
        Xa = s.parameters()[3]
    finally:
        s.close()
        maps.close()
    s = sba.Solver(prob)
    try:
        b = s.solve(o)
        assert b.final_cost == a.final_cost and np.array_equal(s.parameters()[3], Xa)
    finally:
        s.close()
