"""The blocked device Cholesky (csrc/cholesky.cu: potrf_diag / trsm_panel / DMMA syrk_update / back_sweep) that every
reduced camera system with more than 1024 columns takes -- the SPARSE_SCHUR regime of the reference's solver policy
(/root/reference/src/optim/bundle_adjustment.cc:276-282: 50 < images <= 1000), BASELINE config C3."""
import numpy as np
import pytest

from util import oracle_solve

pytestmark = pytest.mark.gpu


def _parity(sba, oracle, prob, iters=4, solver=-1):
    o = sba.capi.default_solver_options()
    o.max_num_iterations, o.linear_solver_type = iters, solver
    rc, so, tro, po = oracle_solve(oracle, sba.capi, prob, o)
    s = sba.Solver(prob)
    sg = s.solve(o)
    trg = s.trace()
    X = s.parameters()[3]
    s.close()
    assert rc == 0 and sg.linear_solver_type_used == so.linear_solver_type_used
    assert sg.initial_cost == pytest.approx(so.initial_cost, rel=1e-12)
    assert abs(sg.final_cost - so.final_cost) <= 1e-6 * so.final_cost
    assert len(trg) == len(tro)
    for a, b in zip(trg, tro):
        if abs(b.cost_change) > 1e-9 * b.cost:
            assert a.step_is_successful == b.step_is_successful and a.cost == pytest.approx(b.cost, rel=1e-9)
    assert np.allclose(X, po[3], rtol=0, atol=1e-5)
    return sg


@pytest.mark.parametrize("n_cam,model,refine,sem", [
    (172, "SIMPLE_PINHOLE", False, None),       # n = 6 * 171 + ... just above 1024, not a multiple of 64
    (130, "SIMPLE_RADIAL", True, None),         # d_c = 8: n = 1033
    (180, "SIMPLE_PINHOLE", False, "distance"),  # three-row blocks into the same factorisation
])
def test_blocked_cholesky_solve_parity(sba, oracle, n_cam, model, refine, sem):
    prob, _ = sba.scene.make_scene(n_cam=n_cam, n_pt=6000, mean_track=6.0, model=model, refine_intr=refine, seed=13,
                                   width=320, height=240, window=24, semantic=sem)
    sg = _parity(sba, oracle, prob)
    n = sg.num_effective_parameters_reduced - 3 * 6000
    assert n > 1024 and n % 64 != 0 and sg.linear_solver_type_used == sba.capi.SPARSE_SCHUR


def test_c3_shape_solve_parity(sba, oracle):
    """C3's reduced system (500 cameras x 8 = ~4000 columns: 63 block columns of the right-looking factorisation) on a
    fifth of C3's points, so that the CPU oracle's dense Cholesky finishes in seconds."""
    cfg = dict(sba.scene.CONFIGS["C3"]); cfg["n_pt"] = 40000
    prob, _ = sba.scene.make_scene(seed=0, **cfg)
    sg = _parity(sba, oracle, prob, iters=3)
    assert sg.num_effective_parameters_reduced - 3 * 40000 > 3900


def test_blocked_cholesky_reports_failure(sba, oracle):
    """A reduced system that is not positive definite (NaN in S) must set chol_fail -> invalid steps -> FAILURE after
    max_num_consecutive_invalid_steps, exactly like the oracle's factorisation does -- not hang, not return garbage."""
    prob, _ = sba.scene.make_scene(n_cam=180, n_pt=4000, mean_track=6.0, model="SIMPLE_PINHOLE", refine_intr=False,
                                   seed=14, width=320, height=240, window=24)
    prob.obs_xy[prob.num_observations // 2, 0] = np.nan
    o = sba.capi.default_solver_options()
    o.max_num_iterations, o.max_num_consecutive_invalid_steps = 20, 4
    rc, so, tro, _ = oracle_solve(oracle, sba.capi, prob, o)
    s = sba.Solver(prob)
    sg = s.solve(o)
    s.close()
    assert so.termination_type == sba.capi.FAILURE and sg.termination_type == sba.capi.FAILURE
    assert sg.num_unsuccessful_steps == so.num_unsuccessful_steps == 4
