"""GPU parity tests proper: the CUDA path through the C ABI against the CPU oracle."""
import numpy as np
import pytest

from util import assert_eval_parity, oracle_evaluate, oracle_solve

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("model,refine", [("SIMPLE_RADIAL", True), ("SIMPLE_PINHOLE", False), ("SIMPLE_PINHOLE", True)])
def test_eval_parity_small(sba, oracle, model, refine):
    prob, _ = sba.scene.make_scene(n_cam=8, n_pt=500, mean_track=5.0, model=model, refine_intr=refine, seed=1,
                                   width=640 if model == "SIMPLE_PINHOLE" else 1000, height=480 if model == "SIMPLE_PINHOLE" else 1000)
    s = sba.Solver(prob)
    ev = s.evaluate(want=("r", "pose", "intr", "point", "qvec"))
    ref = oracle_evaluate(oracle, sba.capi, prob, qvec=True)
    assert_eval_parity(ev, ref, prob)
    assert np.linalg.norm(ev["qvec"] - ref["qvec"]) <= 1e-10 * np.linalg.norm(ref["qvec"])
    s.close()


@pytest.mark.parametrize("loss", [1, 2])
def test_eval_parity_robust_loss(sba, oracle, loss):
    prob, _ = sba.scene.make_scene(n_cam=8, n_pt=500, mean_track=5.0, seed=2)
    s = sba.Solver(prob)
    ev = s.evaluate(loss_type=loss, loss_scale=1.5)
    ref = oracle_evaluate(oracle, sba.capi, prob, loss_type=loss, loss_scale=1.5)
    assert_eval_parity(ev, ref, prob)
    s.close()


@pytest.mark.parametrize("solver", [0, 2])
def test_solve_parity_c1(sba, oracle, solver):
    prob, _ = sba.scene.make_config("C1")
    o = sba.capi.default_solver_options()
    o.max_num_iterations = 10
    o.linear_solver_type = solver
    rc, so, tro, po = oracle_solve(oracle, sba.capi, prob, o)
    s = sba.Solver(prob)
    sg = s.solve(o)
    trg = s.trace()
    assert sg.initial_cost == pytest.approx(so.initial_cost, rel=1e-12)
    assert abs(sg.final_cost - so.final_cost) <= 1e-6 * so.final_cost
    # identical accept / reject sequence while the cost is still moving (below 1e-10 relative the sign of the
    # cost change is rounding noise -- even the oracle differs from itself run to run with OpenMP reductions)
    # (either side may stop early once the cost change is exactly 0.0: function tolerance 0 is then met)
    for a, b in zip(trg, tro):
        if abs(b.cost_change) > 1e-9 * b.cost:
            assert a.step_is_successful == b.step_is_successful
            assert a.linear_solver_iterations == b.linear_solver_iterations
            assert a.cost == pytest.approx(b.cost, rel=1e-9)
    assert sg.num_residuals_reduced == so.num_residuals_reduced
    assert sg.num_effective_parameters_reduced == so.num_effective_parameters_reduced
    q, t, cam, X = s.parameters()
    assert np.allclose(X, po[3], rtol=0, atol=1e-5)
    s.close()


def test_graph_replay_is_bit_identical_to_plain_launches(sba, oracle):
    """A handle that is solved again (sba_reset) replays the LM iteration as CUDA graphs (csrc/lm.cu); a long solve
    does so from its second iteration.  Same kernels, same order: costs and parameters must not change by one bit."""
    prob, _ = sba.scene.make_scene(n_cam=12, n_pt=1500, mean_track=5.0, seed=21)
    o = sba.capi.default_solver_options()
    o.max_num_iterations, o.linear_solver_type = 8, 0
    s = sba.Solver(prob)
    first = s.solve(o)                       # fresh handle, short solve: plain launches
    tr1, par1 = [t.cost for t in s.trace()], [a.copy() for a in s.parameters()]
    s.reset()
    second = s.solve(o)                      # graphs
    tr2, par2 = [t.cost for t in s.trace()], s.parameters()
    assert first.final_cost == second.final_cost and tr1 == tr2
    assert all(np.array_equal(a, b) for a, b in zip(par1, par2))
    s.close()
    o.max_num_iterations = 30                # long solve on a fresh handle: graphs from iteration 2
    s = sba.Solver(prob)
    sg = s.solve(o)
    rc, so, tro, _ = oracle_solve(oracle, sba.capi, prob, o)
    assert abs(sg.final_cost - so.final_cost) <= 1e-6 * so.final_cost
    # (iteration counts differ at the noise floor: with tolerances 0 a solve stops when a cost change rounds to exactly 0)
    s.close()


def test_sharded_handles_see_exactly_their_observations(sba):
    """Multi-GPU layout on one device: each shard sorts the WHOLE problem on the device and keeps its contiguous slice
    (csrc/order.cu).  Per-observation residuals / Jacobians do not depend on the sharding: bit-identical to the
    unsharded handle for owned observations, untouched for the others; the shards partition the observations."""
    prob, _ = sba.scene.make_scene(n_cam=9, n_pt=700, mean_track=4.5, seed=31)
    s = sba.Solver(prob)
    full = s.evaluate()
    s.close()
    covered = np.zeros(prob.num_observations, bool)
    for rank in range(3):
        _, obs_keep = prob.shard_masks(3, rank)
        own = obs_keep.astype(bool)
        sh = sba.Solver(prob, shard_rank=rank, shard_count=3)
        ev = sh.evaluate()
        sh.close()
        for k in ("r", "pose", "intr", "point"):
            assert np.array_equal(ev[k][own], full[k][own]), k
            assert not ev[k][~own].any(), k
        assert not (covered & own).any()
        covered |= own
    assert covered.all()
