"""Explicit block-sparse reduced camera matrix for the iterative solver (csrc/bsr.cu) against the implicit products
(csrc/linalg.cu) it replaces and against the oracle: S is the same matrix either way, so the LM trajectories agree
to rounding.  ITERATIVE_SCHUR + SCHUR_JACOBI as /root/reference/src/optim/bundle_adjustment.cc:283-286 selects."""
import os

import numpy as np
import pytest

from util import oracle_solve

pytestmark = pytest.mark.gpu
ITERATIVE = 2


def _solve(sba, prob, implicit, iters=6, launches=False):
    old = os.environ.get("SBA_IMPLICIT_SCHUR")
    os.environ["SBA_IMPLICIT_SCHUR"] = "1" if implicit else "0"
    if launches:   # CG on the block-sparse S as three launches per iteration (pcg.cu) instead of the persistent kernel
        os.environ["SBA_PCG_LAUNCHES"] = "1"
    try:
        o = sba.capi.default_solver_options()
        o.max_num_iterations, o.linear_solver_type = iters, ITERATIVE
        s = sba.Solver(prob)
        summ = s.solve(o)
        trace, par = s.trace(), [a.copy() for a in s.parameters()]
        try:
            s.bench_stage("matvec_blocks", 1)
            blocks = True
        except Exception:
            blocks = False
        s.close()
    finally:
        os.environ.pop("SBA_PCG_LAUNCHES", None)
        if old is None:
            os.environ.pop("SBA_IMPLICIT_SCHUR", None)
        else:
            os.environ["SBA_IMPLICIT_SCHUR"] = old
    return summ, trace, par, blocks


def _agree(a, b):
    """CG stops on the quadratic-model test (eta = 0.1): a product that differs in the last bits can stop one iteration
    apart, so the two inexact Newton trajectories agree to ~1e-7, not to rounding; the PRODUCTS agree to rounding."""
    (sa, ta, pa, _), (sb, tb, pb, _) = a, b
    assert sa.initial_cost == sb.initial_cost
    assert abs(sa.final_cost - sb.final_cost) <= 1e-6 * sb.final_cost
    for x, y in zip(ta, tb):
        if abs(y.cost_change) > 1e-6 * y.cost:
            assert x.step_is_successful == y.step_is_successful
            assert x.cost == pytest.approx(y.cost, rel=1e-6)
    assert np.allclose(pa[3], pb[3], rtol=0, atol=1e-4)


def _products_agree(sba, prob):
    s = sba.Solver(prob)
    n = s.schur_product()
    rng = np.random.default_rng(11)
    for _ in range(3):
        x = rng.normal(size=n)
        yi, yb = s.schur_product(x, use_blocks=False), s.schur_product(x, use_blocks=True)
        assert np.linalg.norm(yb - yi) <= 1e-12 * np.linalg.norm(yi)
        assert np.abs(yb - yi).max() <= 1e-11 * np.abs(yi).max()
    again = s.schur_product(x, use_blocks=True)
    assert np.array_equal(again, yb)          # fixed summation order: bit-reproducible
    s.close()


def _scene(sba, kind, seed=3):
    kw = dict(n_cam=60, n_pt=6000, mean_track=5.0, window=10, seed=seed)
    if kind == "pinhole_const":
        prob, _ = sba.scene.make_scene(model="SIMPLE_PINHOLE", refine_intr=False, width=640, height=480, **kw)
    else:
        prob, _ = sba.scene.make_scene(**kw)
    if kind == "all_intrinsics_free":   # dc = 10
        prob.camera_param_constant_mask[:] = 0
    if kind == "constant_points":
        prob.point_constant[::7] = 1
    if kind == "duplicates":            # a point seen twice by one image (track element pairs inside an image)
        pick = np.arange(0, prob.num_observations, 11)
        rng = np.random.default_rng(5)
        img = np.concatenate([prob.obs_image, prob.obs_image[pick]])
        pt = np.concatenate([prob.obs_point, prob.obs_point[pick]])
        xy = np.concatenate([prob.obs_xy, prob.obs_xy[pick] + rng.uniform(-1, 1, (len(pick), 2))])
        order = np.argsort(img, kind="stable")
        prob.obs_image, prob.obs_point, prob.obs_xy = (np.ascontiguousarray(img[order]), np.ascontiguousarray(pt[order]),
                                                       np.ascontiguousarray(xy[order]))
    return prob


@pytest.mark.parametrize("kind", ["radial", "pinhole_const", "all_intrinsics_free", "constant_points", "duplicates"])
def test_block_sparse_schur_matches_implicit_products(sba, kind):
    prob = _scene(sba, kind)
    _products_agree(sba, prob)
    explicit, implicit = _solve(sba, prob, False), _solve(sba, prob, True)
    assert explicit[3] and not implicit[3]
    _agree(explicit, implicit)


@pytest.mark.parametrize("kind", ["radial", "pinhole_const", "all_intrinsics_free"])
def test_persistent_pcg_matches_launch_per_step_pcg(sba, kind):
    """Same CG on the same matrix; only the order of the three dot-product sums differs (per-row partials instead of
    one CTA's strided sums), so the iteration counts match and the costs agree far below the solver's own noise."""
    prob = _scene(sba, kind, seed=4)
    one, many = _solve(sba, prob, False), _solve(sba, prob, False, launches=True)
    assert one[3] and many[3]
    _agree(one, many)
    assert [t.linear_solver_iterations for t in one[1]][:3] == [t.linear_solver_iterations for t in many[1]][:3]
    again = _solve(sba, prob, False)
    assert again[0].final_cost == one[0].final_cost and all(np.array_equal(a, b) for a, b in zip(again[2], one[2]))


def test_graph_replay_of_the_iterative_path_is_bit_identical(sba):
    """A re-solved handle replays each LM iteration -- block-sparse S build, the persistent PCG launch, back-substitution,
    step -- as a CUDA graph (csrc/lm.cu); same kernels, same order, so nothing may change by a bit, and the CG iteration
    counts read back after the replay are those of the plain run."""
    prob = _scene(sba, "radial", seed=13)
    o = sba.capi.default_solver_options()
    o.max_num_iterations, o.linear_solver_type = 7, ITERATIVE
    s = sba.Solver(prob)
    first = s.solve(o)
    tr1 = [(t.cost, t.linear_solver_iterations) for t in s.trace()]
    par1 = [a.copy() for a in s.parameters()]
    s.reset()
    second = s.solve(o)
    tr2 = [(t.cost, t.linear_solver_iterations) for t in s.trace()]
    assert first.final_cost == second.final_cost and tr1 == tr2
    assert all(np.array_equal(a, b) for a, b in zip(par1, s.parameters()))
    assert first.num_linear_solver_iterations == second.num_linear_solver_iterations > 0
    s.close()


def test_block_sparse_schur_matches_oracle(sba, oracle):
    prob = _scene(sba, "radial", seed=8)
    o = sba.capi.default_solver_options()
    o.max_num_iterations, o.linear_solver_type = 6, ITERATIVE
    _, so, tro, po = oracle_solve(oracle, sba.capi, prob, o)
    sg, trg, par, blocks = _solve(sba, prob, False)
    assert blocks
    assert abs(sg.final_cost - so.final_cost) <= 1e-6 * so.final_cost
    for a, b in zip(trg, tro):
        if abs(b.cost_change) > 1e-5 * b.cost:   # inexact Newton: see _agree
            assert a.step_is_successful == b.step_is_successful
            assert a.cost == pytest.approx(b.cost, rel=1e-5)
    assert np.allclose(par[3], po[3], rtol=0, atol=1e-3)


def test_shared_intrinsics_keep_the_implicit_products(sba, oracle):
    prob = _scene(sba, "radial", seed=9)
    prob.image_camera[:] = 0            # one camera, intrinsics free: its columns are shared by every image block
    default = _solve(sba, prob, False)
    assert not default[3]
    o = sba.capi.default_solver_options()
    o.max_num_iterations, o.linear_solver_type = 6, ITERATIVE
    _, so, _, _ = oracle_solve(oracle, sba.capi, prob, o)
    assert abs(default[0].final_cost - so.final_cost) <= 1e-6 * so.final_cost


def test_block_product_stage_reports(sba):
    prob = _scene(sba, "radial")
    s = sba.Solver(prob)
    ms, nbytes = s.bench_stage("matvec_blocks", 3)
    ms2, nbytes2 = s.bench_stage("schur_blocks", 3)
    assert ms > 0 and ms2 > 0 and nbytes > 0 and nbytes2 > nbytes
    s.close()
