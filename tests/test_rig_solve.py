"""RigBundleAdjuster as a solve (/root/reference/src/optim/bundle_adjustment.cc:789-1140).

Pinned by the reference's own tests: the structural counts of TestRigTwoView / TestRigFourView / TestConstantRigFourView /
TestRigFourViewPartial (/root/reference/src/optim/bundle_adjustment_test.cc:755-969: 400/316, 800/322, 800/316, 800/328).
The LM trajectory is "parity unpinned" like every other solve: device == CPU restatement, and the
restatement of the rig solve is itself held against the ordinary-BA oracle on problems without rigs."""
import copy
import ctypes as C

import numpy as np
import pytest

from test_filters import oracle_errors
from test_host_logic import generate_reconstruction
from util import oracle_solve


def rig_scene(sba, num_images, rig_layout, seed=0, ref=0, perturb=True):
    """GenerateReconstruction + a rig as the reference's tests build it; rig_layout: list of snapshots (image ids)."""
    rec = generate_reconstruction(sba, num_images, 100, seed=seed)
    for i in range(num_images):
        rec.Image(i).camera_id = i % 2 if num_images > 2 else i
    rig = sba.rig_adjuster.CameraRig()
    rig.AddCamera(0, (1, 0, 0, 0), (0, 0, 0))
    rig.AddCamera(1, (1, 0, 0, 0), (0, 0, 0))
    for snap in rig_layout:
        rig.AddSnapshot(snap)
    rig.SetRefCameraId(ref)
    config = sba.BundleAdjustmentConfig()
    for i in range(num_images):
        config.AddImage(i)
    return rec, [rig], config


def run_oracle(sba, oracle, f, opts, cap=256):
    g = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in f.items()}
    d = sba.rig_adjuster.RigBundleAdjuster.desc(g)
    s = sba.capi.Summary()
    tr = (sba.capi.IterationSummary * cap)()
    n = C.c_int32()
    rc = oracle.oracle_rig_solve(C.byref(d), C.byref(opts), 0, C.byref(s), tr, cap, C.byref(n))
    return rc, s, [tr[i] for i in range(n.value)], g


def run_device(sba, f, opts, cap=256):
    g = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in f.items()}
    d = sba.rig_adjuster.RigBundleAdjuster.desc(g)
    s = sba.capi.Summary()
    tr = (sba.capi.IterationSummary * cap)()
    n = C.c_int32()
    rc = sba.capi.load().sba_rig_solve(C.byref(d), C.byref(opts), C.cast(None, sba.capi.ITERATION_CALLBACK), None, C.byref(s), tr, cap, C.byref(n))
    return rc, s, [tr[i] for i in range(n.value)], g


CASES = [  # (images, snapshots, refine_relative_poses, residuals, effective parameters)  bundle_adjustment_test.cc:755-969
    (2, [[0, 1]], True, 400, 316),
    (4, [[0, 1], [2, 3]], True, 800, 322),
    (4, [[0, 1], [2, 3]], False, 800, 316),
    (4, [[0, 1], [2]], True, 800, 328),
]


@pytest.mark.parametrize("n_img,layout,refine,n_res,n_eff", CASES)
def test_rig_counts_of_the_reference_tests(sba, oracle, n_img, layout, refine, n_res, n_eff):
    ra = sba.rig_adjuster
    rec, rigs, config = rig_scene(sba, n_img, layout)
    ba = ra.RigBundleAdjuster(sba.BundleAdjustmentOptions(), ra.RigBundleAdjusterOptions(refine_relative_poses=refine), config)
    f = ba.Flatten(rec, rigs, errors_fn=oracle_errors(oracle))
    o = ba.SolverOptions()
    o.max_num_iterations = 3
    rc, s, tr, g = run_oracle(sba, oracle, f, o)
    assert rc == 0
    assert (s.num_residuals_reduced, s.num_effective_parameters_reduced) == (n_res, n_eff)
    assert s.final_cost < s.initial_cost
    # the rig pose is the average of the images' poses composed with the inverse relative pose: identity relative
    # poses and identical rotations -> the mean translation (camera_rig.cc:230-258)
    t01 = (np.asarray(generate_reconstruction(sba, n_img, 100).Image(0).Tvec()) + np.asarray(generate_reconstruction(sba, n_img, 100).Image(1).Tvec())) / 2
    assert np.allclose(f["rig_t"][1], t01) and np.allclose(np.abs(f["rig_q"][1]), [1, 0, 0, 0])


def test_rig_oracle_without_rigs_equals_ba_oracle(sba, oracle):
    """No rig: every block hangs on the constant identity pose, and the dense normal equations over all blocks must
    reproduce the Schur-eliminated solve of the ordinary oracle (same cost table to rounding)."""
    ra = sba.rig_adjuster
    rec = generate_reconstruction(sba, 3, 60, seed=2)
    config = sba.BundleAdjustmentConfig()
    for i in range(3):
        config.AddImage(i)
    config.SetConstantPose(0)
    config.SetConstantTvec(1, [0])
    opts = sba.BundleAdjustmentOptions()
    ba_ref = sba.BundleAdjuster(opts, config)
    flat = ba_ref.Flatten(copy.deepcopy(rec))
    o = ba_ref.SolverOptions()
    o.max_num_iterations = 6
    o.linear_solver_type = sba.capi.DENSE_SCHUR
    rc, s_ref, tr_ref, _ = oracle_solve(oracle, sba.capi, flat, o)
    assert rc == 0
    rig_ba = ra.RigBundleAdjuster(opts, ra.RigBundleAdjusterOptions(), config)
    f = rig_ba.Flatten(copy.deepcopy(rec), [], errors_fn=oracle_errors(oracle))
    rc, s, tr, g = run_oracle(sba, oracle, f, o)
    assert rc == 0 and len(tr) == len(tr_ref)
    assert (s.num_residuals_reduced, s.num_effective_parameters_reduced) == (s_ref.num_residuals_reduced, s_ref.num_effective_parameters_reduced)
    for a, b in zip(tr, tr_ref):
        assert a.step_is_successful == b.step_is_successful
        assert abs(a.cost - b.cost) <= 1e-8 * b.cost, (a.iteration, a.cost, b.cost)


@pytest.mark.gpu
@pytest.mark.parametrize("n_img,layout,refine,n_res,n_eff", CASES)
def test_rig_solve_device_matches_oracle(sba, oracle, n_img, layout, refine, n_res, n_eff):
    ra = sba.rig_adjuster
    rec, rigs, config = rig_scene(sba, n_img, layout, seed=1)
    opts = sba.BundleAdjustmentOptions()
    opts.loss_function_type = sba.capi.LOSS_CAUCHY
    ba = ra.RigBundleAdjuster(opts, ra.RigBundleAdjusterOptions(refine_relative_poses=refine), config)
    f = ba.Flatten(rec, rigs)   # gate through the device pass
    o = ba.SolverOptions()
    o.max_num_iterations = 8
    rc_o, s_o, tr_o, g_o = run_oracle(sba, oracle, f, o)
    rc_d, s_d, tr_d, g_d = run_device(sba, f, o)
    assert rc_o == 0 and rc_d == 0
    assert (s_d.num_residuals_reduced, s_d.num_effective_parameters_reduced) == (n_res, n_eff)
    assert s_d.termination_type == s_o.termination_type and len(tr_d) == len(tr_o)
    for a, b in zip(tr_d, tr_o):
        assert a.step_is_successful == b.step_is_successful
        assert abs(a.cost - b.cost) <= 1e-9 * b.cost, (a.iteration, a.cost, b.cost)
    for k in ("rig_q", "rig_t", "rel_q", "rel_t", "cam_params", "points"):
        assert np.allclose(g_d[k], g_o[k], rtol=1e-8, atol=1e-9), k
    assert s_d.final_cost < 0.9 * s_d.initial_cost
    # the reference camera's relative pose never moves; the other one only when asked
    assert np.array_equal(g_d["rel_q"][0], f["rel_q"][0]) and np.array_equal(g_d["rel_t"][0], f["rel_t"][0])
    if not refine:
        assert np.array_equal(g_d["rel_t"], f["rel_t"])
    else:
        assert not np.array_equal(g_d["rel_t"][1], f["rel_t"][1])
    # bit-reproducible
    rc2, s2, tr2, g2 = run_device(sba, f, o)
    assert s2.final_cost == s_d.final_cost and np.array_equal(g2["points"], g_d["points"])


@pytest.mark.gpu
def test_rig_bundle_adjuster_mirror(sba, oracle):
    """RigBundleAdjuster.Solve end to end (TestRigFourViewPartial's shape): counts, poses written back through the
    concatenation (TearDown), the gate, error behaviour."""
    ra = sba.rig_adjuster
    rec, rigs, config = rig_scene(sba, 4, [[0, 1], [2]], seed=3)
    before = copy.deepcopy(rec)
    opts = sba.BundleAdjustmentOptions()
    opts.print_summary = False
    ba = ra.RigBundleAdjuster(opts, ra.RigBundleAdjusterOptions(), config)
    assert ba.Solve(rec, rigs) is True
    s = ba.Summary()
    assert (s.num_residuals_reduced, s.num_effective_parameters_reduced) == (800, 328) and s.final_cost < s.initial_cost
    with pytest.raises(RuntimeError):
        ba.Solve(rec, rigs)
    # images of a snapshot share the rig pose: image pose = relative pose o rig pose (TearDown, :889-903)
    f = ba.flat_
    for image_id, slot in ((0, 1), (1, 1), (2, 2)):
        im = rec.Image(image_id)
        q, t = ra.ConcatenatePoses(f["rig_q"][slot], f["rig_t"][slot], rigs[0].RelativeQvec(im.CameraId()), rigs[0].RelativeTvec(im.CameraId()))
        assert np.allclose(im.Qvec(), q) and np.allclose(im.Tvec(), t)
    assert not np.array_equal(rec.Image(3).Tvec(), before.Image(3).Tvec())      # the image outside the rig moved on its own
    assert not np.array_equal(rec.Point3D(next(iter(rec.points3D))).XYZ(), before.Point3D(next(iter(before.points3D))).XYZ())
    # the gate: with max_reproj_error tiny every rig observation is dropped, only image 3 contributes
    rec2, rigs2, config2 = rig_scene(sba, 4, [[0, 1], [2]], seed=3)
    ba2 = ra.RigBundleAdjuster(opts, ra.RigBundleAdjusterOptions(max_reproj_error=1e-9), config2)
    f2 = ba2.Flatten(rec2, rigs2)
    assert len(f2["b_rig"]) == 100 and (f2["b_rig"] == 0).all()
    # rig images must not be constant (CHECK, :929-932)
    rec3, rigs3, config3 = rig_scene(sba, 2, [[0, 1]])
    config3.SetConstantPose(0)
    with pytest.raises(ValueError):
        ra.RigBundleAdjuster(opts, ra.RigBundleAdjusterOptions(), config3).Flatten(rec3, rigs3)
    # too large for the dense path: refused, not attempted
    big = {k: v for k, v in f.items()}
    d = ra.RigBundleAdjuster.desc(big)
    d.num_blocks = 0
    s0 = sba.capi.Summary(); n0 = C.c_int32()
    o = ba.SolverOptions()
    assert sba.capi.load().sba_rig_solve(C.byref(d), C.byref(o), C.cast(None, sba.capi.ITERATION_CALLBACK), None, C.byref(s0), None, 0, C.byref(n0)) == sba.capi.ERR_EMPTY_PROBLEM


@pytest.mark.parametrize("loss,perturb,refine_focal", [(0, 0.15, True), (2, 0.15, True), (1, 0.3, False), (0, 0.5, True)])
def test_dense_and_schur_loops_of_the_oracle_agree_row_for_row(sba, oracle, loss, perturb, refine_focal):
    """The link of the pinning chain between the LM loop held to Ceres' published table (SolveDense, reached here through
    oracle_rig_solve with no rigs: dense normal equations over ALL blocks) and the Schur-eliminated loop of ba_oracle.cpp
    that the device's bundle-adjustment solves follow: on perturbed scenes -- far enough that steps are REJECTED and the
    radius shrinks -- with the trivial and both robust losses, every row of the two iteration tables agrees: accepted /
    rejected, cost to 1e-8, trust-region radius to 1e-8."""
    ra = sba.rig_adjuster
    rec = generate_reconstruction(sba, 4, 50, seed=5)
    rng = np.random.default_rng(8)
    for im in rec.images.values():
        im.tvec = np.asarray(im.tvec) + rng.normal(0, perturb, 3)
        im.qvec = ra.NormalizeQuaternion(np.asarray(im.qvec) + rng.normal(0, 0.05 * perturb, 4))
    for p in rec.points3D.values():
        p.xyz = np.asarray(p.xyz) + rng.normal(0, perturb, 3)
    config = sba.BundleAdjustmentConfig()
    for i in range(4):
        config.AddImage(i)
    config.SetConstantPose(0)
    config.SetConstantTvec(1, [0])
    opts = sba.BundleAdjustmentOptions()
    opts.loss_function_type = loss
    opts.refine_focal_length = refine_focal
    ba_ref = sba.BundleAdjuster(opts, config)
    flat = ba_ref.Flatten(copy.deepcopy(rec))
    o = ba_ref.SolverOptions()
    o.max_num_iterations = 12
    o.linear_solver_type = sba.capi.DENSE_SCHUR
    rc, s_ref, tr_ref, _ = oracle_solve(oracle, sba.capi, flat, o)
    assert rc == 0
    f = ra.RigBundleAdjuster(opts, ra.RigBundleAdjusterOptions(), config).Flatten(copy.deepcopy(rec), [], errors_fn=oracle_errors(oracle))
    rc, s, tr, g = run_oracle(sba, oracle, f, o)
    assert rc == 0 and len(tr) == len(tr_ref) and s.termination_type == s_ref.termination_type
    for a, b in zip(tr, tr_ref):
        assert a.step_is_successful == b.step_is_successful and a.step_is_valid == b.step_is_valid
        assert abs(a.cost - b.cost) <= 1e-8 * b.cost, (a.iteration, a.cost, b.cost)
        assert abs(a.trust_region_radius - b.trust_region_radius) <= 1e-8 * b.trust_region_radius
    if perturb >= 0.3:
        assert any(not r.step_is_successful for r in tr_ref[1:]), "the scene was meant to produce rejected steps"
