"""CPU tests of the host side: the BundleAdjuster mirror flattens a Reconstruction exactly as
the reference's SetUp does, and both the product's host analysis (sba_problem_counts) and the
oracle reproduce every structural count pinned by src/optim/bundle_adjustment_test.cc:186-645.
Also: the C-ABI library loads and exports every symbol include/sba_b200.h declares, fails loudly
without a device, and shards points deterministically."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def generate_reconstruction(sba, num_images, num_points, seed=0):
    """GenerateReconstruction of bundle_adjustment_test.cc:123-184 (SIMPLE_RADIAL f=1200 on 1000^2,
    identity rotation, t = (U[-1,1], U[-1,1], 10), every point seen by every image, noise U[-2,2])."""
    rng = np.random.default_rng(seed)
    rec = sba.Reconstruction()
    xyz = rng.uniform(-1, 1, (num_points, 3))
    obs = {}
    for i in range(num_images):
        rec.AddCamera(sba.Camera(i, "SIMPLE_RADIAL", [1200.0, 500.0, 500.0, 0.0]))
        im = sba.Image(i, i, (1, 0, 0, 0), (rng.uniform(-1, 1), rng.uniform(-1, 1), 10.0))
        p = xyz + im.tvec
        uv = p[:, :2] / p[:, 2:3]
        xy = 1200.0 * uv + 500.0 + rng.uniform(-2, 2, uv.shape)
        im.points2D = [sba.Point2D(v) for v in xy]
        rec.AddImage(im)
    for j in range(num_points):
        rec.AddPoint3D(xyz[j], [(i, j) for i in range(num_images)])
    return rec


def delete_observation(rec, image_id, point2D_idx):
    p2 = rec.Image(image_id).Point2D(point2D_idx)
    pid = p2.Point3DId()
    tr = rec.Point3D(pid).Track()
    tr.elements = [e for e in tr.elements if not (e.image_id == image_id and e.point2D_idx == point2D_idx)]
    from importlib import import_module
    p2.point3D_id = import_module("semantic-bundle-adjustment-colmap_b200").bundle_adjustment.kInvalidPoint3DId
    return pid


def counts(sba, oracle, rec, config, options=None):
    ba = sba.BundleAdjuster(options or sba.BundleAdjustmentOptions(), config)
    flat = ba.Flatten(rec)
    s = flat.counts()
    so = sba.capi.Summary()
    d = flat.desc()
    oracle.oracle_ba_counts(C.byref(d), C.byref(so))
    assert (s.num_residuals_reduced, s.num_effective_parameters_reduced, s.num_parameters_reduced) == \
           (so.num_residuals_reduced, so.num_effective_parameters_reduced, so.num_parameters_reduced)
    return s.num_residuals_reduced, s.num_effective_parameters_reduced, flat


def two_view_config(sba, n=2):
    c = sba.BundleAdjustmentConfig()
    for i in range(n):
        c.AddImage(i)
    c.SetConstantPose(0)
    c.SetConstantTvec(1, [0])
    return c


def test_config_num_observations(sba):  # :186-208
    rec = generate_reconstruction(sba, 4, 100)
    c = sba.BundleAdjustmentConfig()
    c.AddImage(0); c.AddImage(1)
    assert c.NumResiduals(rec) == 400
    c.AddVariablePoint(1)
    assert c.NumResiduals(rec) == 404
    c.AddConstantPoint(2)
    assert c.NumResiduals(rec) == 408
    c.AddImage(2)
    assert c.NumResiduals(rec) == 604
    c.AddImage(3)
    assert c.NumResiduals(rec) == 800


def test_two_view(sba, oracle):  # :210-246 -> 400 / 309
    rec = generate_reconstruction(sba, 2, 100)
    assert counts(sba, oracle, rec, two_view_config(sba))[:2] == (400, 309)


def test_two_view_constant_camera(sba, oracle):  # :248-284 -> 400 / 302
    rec = generate_reconstruction(sba, 2, 100)
    c = sba.BundleAdjustmentConfig()
    c.AddImage(0); c.AddImage(1); c.SetConstantPose(0); c.SetConstantPose(1); c.SetConstantCamera(0)
    assert counts(sba, oracle, rec, c)[:2] == (400, 302)


def test_partially_contained_tracks(sba, oracle):  # :286-330 -> 400 / 7
    rec = generate_reconstruction(sba, 3, 100)
    delete_observation(rec, 2, 0)
    c = sba.BundleAdjustmentConfig()
    c.AddImage(0); c.AddImage(1); c.SetConstantPose(0); c.SetConstantPose(1)
    nres, npar, flat = counts(sba, oracle, rec, c)
    assert (nres, npar) == (400, 7)
    assert flat.point_constant.sum() == 99


def test_partially_contained_tracks_force_to_optimize_point(sba, oracle):  # :332-388 -> 402 / 10
    rec = generate_reconstruction(sba, 3, 100)
    add_variable = rec.Image(2).Point2D(1).Point3DId()
    add_constant = rec.Image(2).Point2D(2).Point3DId()
    delete_observation(rec, 2, 0)
    c = sba.BundleAdjustmentConfig()
    c.AddImage(0); c.AddImage(1); c.SetConstantPose(0); c.SetConstantPose(1)
    c.AddVariablePoint(add_variable); c.AddConstantPoint(add_constant)
    nres, npar, flat = counts(sba, oracle, rec, c)
    assert (nres, npar) == (402, 10)
    # camera 2 entered through AddPointToProblem and was forced constant (bundle_adjustment.cc:456-459)
    assert flat.camera_param_constant_mask[2] == 0b1111 and flat.image_pose_constant[2] == 1


def test_constant_points(sba, oracle):  # :390-435 -> 400 / 298
    rec = generate_reconstruction(sba, 2, 100)
    c = sba.BundleAdjustmentConfig()
    c.AddImage(0); c.AddImage(1); c.SetConstantPose(0); c.SetConstantPose(1)
    c.AddConstantPoint(1); c.AddConstantPoint(2)
    assert counts(sba, oracle, rec, c)[:2] == (400, 298)


def test_variable_image(sba, oracle):  # :437-477 -> 600 / 317
    rec = generate_reconstruction(sba, 3, 100)
    assert counts(sba, oracle, rec, two_view_config(sba, 3))[:2] == (600, 317)


@pytest.mark.parametrize("flag,value,expect", [("refine_focal_length", False, 307),   # :479-529
                                               ("refine_principal_point", True, 313),  # :531-593
                                               ("refine_extra_params", False, 307)])   # :595-645
def test_intrinsics_groups(sba, oracle, flag, value, expect):
    rec = generate_reconstruction(sba, 2, 100)
    o = sba.BundleAdjustmentOptions()
    setattr(o, flag, value)
    assert counts(sba, oracle, rec, two_view_config(sba), o)[:2] == (400, expect)


def test_residual_order_is_the_reference_loop_nest(sba):
    """config.Images() outer, Points2D index inner, then out-of-config track elements (a7)."""
    rec = generate_reconstruction(sba, 3, 5)
    c = sba.BundleAdjustmentConfig()
    c.AddImage(1); c.AddImage(0)  # insertion order is the iteration order of the mirror's container
    c.AddVariablePoint(rec.Image(2).Point2D(3).Point3DId())
    flat = sba.BundleAdjuster(sba.BundleAdjustmentOptions(), c).Flatten(rec)
    assert flat.obs_image.tolist() == [0] * 5 + [1] * 5 + [2]
    assert flat.obs_point.tolist() == [0, 1, 2, 3, 4] * 2 + [3]
    assert np.array_equal(flat.obs_xy[5:10], np.array([p.XY() for p in rec.Image(0).Points2D()]))
    assert flat.image_pose_constant.tolist() == [0, 0, 1]


def test_adjuster_is_one_shot_and_checks_options(sba):
    o = sba.BundleAdjustmentOptions()
    o.loss_function_scale = -1.0
    with pytest.raises(ValueError):
        sba.BundleAdjuster(o, sba.BundleAdjustmentConfig())
    ba = sba.BundleAdjuster(sba.BundleAdjustmentOptions(), sba.BundleAdjustmentConfig())
    assert ba.Solve(sba.Reconstruction()) is False  # empty problem, bundle_adjustment.cc:267-269
    with pytest.raises(RuntimeError):
        ba.Solve(sba.Reconstruction())               # CHECK(!problem_), :260
    with pytest.raises(ValueError):
        sba.Camera(0, 99, [1.0])                     # unknown model, camera_models.h:140
