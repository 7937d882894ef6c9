"""The reference's own BundleAdjuster integration tests (src/optim/bundle_adjustment_test.cc:210-477),
run through the host mirror on the GPU: counts plus which parameters did / did not change."""
import copy

import numpy as np
import pytest

from test_host_logic import delete_observation, generate_reconstruction, two_view_config

pytestmark = pytest.mark.gpu


def changed(a, b):
    return not np.array_equal(a, b)


def test_two_view(sba):  # :210-246
    rec = generate_reconstruction(sba, 2, 100)
    orig = copy.deepcopy(rec)
    ba = sba.BundleAdjuster(sba.BundleAdjustmentOptions(), two_view_config(sba))
    assert ba.Solve(rec)
    s = ba.Summary()
    assert (s.num_residuals_reduced, s.num_effective_parameters_reduced) == (400, 309)
    assert s.final_cost < s.initial_cost
    for c in (0, 1):  # CheckVariableCamera: focal + extra change, principal point fixed
        a, b = rec.Camera(c).Params(), orig.Camera(c).Params()
        assert a[0] != b[0] and a[3] != b[3] and a[1] == b[1] and a[2] == b[2]
    assert not changed(rec.Image(0).Qvec(), orig.Image(0).Qvec()) and not changed(rec.Image(0).Tvec(), orig.Image(0).Tvec())
    assert changed(rec.Image(1).Qvec(), orig.Image(1).Qvec())           # CheckConstantXImage
    assert rec.Image(1).Tvec()[0] == orig.Image(1).Tvec()[0] and rec.Image(1).Tvec()[1] != orig.Image(1).Tvec()[1]
    for pid, p in rec.points3D.items():
        assert changed(p.XYZ(), orig.Point3D(pid).XYZ())


def test_partially_contained_tracks(sba):  # :286-330
    rec = generate_reconstruction(sba, 3, 100)
    variable_pid = delete_observation(rec, 2, 0)
    orig = copy.deepcopy(rec)
    c = sba.BundleAdjustmentConfig()
    c.AddImage(0); c.AddImage(1); c.SetConstantPose(0); c.SetConstantPose(1)
    ba = sba.BundleAdjuster(sba.BundleAdjustmentOptions(), c)
    assert ba.Solve(rec)
    s = ba.Summary()
    assert (s.num_residuals_reduced, s.num_effective_parameters_reduced) == (400, 7)
    assert not changed(rec.Camera(2).Params(), orig.Camera(2).Params())
    for i in range(3):
        assert not changed(rec.Image(i).Qvec(), orig.Image(i).Qvec())
    for pid, p in rec.points3D.items():
        assert changed(p.XYZ(), orig.Point3D(pid).XYZ()) == (pid == variable_pid)


def test_callbacks_and_state_updates(sba):
    """ceres::IterationCallback contract used by the controller (src/controllers/bundle_adjustment.cc:43-61)."""
    rec = generate_reconstruction(sba, 3, 100)
    o = sba.BundleAdjustmentOptions()
    o.solver_options.update_state_every_iteration = True
    seen = []

    def cb(it):
        seen.append((it.iteration, it.cost, rec.Image(2).Tvec().copy()))
        return 1 if it.iteration == 3 else 0  # SOLVER_TERMINATE_SUCCESSFULLY
    o.solver_options.callbacks.append(cb)
    ba = sba.BundleAdjuster(o, two_view_config(sba, 3))
    assert ba.Solve(rec)
    assert [s[0] for s in seen] == [0, 1, 2, 3]
    assert ba.Summary().termination_type == sba.capi.USER_SUCCESS
    assert changed(seen[0][2], seen[2][2])  # the callback saw updated parameters


def test_soft_l1_local_ba_preset(sba):
    rec = generate_reconstruction(sba, 4, 100)
    o = sba.BundleAdjustmentOptions()
    o.loss_function_type = o.SOFT_L1      # src/controllers/incremental_mapper.cc:237-238
    o.solver_options.max_num_iterations = 25
    ba = sba.BundleAdjuster(o, two_view_config(sba, 4))
    assert ba.Solve(rec)
    s = ba.Summary()
    assert s.num_residuals_reduced == 800 and s.final_cost < s.initial_cost
