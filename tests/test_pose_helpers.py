"""The pose helpers and the CameraRig mirror the rig adjuster's host side uses (rig_adjuster.py) against the reference's
own tests: /root/reference/src/base/pose_test.cc:138-226, 313-361 and /root/reference/src/base/camera_rig_test.cc:288-330
(statement by statement; BOOST_CHECK_EQUAL -> exact, isApprox -> 1e-6 as written there or Eigen's default 1e-12)."""
import numpy as np
import pytest

I = [1.0, 0.0, 0.0, 0.0]


def test_invert_quaternion(sba):  # pose_test.cc:138-146
    ra = sba.rig_adjuster
    assert np.array_equal(ra.InvertQuaternion(np.array(I)), [1, 0, 0, 0])
    assert np.array_equal(ra.InvertQuaternion(np.array([2.0, 0, 0, 0])), [2, 0, 0, 0])
    q = np.array([0.1, 0.2, 0.3, 0.4])
    assert np.array_equal(ra.InvertQuaternion(ra.InvertQuaternion(q)), q)


def test_concatenate_quaternions(sba):  # pose_test.cc:148-166
    ra = sba.rig_adjuster
    assert np.array_equal(ra.ConcatenateQuaternions(I, I), I)
    assert np.array_equal(ra.ConcatenateQuaternions([2.0, 0, 0, 0], I), I)
    assert np.array_equal(ra.ConcatenateQuaternions(I, [2.0, 0, 0, 0]), I)
    q = np.array([0.1, 0.2, 0.3, 0.4])
    assert np.allclose(ra.ConcatenateQuaternions(q, ra.InvertQuaternion(q)), I, rtol=0, atol=1e-12)
    assert np.allclose(ra.ConcatenateQuaternions(ra.InvertQuaternion(q), q), I, rtol=0, atol=1e-12)


def test_quaternion_rotate_point(sba):  # pose_test.cc:168-186
    ra = sba.rig_adjuster
    assert np.array_equal(ra.QuaternionRotatePoint(I, [0.0, 0, 0]), [0, 0, 0])
    assert np.array_equal(ra.QuaternionRotatePoint([0.1, 0, 0, 0], [0.0, 0, 0]), [0, 0, 0])
    assert np.array_equal(ra.QuaternionRotatePoint(I, [1.0, 1, 0]), [1, 1, 0])
    assert np.array_equal(ra.QuaternionRotatePoint([0.1, 0, 0, 0], [1.0, 1, 0]), [1, 1, 0])
    # EulerAnglesToRotationMatrix(pi, 0, 0) = rotation by pi about x = quaternion (0, 1, 0, 0)
    assert np.allclose(ra.QuaternionRotatePoint([0.0, 1, 0, 0], [1.0, 1, 0]), [1, -1, 0], rtol=0, atol=1e-12)


def test_average_quaternions(sba):  # pose_test.cc:188-226
    ra = sba.rig_adjuster
    avg = ra.AverageQuaternions
    assert np.array_equal(avg([I], [1.0]), I) and np.array_equal(avg([I], [2.0]), I)
    assert np.allclose(avg([I, I], [1.0, 1.0]), I, rtol=0, atol=1e-15)
    assert np.allclose(avg([I, I], [1.0, 2.0]), I, rtol=0, atol=1e-15)
    assert np.allclose(avg([I, [2.0, 0, 0, 0]], [1.0, 2.0]), I, rtol=0, atol=1e-15)
    assert np.allclose(avg([I, [1.0, 1, 0, 0]], [1.0, 1.0]), [0.92388, 0.382683, 0, 0], rtol=1e-6, atol=1e-6)
    assert np.allclose(avg([I, [1.0, 1, 0, 0]], [1.0, 2.0]), [0.850651, 0.525731, 0, 0], rtol=1e-6, atol=1e-6)


def test_concatenate_and_invert_poses(sba):  # pose_test.cc:313-361
    ra = sba.rig_adjuster
    q, t = ra.ConcatenatePoses(I, [0.0, 0, 0], I, [0.0, 0, 0])
    assert np.array_equal(q, I) and np.array_equal(t, [0, 0, 0])
    q, t = ra.ConcatenatePoses(I, [0.0, 0, 0], I, [0.0, 1, 2])
    assert np.array_equal(q, I) and np.array_equal(t, [0, 1, 2])
    q, t = ra.ConcatenatePoses(I, [0.0, 1, 2], I, [3.0, 4, 5])
    assert np.array_equal(q, I) and np.array_equal(t, [3, 5, 7])
    # ComputeRelativePose (pose.cc:176-181) then ConcatenatePoses returns the second pose
    q1, t1, q2, t2 = np.array([1.0, 1, 0, 0]), np.array([0.0, 1, 2]), np.array([1.0, 3, 0, 0]), np.array([3.0, 4, 5])
    rq = ra.ConcatenateQuaternions(ra.InvertQuaternion(q1), q2)
    rt = t2 - ra.QuaternionRotatePoint(rq, t1)
    q, t = ra.ConcatenatePoses(q1, t1, rq, rt)
    assert np.allclose(q, ra.NormalizeQuaternion(q2), rtol=0, atol=1e-12) and np.allclose(t, t2, rtol=0, atol=1e-12)
    iq, it = ra.InvertPose(I, [0.0, 0, 0])
    assert np.array_equal(iq, I) and np.array_equal(it, [0, 0, 0])
    iq, it = ra.InvertPose(np.array([0.0, 1, 2, 3]), np.array([0.0, 1, 2]))
    iiq, iit = ra.InvertPose(iq, it)
    assert np.array_equal(iiq, [0, 1, 2, 3]) and np.allclose(iit, [0, 1, 2], rtol=0, atol=1e-12)


def test_camera_rig_compute_absolute_pose(sba):  # camera_rig_test.cc:288-330
    ra = sba.rig_adjuster
    rig = ra.CameraRig()
    rig.AddCamera(0, I, [0.0, 1, 2])
    rig.AddCamera(1, I, [3.0, 4, 5])
    rig.AddSnapshot([0, 1])
    rec = sba.Reconstruction()
    rec.AddCamera(sba.Camera(0, "PINHOLE", [1.0, 1.0, 0.5, 0.5], 1, 1))
    rec.AddCamera(sba.Camera(1, "PINHOLE", [1.0, 1.0, 0.5, 0.5], 1, 1))
    rec.AddImage(sba.Image(0, 0, I, [0.0, 0, 0], name="a"))
    rec.AddImage(sba.Image(1, 1, I, [3.0, 3, 3], name="b"))
    rig.SetRefCameraId(0)
    rig.Check(rec)
    q, t = rig.ComputeAbsolutePose(0, rec)
    assert np.array_equal(q, I) and np.array_equal(t, [0, -1, -2])
    # Check's failure cases (camera_rig_test.cc:112-157 exercises them through CHECK aborts): a snapshot without the
    # reference camera, an image in two snapshots
    bad = ra.CameraRig()
    bad.AddCamera(0, I, [0.0, 0, 0]); bad.AddCamera(1, I, [0.0, 0, 0])
    bad.AddSnapshot([1])
    bad.SetRefCameraId(0)
    with pytest.raises(ValueError):
        bad.Check(rec)
    twice = ra.CameraRig()
    twice.AddCamera(0, I, [0.0, 0, 0]); twice.AddCamera(1, I, [0.0, 0, 0])
    twice.AddSnapshot([0, 1]); twice.AddSnapshot([0])
    twice.SetRefCameraId(0)
    with pytest.raises(ValueError):
        twice.Check(rec)


def test_camera_param_idxs(sba):  # camera_test.cc:130-138, 254-256: group sizes of FULL_OPENCV and SIMPLE_PINHOLE
    cam = sba.Camera(1, "FULL_OPENCV", [1.0] * 12, 1, 1)
    assert (len(cam.FocalLengthIdxs()), len(cam.PrincipalPointIdxs()), len(cam.ExtraParamsIdxs())) == (2, 2, 8)
    cam = sba.Camera(1, "SIMPLE_PINHOLE", [1.0, 0.5, 0.5], 1, 1)
    assert (len(cam.FocalLengthIdxs()), len(cam.PrincipalPointIdxs()), len(cam.ExtraParamsIdxs())) == (1, 2, 0)
    # every parameter of every model belongs to exactly one group (camera_models.h:189-349)
    for mid, n in sba.capi.MODEL_NUM_PARAMS.items() if isinstance(sba.capi.MODEL_NUM_PARAMS, dict) else enumerate(sba.capi.MODEL_NUM_PARAMS):
        cam = sba.Camera(1, int(mid), [1.0] * int(n), 1, 1)
        groups = list(cam.FocalLengthIdxs()) + list(cam.PrincipalPointIdxs()) + list(cam.ExtraParamsIdxs())
        assert sorted(groups) == list(range(int(n)))


def test_normalize_quaternion(sba):  # pose_test.cc:121-136, on the helper and on Image::NormalizeQvec of the mirror
    ra = sba.rig_adjuster
    half = np.sqrt(2) / 2

    def through_image(q):
        im = sba.Image(1, 1, q, [0.0, 0, 0])
        im.NormalizeQvec()
        return np.asarray(im.Qvec())
    for fn in (ra.NormalizeQuaternion, through_image):
        for q in (I, [2.0, 0, 0, 0], [0.5, 0, 0, 0], [0.0, 0, 0, 0]):
            assert np.array_equal(fn(np.array(q)), I)
        assert np.allclose(fn(np.array([1.0, 1, 0, 0])), [half, half, 0, 0], rtol=0, atol=1e-15)
        assert np.allclose(fn(np.array([0.5, 0.5, 0, 0])), [half, half, 0, 0], rtol=0, atol=1e-15)
