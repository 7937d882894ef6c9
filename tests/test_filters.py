"""Mapper-side filters around global BA: per-observation squared reprojection error and depth
on the device (sba_observation_errors) + the reference's bookkeeping loops (filters.py).
Reference: src/base/projection.cc:111-128, 191-195; src/base/reconstruction.cc:236-277, 647-665, 1472-1523."""
import ctypes as C

import numpy as np
import pytest

dp = C.POINTER(C.c_double)


def scene_reconstruction(sba, prob):
    rec = sba.Reconstruction()
    for c in range(len(prob.camera_model)):
        mid = int(prob.camera_model[c])
        rec.AddCamera(sba.Camera(c + 1, mid, prob.camera_params[c, :sba.capi.MODEL_NUM_PARAMS[mid]], width=1000, height=1000))
    for i in range(len(prob.image_camera)):
        rec.AddImage(sba.Image(i + 1, int(prob.image_camera[i]) + 1, prob.image_qvec[i], prob.image_tvec[i], name="im%d.png" % i))
    tracks = {}
    for k in range(prob.num_observations):
        img = rec.Image(int(prob.obs_image[k]) + 1)
        img.points2D.append(sba.Point2D(prob.obs_xy[k].copy()))
        tracks.setdefault(int(prob.obs_point[k]), []).append((img.image_id, len(img.points2D) - 1))
    for q in range(len(prob.points)):
        rec.AddPoint3D(prob.points[q], tracks.get(q, []))
    return rec


def oracle_errors(oracle):
    def fn(prob):
        n = prob.num_observations
        err, depth = np.zeros(n), np.zeros(n)
        e, z = C.c_double(), C.c_double()
        for k in range(n):
            i, q = int(prob.obs_image[k]), int(prob.obs_point[k])
            c = int(prob.image_camera[i])
            a = [np.ascontiguousarray(v) for v in (prob.image_qvec[i], prob.image_tvec[i], prob.points[q], prob.camera_params[c], prob.obs_xy[k])]
            oracle.oracle_observation_error(int(prob.camera_model[c]), *[v.ctypes.data_as(dp) for v in a], C.byref(e), C.byref(z))
            err[k], depth[k] = e.value, z.value
        return err, depth
    return fn


def planted(sba, seed=2):
    prob, _ = sba.scene.make_scene(n_cam=8, n_pt=120, mean_track=4.0, seed=seed, perturb=False, noise_px=0.5)
    rec = scene_reconstruction(sba, prob)
    return rec


def snapshot(rec):
    return ({pid: [(e.image_id, e.point2D_idx) for e in p.track.Elements()] for pid, p in rec.points3D.items()},
            {iid: [p.point3D_id for p in im.points2D] for iid, im in rec.images.items()},
            {pid: p.error for pid, p in rec.points3D.items()})


def test_known_answers_of_the_error_function(oracle):
    """A point on the optical axis projects to the principal point; behind the camera the error is DBL_MAX."""
    e, z = C.c_double(), C.c_double()
    q, t = np.array([2.0, 0, 0, 0]), np.array([0.0, 0, 0])          # non-unit quaternion: normalised first
    cam, xy = np.array([100.0, 320, 240] + [0.0] * 9), np.array([323.0, 244.0])
    X = np.array([0.0, 0, 5])
    oracle.oracle_observation_error(0, *[v.ctypes.data_as(dp) for v in (q, t, X, cam, xy)], C.byref(e), C.byref(z))
    assert e.value == 25.0 and z.value == 5.0
    X = np.array([0.0, 0, -5])
    oracle.oracle_observation_error(0, *[v.ctypes.data_as(dp) for v in (q, t, X, cam, xy)], C.byref(e), C.byref(z))
    assert e.value == np.finfo(np.float64).max and z.value == -5.0
    X = np.array([0.1, -0.2, 2.0])                                  # f * (0.05, -0.1) + pp = (325, 230)
    oracle.oracle_observation_error(0, *[v.ctypes.data_as(dp) for v in (q, t, X, cam, xy)], C.byref(e), C.byref(z))
    assert e.value == pytest.approx(2.0 ** 2 + 14.0 ** 2, rel=1e-14)


def test_negative_depth_filter_follows_the_reference_loop(sba, oracle):
    rec = planted(sba)
    far = [pid for pid, p in rec.points3D.items() if p.track.Length() >= 4][:3]
    short = [pid for pid, p in rec.points3D.items() if p.track.Length() == 2][:1]
    for pid in far + short:
        rec.points3D[pid].xyz[:] = rec.points3D[pid].xyz * -40.0     # far on the other side: behind most cameras
    before = snapshot(rec)
    fn = oracle_errors(oracle)
    prob, keys = sba.filters.flatten_observations(rec)
    depth = dict(zip(keys, fn(prob)[1]))
    n = sba.filters.filter_observations_with_negative_depth(rec, errors_fn=fn)
    assert n > 0
    tracks, p2d, _ = snapshot(rec)
    for pid, tr in before[0].items():
        bad = [k for k in tr if not depth[k] >= np.finfo(float).eps]
        if not bad:
            assert tracks[pid] == tr
        elif len(tr) - len(bad) < 2:
            assert pid not in tracks                                  # DeleteObservation on a 2-track deletes the point
            assert all(p2d[i][j] == sba.bundle_adjustment.kInvalidPoint3DId for i, j in tr)
        else:
            assert tracks[pid] == [k for k in tr if k not in bad]
    assert sba.filters.filter_observations_with_negative_depth(rec, errors_fn=fn) == 0   # idempotent


def test_reprojection_error_filter_follows_the_reference_loop(sba, oracle):
    rec = planted(sba, seed=4)
    pids = [pid for pid, p in rec.points3D.items() if p.track.Length() >= 4]
    one_bad, all_but_one_bad = pids[0], pids[1]
    el = rec.points3D[one_bad].track.Elements()[1]
    rec.images[el.image_id].points2D[el.point2D_idx].xy += 60.0
    for el in rec.points3D[all_but_one_bad].track.Elements()[1:]:
        rec.images[el.image_id].points2D[el.point2D_idx].xy -= 45.0
    lonely = [pid for pid, p in rec.points3D.items() if p.track.Length() == 2][0]
    gone = rec.points3D[lonely].track.elements.pop()
    rec.images[gone.image_id].points2D[gone.point2D_idx].point3D_id = sba.bundle_adjustment.kInvalidPoint3DId
    before = snapshot(rec)
    fn = oracle_errors(oracle)
    n = sba.filters.filter_points3D_with_large_reprojection_error(rec, 4.0, errors_fn=fn)
    tracks, p2d, errs = snapshot(rec)
    assert lonely not in tracks and all_but_one_bad not in tracks                  # track < 2, and >= length - 1 bad elements
    assert tracks[one_bad] == [k for j, k in enumerate(before[0][one_bad]) if j != 1]
    assert n == 1 + len(before[0][all_but_one_bad]) + 1
    assert 0 < errs[one_bad] < 4.0                                                   # mean error over the remaining elements
    untouched = [pid for pid in tracks if pid not in (one_bad,)]
    assert all(tracks[pid] == before[0][pid] for pid in untouched)
    assert all(0 <= errs[pid] < 4.0 for pid in untouched)


def test_observation_errors_need_a_device(sba):
    prob, _ = sba.scene.make_scene(n_cam=3, n_pt=20, mean_track=3.0, seed=1)
    lib = sba.capi.load()
    d = prob.desc()
    err = np.zeros(prob.num_observations)
    rc = lib.sba_observation_errors(C.byref(d), err.ctypes.data_as(dp), None)
    assert rc == (0 if lib.sba_device_count() >= 1 else 2)
    assert lib.sba_observation_errors(None, err.ctypes.data_as(dp), None) == 1


@pytest.mark.gpu
def test_device_errors_match_the_oracle_for_every_model(sba, oracle):
    from test_rig import MODEL_PARAMS
    rng = np.random.default_rng(17)
    n_img, n_pt, n = 11, 60, 900
    q = rng.normal(size=(n_img, 4)) * [1, 0.1, 0.1, 0.1] * 3.0           # not normalised on purpose
    t = rng.normal(0, 0.2, (n_img, 3)) + [0, 0, 4.0]
    cam = np.zeros((11, 12))
    for m, p in MODEL_PARAMS.items():
        cam[m, :len(p)] = p
    pts = rng.uniform(-1, 1, (n_pt, 3))
    pts[:5, 2] -= 30.0                                                      # some behind every camera
    prob = sba.FlatProblem(q, t, np.arange(n_img), np.arange(11), cam, pts, rng.integers(0, n_img, n), rng.integers(0, n_pt, n),
                           rng.uniform(100, 700, (n, 2)))
    err, depth = sba.filters.device_observation_errors(prob)
    ref_err, ref_depth = oracle_errors(oracle)(prob)
    assert np.array_equal(err == np.finfo(float).max, ref_err == np.finfo(float).max) and (err == np.finfo(float).max).any()
    ok = ref_err != np.finfo(float).max
    assert np.allclose(depth, ref_depth, rtol=1e-13, atol=1e-13)
    assert np.allclose(err[ok], ref_err[ok], rtol=1e-10, atol=0)


@pytest.mark.gpu
def test_filters_on_the_device_equal_the_oracle_run(sba, oracle):
    a, b = planted(sba, seed=6), planted(sba, seed=6)
    for rec in (a, b):
        pids = [pid for pid, p in rec.points3D.items() if p.track.Length() >= 3]
        for pid in pids[:4]:
            el = rec.points3D[pid].track.Elements()[0]
            rec.images[el.image_id].points2D[el.point2D_idx].xy += 25.0
        rec.points3D[pids[5]].xyz[:] *= -30.0
    na = (sba.filters.filter_observations_with_negative_depth(a), sba.filters.filter_points3D_with_large_reprojection_error(a, 4.0))
    fn = oracle_errors(oracle)
    nb = (sba.filters.filter_observations_with_negative_depth(b, errors_fn=fn),
          sba.filters.filter_points3D_with_large_reprojection_error(b, 4.0, errors_fn=fn))
    assert na == nb and na[0] > 0 and na[1] > 0
    sa, sb = snapshot(a), snapshot(b)
    assert sa[0] == sb[0] and sa[1] == sb[1]
    assert all(abs(sa[2][k] - sb[2][k]) <= 1e-10 * max(1.0, abs(sb[2][k])) for k in sb[2])


# ------------------------------------------------------------------ Reconstruction::Normalize
def _normalize_inputs(sba, seed=0, n_img=40, n_pt=3000):
    rng = np.random.default_rng(seed)
    q = rng.normal(size=(n_img, 4)); q[::7] *= 3.0          # not normalised on purpose; some w < 0
    t = rng.normal(scale=4.0, size=(n_img, 3)) + np.array([30.0, -12.0, 5.0])
    X = rng.normal(scale=6.0, size=(n_pt, 3)) + np.array([-20.0, 8.0, 40.0])
    reg = np.ones(n_img, np.uint8); reg[3] = 0; reg[min(11, n_img - 1)] = 0
    return q, t, reg, X


def test_normalize_oracle_properties(sba):
    """The restatement itself: extent of the robust box of the transformed centres = `extent`, centroid at the origin,
    projections unchanged (a similarity of the world does not move image points)."""
    from oracle import normalize_oracle as no
    q, t, reg, X = _normalize_inputs(sba)
    q0, t0, X0 = q.copy(), t.copy(), X.copy()
    s, tr = no.normalize(q, t, reg, X, 10.0, 0.1, 0.9, True)
    centres = np.array([-(no.quat_to_rot(q[i]).T @ t[i]) for i in np.nonzero(reg)[0]])
    lo, hi, mean = no.bounds_and_centroid(centres, 0.1, 0.9)
    assert abs(np.linalg.norm(hi - lo) - 10.0) < 1e-5 and np.linalg.norm(mean) < 1e-5   # float-cast coordinates
    for i in (0, 3, 17):
        a = no.quat_to_rot(q0[i]) @ X0[5] + t0[i]
        b = no.quat_to_rot(q[i]) @ X[5] + t[i]
        assert np.allclose(a[:2] / a[2], b[:2] / b[2], rtol=1e-12) and np.isclose(b[2], s * a[2], rtol=1e-12)
        assert np.isclose(np.linalg.norm(q[i]), 1.0, rtol=1e-14)
    # fewer than two elements: untouched
    q1, t1, X1 = q0[:1].copy(), t0[:1].copy(), X0.copy()
    assert no.normalize(q1, t1, np.ones(1, np.uint8), X1, 10.0, 0.1, 0.9, True)[0] == 1.0 and np.array_equal(X1, X0)


@pytest.mark.gpu
@pytest.mark.parametrize("use_images", [True, False])
def test_normalize_device_matches_oracle(sba, use_images):
    from oracle import normalize_oracle as no
    q, t, reg, X = _normalize_inputs(sba, seed=4)
    qo, to, Xo = q.copy(), t.copy(), X.copy()
    s_ref, tr_ref = no.normalize(qo, to, reg, Xo, 10.0, 0.1, 0.9, use_images)
    s, tr = sba.filters.device_normalize(q, t, reg, X, 10.0, 0.1, 0.9, use_images)
    assert abs(s - s_ref) <= 1e-13 * s_ref and np.allclose(tr, tr_ref, rtol=1e-12, atol=1e-12)
    assert np.allclose(X, Xo, rtol=1e-12, atol=1e-12) and np.allclose(t, to, rtol=1e-12, atol=1e-12)
    assert np.allclose(q, qo, rtol=0, atol=1e-14)
    # through the Reconstruction mirror, and the degenerate cases
    rec = sba.Reconstruction()
    rec.AddCamera(sba.Camera(1, "SIMPLE_PINHOLE", [500.0, 320.0, 240.0], 640, 480))
    q2, t2, _, X2 = _normalize_inputs(sba, seed=9, n_img=5, n_pt=50)
    for i in range(5):
        rec.AddImage(sba.Image(i + 1, 1, q2[i], t2[i], name="i%d.png" % i))
    for k in range(50):
        rec.AddPoint3D(X2[k], [])
    s2, _ = sba.filters.normalize(rec)
    qo2, to2, Xo2 = q2.copy(), t2.copy(), X2.copy()
    s2_ref, _ = no.normalize(qo2, to2, np.ones(5, np.uint8), Xo2, 10.0, 0.1, 0.9, True)
    assert abs(s2 - s2_ref) <= 1e-13 * s2_ref and np.allclose(rec.Image(3).Tvec(), to2[2], rtol=1e-12, atol=1e-12)
    lib = sba.capi.load()
    assert lib.sba_normalize(0, None, None, None, 0, None, 0.0, 0.1, 0.9, 1, None, None) == sba.capi.ERR_INVALID_ARGUMENT
    one_q, one_t = q[:1].copy(), t[:1].copy()
    assert sba.filters.device_normalize(one_q, one_t, np.ones(1, np.uint8), X.copy(), 10.0, 0.1, 0.9, True)[0] == 1.0
    assert np.array_equal(one_q, q[:1])


def _replay_reference_normalize_test(normalize):
    """TestNormalize of the reference, /root/reference/src/base/reconstruction_test.cc:230-286, statement by statement:
    three images with identity rotation at tvec z = -10, 0, 10 (no 3-D points), normalised while unregistered, registered,
    with another extent, with the full percentile range, and after four more images joined."""
    q = np.tile([1.0, 0.0, 0.0, 0.0], (3, 1))
    t = np.zeros((3, 3))
    t[:, 2] = [-10.0, 0.0, 10.0]
    pts = np.zeros((0, 3))
    reg = np.zeros(3, np.uint8)
    normalize(q, t, reg, pts, 10.0, 0.1, 0.9, True)                      # :237-243 nothing registered: untouched
    assert np.allclose(t[:, 2], [-10, 0, 10], atol=1e-6)
    reg[:] = 1
    normalize(q, t, reg, pts, 10.0, 0.1, 0.9, True)                      # :244-250
    assert np.allclose(t[:, 2], [-5, 0, 5], atol=1e-6)
    normalize(q, t, reg, pts, 5.0, 0.1, 0.9, True)                       # :251-254
    assert np.allclose(t[:, 2], [-2.5, 0, 2.5], atol=1e-6)
    normalize(q, t, reg, pts, 10.0, 0.0, 1.0, True)                      # :255-258
    assert np.allclose(t[:, 2], [-5, 0, 5], atol=1e-6)
    normalize(q, t, reg, pts, 20.0, 0.1, 0.9, True)                      # :259
    q = np.vstack([q, np.tile([1.0, 0.0, 0.0, 0.0], (4, 1))])           # :260-277 images 4..7
    t = np.vstack([t, np.zeros((4, 3))])
    t[3:, 2] = [-7.5, -5.0, 5.0, 7.5]
    reg = np.ones(7, np.uint8)
    normalize(q, t, reg, pts, 10.0, 0.0, 1.0, True)                      # :278-285
    assert np.allclose(t[:, 2], [-5, 0, 5, -3.75, -2.5, 2.5, 3.75], atol=1e-6)
    assert np.allclose(t[:, :2], 0.0, atol=1e-12) and np.allclose(q, [1, 0, 0, 0], atol=1e-12)


def test_normalize_oracle_passes_the_reference_test():
    from oracle import normalize_oracle as no
    _replay_reference_normalize_test(no.normalize)


@pytest.mark.gpu
def test_normalize_device_passes_the_reference_test(sba):
    _replay_reference_normalize_test(sba.filters.device_normalize)


def test_error_and_depth_pass_the_reference_projection_tests(oracle):
    """/root/reference/src/base/projection_test.cc:95-124 (TestCalculateSquaredReprojectionError, the (qvec, tvec) overloads),
    :180-198 (TestCalculateDepth) and :200-222 (TestHasPointPositiveDepth) on the oracle's per-observation pass: identity
    pose, SIMPLE_PINHOLE f = 1, c = (0, 0); a point's own projection has error exactly 0, the projection shifted by (1, 1)
    has squared error 2 (1e-6 %), depths 0 / 0 / 1 / -1 and the positive-depth predicate !, !, ok, !."""
    e, z = C.c_double(), C.c_double()
    q, t = np.array([1.0, 0, 0, 0]), np.array([0.0, 0, 0])
    cam = np.array([1.0, 0.0, 0.0] + [0.0] * 9)

    def run(X, xy):
        X, xy = np.ascontiguousarray(X, dtype=np.float64), np.ascontiguousarray(xy, dtype=np.float64)
        oracle.oracle_observation_error(0, *[v.ctypes.data_as(dp) for v in (q, t, X, cam, xy)], C.byref(e), C.byref(z))
        return e.value, z.value
    rng = np.random.default_rng(3)
    for _ in range(20):                                  # Eigen::Vector3d::Random().cwiseAbs()
        X = np.abs(rng.uniform(-1, 1, 3)) + 1e-3
        xy = X[:2] / X[2]
        assert run(X, xy)[0] == 0.0                      # BOOST_CHECK_EQUAL(error1, 0)
        assert abs(run(X, xy + 1)[0] - 2.0) <= 2.0 * 1e-8  # BOOST_CHECK_CLOSE(error3, 2, 1e-6): tolerance in percent
    depths = [run(X, [0.0, 0.0])[1] for X in ([0.0, 0, 0], [0.0, 2, 0], [0.0, 0, 1], [0.0, 0, -1])]
    assert depths == [0.0, 0.0, 1.0, -1.0]
    assert [d >= np.finfo(np.float64).eps for d in depths] == [False, False, True, False]  # projection.cc:191-195
    # a point that is not in front of the camera has the largest error (projection.cc:117-119)
    assert run([0.0, 0, -1], [0.0, 0.0])[0] == np.finfo(np.float64).max and run([0.0, 2, 0], [0.0, 0.0])[0] == np.finfo(np.float64).max


def _reference_test_reconstruction(sba, num_images):
    """GenerateReconstruction of reconstruction_test.cc:42-66: one PINHOLE camera (f = 1 on a 1 x 1 image: params
    1, 1, 0.5, 0.5), identity poses, ten features at (0, 0) per image."""
    rec = sba.Reconstruction()
    rec.AddCamera(sba.Camera(1, "PINHOLE", [1.0, 1.0, 0.5, 0.5], 1, 1))
    for i in range(1, num_images + 1):
        im = sba.Image(i, 1, [1.0, 0, 0, 0], [0.0, 0, 0], name="image%d" % i)
        im.points2D = [sba.Point2D(np.zeros(2)) for _ in range(10)]
        rec.AddImage(im)
    return rec


def _add_observation(rec, pid, image_id, idx):  # Reconstruction::AddObservation, reconstruction.cc:326-339
    rec.Point3D(pid).track.AddElement(image_id, idx)
    rec.Image(image_id).points2D[idx].point3D_id = pid


def test_negative_depth_filter_passes_the_reference_test(sba, oracle):
    """TestFilterObservationsWithNegativeDepth, reconstruction_test.cc:510-532, statement by statement."""
    fl = sba.filters
    fn = oracle_errors(oracle)
    rec = _reference_test_reconstruction(sba, 2)
    pid = rec.AddPoint3D([0.0, 0.0, 1.0], [])
    for z in (1.0, 0.001, 0.0):                      # no observation yet: nothing can be filtered
        rec.Point3D(pid).xyz[2] = z
        fl.filter_observations_with_negative_depth(rec, errors_fn=fn)
        assert len(rec.points3D) == 1
    _add_observation(rec, pid, 1, 0)
    rec.Point3D(pid).xyz[2] = 0.001
    fl.filter_observations_with_negative_depth(rec, errors_fn=fn)
    assert len(rec.points3D) == 1
    rec.Point3D(pid).xyz[2] = 0.0                    # in the image plane: not in front of the camera
    fl.filter_observations_with_negative_depth(rec, errors_fn=fn)
    assert len(rec.points3D) == 0 and not rec.Image(1).points2D[0].HasPoint3D()


@pytest.mark.parametrize("all_points", [False, True])
def test_reprojection_filter_passes_the_reference_tests(sba, oracle, all_points):
    """The reprojection-error half of TestFilterPoints3D / TestFilterAllPoints (reconstruction_test.cc:394-435, 479-508;
    the statements with min_tri_angle = 0, where FilterPoints3D is FilterPoints3DWithLargeReprojectionError followed by
    an angle filter that deletes nothing a two-view point has): points without a pair of observations go, a point whose
    projection error is 0.1 px stays at max_reproj_error 0.1 and goes at 0.09."""
    fl = sba.filters
    fn = oracle_errors(oracle)
    rng = np.random.default_rng(1)

    def run(max_err, ids):
        fl.filter_points3D_with_large_reprojection_error(rec, max_err, point3D_ids=None if all_points else ids, errors_fn=fn)
        return len(rec.points3D)
    rec = _reference_test_reconstruction(sba, 2)
    p1 = rec.AddPoint3D(rng.uniform(-1, 1, 3), [])
    if not all_points:
        assert run(0.0, []) == 1 and run(0.0, [p1 + 1]) == 1   # no id / an id that does not exist
    assert run(0.0, [p1]) == 0                                 # no observation
    p2 = rec.AddPoint3D(rng.uniform(-1, 1, 3), [])
    _add_observation(rec, p2, 1, 0)
    assert run(0.0, [p2]) == 0                                 # one observation
    p3 = rec.AddPoint3D([-0.5, -0.5, 1.0], [])
    _add_observation(rec, p3, 1, 0)
    _add_observation(rec, p3, 2, 0)
    assert run(0.0, [p3]) == 1                                 # projects onto its features: error 0 is not > 0
    fl.delete_point3D(rec, p3)                                 # (the reference's next statement removes it by triangulation angle)
    p4 = rec.AddPoint3D([-0.6, -0.5, 1.0], [])
    _add_observation(rec, p4, 1, 0)
    _add_observation(rec, p4, 2, 0)
    assert run(0.1, [p4]) == 1
    assert run(0.09, [p4]) == 0


def test_bounds_and_centroid_pass_the_reference_test():
    """TestComputeBoundsAndCentroid, reconstruction_test.cc:288-319: points (3,0,0), (0,3,0), (0,0,3) over the full
    percentile range have centroid (1,1,1) and bounding box (0,0,0) - (3,3,3)."""
    from oracle import normalize_oracle as no
    lo, hi, mean = no.bounds_and_centroid(np.array([[3.0, 0, 0], [0, 3.0, 0], [0, 0, 3.0]]), 0.0, 1.0)
    assert np.allclose(mean, 1.0, atol=1e-6) and np.allclose(lo, 0.0, atol=1e-6) and np.allclose(hi, 3.0, atol=1e-6)
