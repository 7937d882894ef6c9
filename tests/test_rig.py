"""Camera-rig reprojection functor: RigBundleAdjustmentCostFunction,
/root/reference/src/base/cost_functions.h:160-214.

CPU : the oracle's Jet Jacobian (2 x 29: rig_q, rig_t, rel_q, rel_t, X, params) against central differences.
GPU : sba_rig_evaluate (analytic chain rule, csrc/rig.cu) against the oracle for all 11 camera models -- residuals to
      1e-10 of the pixel magnitudes, every Jacobian block to 1e-10 relative -- and the reference's own known answers
      (src/base/cost_functions_test.cc:101-132), exactly."""
import ctypes as C

import numpy as np
import pytest

dp = C.POINTER(C.c_double)
MODEL_PARAMS = {  # camera_models_test.cc:133-218 style parameter vectors
    0: [655.123, 386.123, 511.123], 1: [651.123, 655.123, 386.123, 511.123], 2: [651.123, 386.123, 511.123, 0.1],
    3: [651.123, 386.123, 511.123, 0.05, 0.03], 4: [651.123, 655.123, 386.123, 511.123, -0.471, 0.223, -0.001, 0.001],
    5: [651.123, 655.123, 386.123, 511.123, -0.471, 0.223, -0.001, 0.001],
    6: [651.123, 655.123, 386.123, 511.123, -0.471, 0.223, -0.001, 0.001, 0.001, 0.02, -0.02, 0.001],
    7: [651.123, 655.123, 386.123, 511.123, 0.9], 8: [651.123, 386.123, 511.123, 0.1], 9: [651.123, 386.123, 511.123, 0.05, 0.03],
    10: [651.123, 655.123, 386.123, 511.123, -0.471, 0.223, -0.001, 0.001, 0.001, 0.02, -0.02, 0.001]}


def _scene(rng, n_blocks, models):
    n_rig, n_rel, n_pt = 5, len(models), 40
    def quats(n, spread):
        q = np.concatenate([np.ones((n, 1)), rng.normal(0, spread, (n, 3))], axis=1)
        return q / np.linalg.norm(q, axis=1, keepdims=True)
    s = dict(rig_q=quats(n_rig, 0.1), rig_t=rng.normal(0, 0.3, (n_rig, 3)) + [0, 0, 6.0],
             rel_q=quats(n_rel, 0.05), rel_t=rng.normal(0, 0.1, (n_rel, 3)), pts=rng.uniform(-1, 1, (n_pt, 3)),
             model=np.asarray(models, np.int32), cam=np.zeros((n_rel, 12)))
    for c, m in enumerate(models):
        s["cam"][c, :len(MODEL_PARAMS[m])] = MODEL_PARAMS[m]
    s["b_rig"] = rng.integers(0, n_rig, n_blocks).astype(np.int32)
    s["b_rel"] = rng.integers(0, n_rel, n_blocks).astype(np.int32)
    s["b_cam"] = s["b_rel"].copy()  # camera c of the rig has relative pose c
    s["b_pt"] = rng.integers(0, n_pt, n_blocks).astype(np.int32)
    s["xy"] = rng.uniform(100, 700, (n_blocks, 2))
    return s


def _oracle_block(oracle, s, k, jac=True):
    r, J = np.zeros(2), np.zeros((2, 29))
    a = [np.ascontiguousarray(v) for v in (s["rig_q"][s["b_rig"][k]], s["rig_t"][s["b_rig"][k]], s["rel_q"][s["b_rel"][k]],
                                           s["rel_t"][s["b_rel"][k]], s["pts"][s["b_pt"][k]], s["cam"][s["b_cam"][k]], s["xy"][k])]
    oracle.oracle_eval_rig_observation_jac(int(s["model"][s["b_cam"][k]]), *[v.ctypes.data_as(dp) for v in a],
                                           r.ctypes.data_as(dp), J.ctypes.data_as(dp) if jac else None)
    return r, J


def test_oracle_rig_jacobian_matches_central_differences(oracle):
    rng = np.random.default_rng(5)
    s = _scene(rng, 22, list(range(11)))
    for k in range(22):
        r0, J = _oracle_block(oracle, s, k)
        blocks = [("rig_q", s["b_rig"][k], 0, 4), ("rig_t", s["b_rig"][k], 4, 3), ("rel_q", s["b_rel"][k], 7, 4),
                  ("rel_t", s["b_rel"][k], 11, 3), ("pts", s["b_pt"][k], 14, 3),
                  ("cam", s["b_cam"][k], 17, len(MODEL_PARAMS[int(s["model"][s["b_cam"][k]])]))]
        for name, idx, col, n in blocks:
            for j in range(n):
                h = 1e-6 * max(1.0, abs(s[name][idx, j]))
                keep = s[name][idx, j]
                s[name][idx, j] = keep + h; rp, _ = _oracle_block(oracle, s, k, jac=False)
                s[name][idx, j] = keep - h; rm, _ = _oracle_block(oracle, s, k, jac=False)
                s[name][idx, j] = keep
                fd = (rp - rm) / (2 * h)
                assert np.allclose(J[:, col + j], fd, rtol=2e-5, atol=1e-5 * max(1.0, np.abs(J[:, col + j]).max())), (name, j, k)


def _desc(sba, s):
    capi = sba.capi
    d = capi.RigDesc()
    d.num_rig_poses, d.num_rel_poses = len(s["rig_q"]), len(s["rel_q"])
    d.num_cameras, d.num_points, d.num_blocks = len(s["model"]), len(s["pts"]), len(s["b_rig"])
    d.rig_qvec, d.rig_tvec = capi.ptr(s["rig_q"], capi._dp), capi.ptr(s["rig_t"], capi._dp)
    d.rel_qvec, d.rel_tvec = capi.ptr(s["rel_q"], capi._dp), capi.ptr(s["rel_t"], capi._dp)
    d.camera_model, d.camera_params = capi.ptr(s["model"], capi._ip), capi.ptr(s["cam"], capi._dp)
    d.points, d.block_xy = capi.ptr(s["pts"], capi._dp), capi.ptr(s["xy"], capi._dp)
    d.block_rig, d.block_rel = capi.ptr(s["b_rig"], capi._ip), capi.ptr(s["b_rel"], capi._ip)
    d.block_camera, d.block_point = capi.ptr(s["b_cam"], capi._ip), capi.ptr(s["b_pt"], capi._ip)
    return d


def test_rig_evaluate_needs_a_device_or_validates(sba):
    lib = sba.capi.load()
    s = _scene(np.random.default_rng(1), 4, [0, 2])
    d = _desc(sba, s)
    res = np.zeros((4, 2))
    rc = lib.sba_rig_evaluate(C.byref(d), res.ctypes.data_as(dp), None, None)
    if lib.sba_device_count() < 1:
        assert rc == 2  # SBA_ERR_NO_DEVICE: no CPU fallback
    else:
        assert rc == 0
    assert lib.sba_rig_evaluate(None, res.ctypes.data_as(dp), None, None) == 1


@pytest.mark.gpu
def test_rig_known_answers_of_the_reference_test(sba):
    """cost_functions_test.cc:101-132: rig t = (0,0,-1), rel t = (0,0,1), identity rotations, SIMPLE_PINHOLE."""
    cases = [((0, 0, 1), 1, (0, 0)), ((0, 1, 1), 1, (0, 1)), ((0, 1, 1), 2, (0, 2)), ((-1, 1, 1), 2, (-2, 2))]
    s = dict(rig_q=np.array([[1.0, 0, 0, 0]]), rig_t=np.array([[0.0, 0, -1]]), rel_q=np.array([[1.0, 0, 0, 0]]),
             rel_t=np.array([[0.0, 0, 1]]), pts=np.array([c[0] for c in cases], float), model=np.zeros(2, np.int32),
             cam=np.zeros((2, 12)), b_rig=np.zeros(4, np.int32), b_rel=np.zeros(4, np.int32),
             b_cam=np.array([0, 0, 1, 1], np.int32), b_pt=np.arange(4, dtype=np.int32), xy=np.zeros((4, 2)))
    s["cam"][0, 0], s["cam"][1, 0] = 1.0, 2.0
    d = _desc(sba, s)
    res = np.zeros((4, 2))
    sba.capi.check(sba.capi.load().sba_rig_evaluate(C.byref(d), res.ctypes.data_as(dp), None, None))
    assert res.tolist() == [list(map(float, c[2])) for c in cases]


@pytest.mark.gpu
def test_rig_evaluate_matches_oracle_all_models(sba, oracle):
    rng = np.random.default_rng(9)
    n = 600
    s = _scene(rng, n, list(range(11)))
    d = _desc(sba, s)
    res, jac = np.zeros((n, 2)), np.zeros((n, 2, sba.capi.RIG_JAC_COLS))
    ms = C.c_double()
    sba.capi.check(sba.capi.load().sba_rig_evaluate(C.byref(d), res.ctypes.data_as(dp), jac.ctypes.data_as(dp), C.byref(ms)))
    for k in range(n):
        r, J = _oracle_block(oracle, s, k)
        scale = np.maximum(np.abs(s["xy"][k]), np.maximum(np.abs(r), 1.0))
        assert np.all(np.abs(res[k] - r) <= 1e-10 * 10 * scale), (k, res[k], r)
        for lo, hi in ((0, 4), (4, 7), (7, 11), (11, 14), (14, 17), (17, 29)):
            ref = J[:, lo:hi]
            assert np.linalg.norm(jac[k][:, lo:hi] - ref) <= 1e-10 * max(np.linalg.norm(ref), 1e-300), (k, lo)
            assert not jac[k][:, lo:hi][ref == 0].any()
