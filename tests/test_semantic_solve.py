"""The two semantic adjusters of the fork as SOLVES: the device paths
sba_semantic_solve / sba_cylinder_solve against the CPU oracle's restatement of
/root/reference/src/optim/semantic_bundle_adjustment.cc:430-544, 646-906 and
/root/reference/src/optim/geometric_semantic_bundle_adjustment.cc:481-604, 714-996, 1126-1213.
Both are "parity unpinned": the reference has no test for them; what is checked is CUDA == CPU restatement."""
import ctypes as C

import numpy as np
import pytest

from test_gpu_semantic import cylinder_scene, plane_scene


def _opts(capi, oracle, iters=6):
    o = capi.SolverOptions()
    oracle.oracle_default_solver_options(C.byref(o))
    o.max_num_iterations = iters
    # SemanticBundleAdjustmentOptions / GeometricSemanticBundleAdjustmentOptions (semantic_bundle_adjustment.h:116-118)
    o.function_tolerance = o.gradient_tolerance = o.parameter_tolerance = 1e-8
    return o


def semantic_problem(sba, q, t, cams, depth, label, pose_const=None, tvec_mask=None, step=4):
    capi = sba.capi
    p = capi.SemanticProblem()
    n = len(q)
    keep = [np.ascontiguousarray(q, np.float64).copy(), np.ascontiguousarray(t, np.float64).copy(), np.ascontiguousarray(cams, np.float64),
            np.ascontiguousarray(depth, np.float32), np.ascontiguousarray(label, np.float32),
            np.zeros(n, np.uint8) if pose_const is None else np.ascontiguousarray(pose_const, np.uint8),
            np.zeros(n, np.uint8) if tvec_mask is None else np.ascontiguousarray(tvec_mask, np.uint8)]
    p.num_images, p.height, p.width = depth.shape
    p.image_qvec, p.image_tvec, p.camera_params = (capi.ptr(keep[k], capi._dp) for k in range(3))
    p.depth_maps, p.semantic_maps = capi.ptr(keep[3], capi._fp), capi.ptr(keep[4], capi._fp)
    p.image_pose_constant, p.image_tvec_constant_mask = capi.ptr(keep[5], capi._u8p), capi.ptr(keep[6], capi._u8p)
    p.pixel_step, p.depth_error_threshold, p.relative_step_size = step, 2.0, 1e-3
    return p, keep


def _run(fn, prob, opts, *extra, cap=64):
    s_type = type(opts)
    import importlib
    capi = importlib.import_module("semantic-bundle-adjustment-colmap_b200").capi
    s = capi.Summary()
    tr = (capi.IterationSummary * cap)()
    n = C.c_int32()
    rc = fn(C.byref(prob), C.byref(opts), *extra, C.byref(s), tr, cap, C.byref(n))
    return rc, s, [tr[i] for i in range(n.value)]


def _perturbed_scene(sba, n_img=4, seed=0, **kw):
    """plane_scene with the translations off by ~0.05 (pose 0 exact): the 0/1 label-mismatch objective with central
    differences only descends from a mild perturbation -- as in the reference, whose objective this is."""
    q, t, cams, depth, label = plane_scene(sba, n_img=n_img, seed=seed, **kw)
    rng = np.random.default_rng(seed + 100)
    t0 = t + rng.normal(0, 0.05, t.shape)
    t0[0] = t[0]
    return q, t0, cams, depth, label


# ------------------------------------------------------------------ CPU
def test_oracle_semantic_solve_structure_and_descent(sba, oracle):
    q, t, cams, depth, label = _perturbed_scene(sba)
    pc = np.array([1, 0, 0, 0], np.uint8)
    tm = np.array([0, 1, 0, 0], np.uint8)
    p, keep = semantic_problem(sba, q, t, cams, depth, label, pc, tm)
    rc, s, tr = _run(oracle.oracle_semantic_solve, p, _opts(sba.capi, oracle, 14), 0)
    assert rc == 0
    H, W = depth.shape[1:]
    per_image = [(depth[i, ::4, ::4] >= 1e-4).sum() for i in range(4)]
    assert s.num_residual_blocks == sum(per_image[a] for a in range(4) for b in range(4) if a != b)
    assert s.num_effective_parameters_reduced == 3 * 6 - 1 and s.num_parameters_reduced == 3 * 7
    assert s.final_cost < s.initial_cost and s.num_successful_steps >= 1
    assert np.allclose(np.linalg.norm(keep[0], axis=1), 1.0, atol=1e-2)  # poses written back
    assert np.array_equal(keep[0][0], q[0] / np.linalg.norm(q[0])) and np.array_equal(keep[1][0], t[0])  # constant pose untouched
    assert keep[1][1, 0] == t[1, 0]                                        # constant tvec component untouched


def test_oracle_semantic_solve_empty_when_all_constant(sba, oracle):
    q, t, cams, depth, label = _perturbed_scene(sba, n_img=3)
    p, keep = semantic_problem(sba, q, t, cams, depth, label, np.ones(3, np.uint8))
    rc, s, tr = _run(oracle.oracle_semantic_solve, p, _opts(sba.capi, oracle), 0)
    assert rc == sba.capi.ERR_EMPTY_PROBLEM if hasattr(sba.capi, "ERR_EMPTY_PROBLEM") else rc == 5


def cylinder_problem(sba, oracle, kind=0, refine_geometry=True, pose_const=None, seed=0, H=120, W=160):
    capi = sba.capi
    q, t, cams, cq, ct, cr, ch = cylinder_scene(sba, H=H, W=W, seed=seed)
    n_img, n_cyl = len(q), len(cq)
    # masks: the oracle's rasterisation of the true cylinders seen from slightly different poses
    masks = np.zeros((n_img, H, W), np.uint8)
    pts = np.zeros(8)
    for i in range(n_img):
        for c in range(n_cyl):
            ok = oracle.oracle_cylinder_quadrilateral(capi.ptr(q[i].copy(), capi._dp), capi.ptr((t[i] + 0.03).copy(), capi._dp),
                                                      capi.ptr(cams[i].copy(), capi._dp), capi.ptr(cq[c].copy(), capi._dp),
                                                      capi.ptr(ct[c].copy(), capi._dp), cr[c] * 1.1, ch[c], capi.ptr(pts, capi._dp))
            assert ok == 1
            m = np.zeros((H, W), np.uint8)
            oracle.oracle_draw_quadrilateral(capi.ptr(pts, capi._dp), H, W, m.ctypes.data_as(C.POINTER(C.c_uint8)))
            masks[i] |= m
    cyl = np.zeros((n_cyl, 9))
    if kind == 0:
        cyl[:, :4], cyl[:, 4:7], cyl[:, 7], cyl[:, 8] = cq, ct, cr, ch
    else:  # two points: bottom centre and top centre (axis = R(q) z)
        R = sba.scene.rotmat_from_quat(cq / np.linalg.norm(cq, axis=1, keepdims=True))
        cyl[:, :3] = ct
        cyl[:, 3:6] = ct + R[:, :, 2] * ch[:, None]
        cyl[:, 6] = cr
    p = capi.CylinderProblem()
    keep = [np.ascontiguousarray(q).copy(), np.ascontiguousarray(t).copy(), np.ascontiguousarray(cams), masks,
            np.zeros(n_img, np.uint8) if pose_const is None else np.ascontiguousarray(pose_const, np.uint8),
            np.zeros(n_img, np.uint8), np.ascontiguousarray(cyl)]
    p.num_images, p.height, p.width = n_img, H, W
    p.image_qvec, p.image_tvec, p.camera_params = (capi.ptr(keep[k], capi._dp) for k in range(3))
    p.masks = capi.ptr(keep[3], capi._u8p)
    p.image_pose_constant, p.image_tvec_constant_mask = capi.ptr(keep[4], capi._u8p), capi.ptr(keep[5], capi._u8p)
    p.num_cylinders, p.cylinder_kind, p.cylinder_params = n_cyl, kind, capi.ptr(keep[6], capi._dp)
    p.refine_geometry, p.relative_step_size = 1 if refine_geometry else 0, 1e-3
    return p, keep


@pytest.mark.parametrize("kind", [0, 1])
def test_oracle_cylinder_solve_structure_and_descent(sba, oracle, kind):
    p, keep = cylinder_problem(sba, oracle, kind=kind, pose_const=[1, 0, 0])
    rc, s, tr = _run(oracle.oracle_cylinder_solve, p, _opts(sba.capi, oracle, 8), 0)
    assert rc == 0 and s.num_residual_blocks == 9
    assert s.num_effective_parameters_reduced == 2 * 6 + 3 * (8 if kind == 0 else 7)
    assert 0 < s.initial_cost < 0.5 * 9 / 3 and s.final_cost <= s.initial_cost
    assert (keep[6][:, 7 if kind == 0 else 6] >= 0).all()  # radius lower bound


def test_oracle_cylinder_constant_geometry_and_constant_cameras(sba, oracle):
    p, keep = cylinder_problem(sba, oracle, refine_geometry=False, pose_const=[1, 0, 0])
    cyl0 = keep[6].copy()
    rc, s, tr = _run(oracle.oracle_cylinder_solve, p, _opts(sba.capi, oracle, 5), 0)
    assert rc == 0 and s.num_residual_blocks == 6 and s.num_effective_parameters_reduced == 12  # constant camera skipped
    assert np.array_equal(keep[6], cyl0)
    p, keep = cylinder_problem(sba, oracle, refine_geometry=True, pose_const=[1, 1, 1])
    rc, s, tr = _run(oracle.oracle_cylinder_solve, p, _opts(sba.capi, oracle, 5), 0)
    assert rc == 0 and s.num_residual_blocks == 9 and s.num_effective_parameters_reduced == 24


# ------------------------------------------------------------------ GPU
def _same_trace(trg, tro, rel=1e-9):
    assert len(trg) == len(tro)
    for a, b in zip(trg, tro):
        assert a.step_is_valid == b.step_is_valid and a.step_is_successful == b.step_is_successful
        assert a.cost == pytest.approx(b.cost, rel=rel, abs=1e-12)
        assert a.trust_region_radius == pytest.approx(b.trust_region_radius, rel=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("variant", ["all_variable", "constant_first", "tvec_subset", "soft_l1"])
def test_semantic_solve_parity(sba, oracle, variant):
    q, t, cams, depth, label = _perturbed_scene(sba, n_img=5, seed=3)
    pc = tm = None
    if variant == "constant_first":
        pc = np.array([1, 0, 0, 1, 0], np.uint8)
    if variant == "tvec_subset":
        pc = np.array([1, 0, 0, 0, 0], np.uint8); tm = np.array([0, 1, 0, 6, 0], np.uint8)
    o = _opts(sba.capi, oracle, 12)
    if variant == "soft_l1":
        o.loss_function_type, o.loss_function_scale = 1, 0.7
    po, ko = semantic_problem(sba, q, t, cams, depth, label, pc, tm)
    pg, kg = semantic_problem(sba, q, t, cams, depth, label, pc, tm)
    rco, so, tro = _run(oracle.oracle_semantic_solve, po, o, 0)
    lib = sba.capi.load()
    rcg, sg, trg = _run(lib.sba_semantic_solve, pg, o, C.cast(None, sba.capi.ITERATION_CALLBACK), None)
    assert rco == 0 and rcg == 0
    assert (sg.num_residual_blocks, sg.num_effective_parameters_reduced, sg.num_parameters_reduced) == \
           (so.num_residual_blocks, so.num_effective_parameters_reduced, so.num_parameters_reduced)
    assert sg.initial_cost == pytest.approx(so.initial_cost, rel=1e-12) and so.num_successful_steps >= 1
    _same_trace(trg, tro)
    assert np.allclose(kg[0], ko[0], atol=1e-9) and np.allclose(kg[1], ko[1], atol=1e-9)


@pytest.mark.gpu
@pytest.mark.parametrize("kind,refine,pc", [(0, True, [1, 0, 0]), (1, True, [1, 0, 0]), (0, False, [0, 0, 0]), (1, True, [1, 1, 1]),
                                            (0, True, [0, 0, 0])])
def test_cylinder_solve_parity(sba, oracle, kind, refine, pc):
    """All five functors: constant camera (both cylinder kinds), both variable (both kinds), constant geometry."""
    o = _opts(sba.capi, oracle, 6)
    po, ko = cylinder_problem(sba, oracle, kind=kind, refine_geometry=refine, pose_const=pc)
    pg, kg = cylinder_problem(sba, oracle, kind=kind, refine_geometry=refine, pose_const=pc)
    rco, so, tro = _run(oracle.oracle_cylinder_solve, po, o, 0)
    lib = sba.capi.load()
    rcg, sg, trg = _run(lib.sba_cylinder_solve, pg, o, C.cast(None, sba.capi.ITERATION_CALLBACK), None)
    assert rco == 0 and rcg == 0
    assert (sg.num_residual_blocks, sg.num_effective_parameters_reduced) == (so.num_residual_blocks, so.num_effective_parameters_reduced)
    assert sg.initial_cost == pytest.approx(so.initial_cost, rel=1e-12)
    _same_trace(trg, tro, rel=1e-7)
    assert np.allclose(kg[6], ko[6], atol=1e-6) and np.allclose(kg[1], ko[1], atol=1e-6)


@pytest.mark.gpu
def test_semantic_solve_callback_and_errors(sba, oracle):
    q, t, cams, depth, label = _perturbed_scene(sba, n_img=3, seed=5)
    capi = sba.capi
    lib = capi.load()
    o = _opts(capi, oracle, 6)
    rows = []

    def cb(it, _):
        rows.append(it.contents.iteration)
        return 1 if it.contents.iteration >= 2 else 0
    pg, kg = semantic_problem(sba, q, t, cams, depth, label)
    rc, s, tr = _run(lib.sba_semantic_solve, pg, o, capi.ITERATION_CALLBACK(cb), None)
    assert rc == 0 and s.termination_type == capi.USER_SUCCESS and rows == [0, 1, 2]
    pg, kg = semantic_problem(sba, q, t, cams, depth, label, np.ones(3, np.uint8))
    rc, s, tr = _run(lib.sba_semantic_solve, pg, o, C.cast(None, capi.ITERATION_CALLBACK), None)
    assert rc == 5  # SBA_ERR_EMPTY_PROBLEM: Solve() returns false (semantic_bundle_adjustment.cc:484-486)


@pytest.mark.gpu
def test_python_wrappers(sba, oracle):
    """semantic_adjusters.SemanticProblem / CylinderProblem: the Python mirror over the same two entry points."""
    q, t, cams, depth, label = _perturbed_scene(sba, n_img=3, seed=1)
    prob = sba.semantic_adjusters.SemanticProblem(q, t, cams, depth, label, image_pose_constant=[1, 0, 0], pixel_step=4)
    s, rows = prob.solve(sba.semantic_adjusters.semantic_solver_options(10))
    assert s is not None and s.final_cost <= s.initial_cost and len(rows) == s.num_iterations
    assert np.array_equal(prob.image_tvec[0], t[0]) and not np.array_equal(prob.image_tvec, t) or s.num_successful_steps == 0
    pc, keep = cylinder_problem(sba, oracle, kind=0)
    cp = sba.semantic_adjusters.CylinderProblem(keep[0], keep[1], keep[2], keep[3], keep[6], image_pose_constant=[1, 0, 0])
    s, rows = cp.solve(sba.semantic_adjusters.semantic_solver_options(6))
    assert s is not None and s.num_residual_blocks == 9 and s.final_cost <= s.initial_cost
    none, _ = sba.semantic_adjusters.SemanticProblem(q, t, cams, depth, label, image_pose_constant=[1, 1, 1]).solve()
    assert none is None


@pytest.mark.gpu
def test_per_iteration_dump_directories(sba, oracle, tmp_path):
    """SBACallbackFunctor / the GSBA callback (semantic_bundle_adjustment.cc:1104-1119,
    geometric_semantic_bundle_adjustment.cc:1546-1556): step_k/ holds the binary model, step_k/text the text model (and
    cylinders.txt), with the poses of iteration k -- the last one equals the solve's result."""
    ba, io, sa = sba.bundle_adjustment, sba.model_io, sba.semantic_adjusters
    q, t, cams, depth, label = _perturbed_scene(sba, n_img=3, seed=1)
    rec = ba.Reconstruction()
    H, W = depth.shape[1:]
    for k in range(3):
        rec.AddCamera(ba.Camera(k + 1, "SIMPLE_PINHOLE", list(cams[k]), W, H))
        rec.AddImage(ba.Image(10 + k, k + 1, q[k], t[k], name="frame_%d.png" % k))
    prob = sa.SemanticProblem(q, t, cams, depth, label, image_pose_constant=[1, 0, 0], pixel_step=4)
    dump = sa.StepDumper(prob, rec, [10, 11, 12], str(tmp_path / "optim_steps"))
    s, rows = prob.solve(sa.semantic_solver_options(5), callback=dump)
    assert s is not None and dump.steps == [r.iteration for r in rows] and dump.steps[0] == 0
    first = io.read_model(str(tmp_path / "optim_steps" / "step_0"))
    assert np.allclose(first.Image(11).Tvec(), t[1], rtol=0, atol=0)          # iteration 0 = the starting point
    last = io.read_model(str(tmp_path / "optim_steps" / ("step_%d" % dump.steps[-1])))
    txt = io.read_model(str(tmp_path / "optim_steps" / ("step_%d" % dump.steps[-1]) / "text"))
    for k in range(3):
        assert np.array_equal(last.Image(10 + k).Tvec(), prob.image_tvec[k])   # binary: exact
        assert np.array_equal(last.Image(10 + k).Qvec(), prob.image_qvec[k])
        assert np.allclose(txt.Image(10 + k).Tvec(), prob.image_tvec[k], rtol=1e-15, atol=0)
    if s.num_successful_steps > 0:
        assert not np.array_equal(last.Image(11).Tvec(), first.Image(11).Tvec())
    # cylinders
    pc, keep = cylinder_problem(sba, oracle, kind=0)
    cp = sa.CylinderProblem(keep[0], keep[1], keep[2], keep[3], keep[6], image_pose_constant=[1, 0, 0])
    rec2 = ba.Reconstruction()
    for k in range(len(keep[0])):
        rec2.AddCamera(ba.Camera(k + 1, "SIMPLE_PINHOLE", list(np.asarray(keep[2]).reshape(-1, 3)[k]), keep[3].shape[2], keep[3].shape[1]))
        rec2.AddImage(ba.Image(k + 1, k + 1, np.asarray(keep[0]).reshape(-1, 4)[k], np.asarray(keep[1]).reshape(-1, 3)[k], name="c%d.png" % k))
    dump2 = sa.StepDumper(cp, rec2, [k + 1 for k in range(len(keep[0]))], str(tmp_path / "gsba_steps"))
    s2, rows2 = cp.solve(sa.semantic_solver_options(4), callback=dump2)
    assert s2 is not None and len(dump2.steps) == len(rows2)
    cyl = io.read_cylinders(str(tmp_path / "gsba_steps" / ("step_%d" % dump2.steps[-1]) / "cylinders.txt"))
    assert np.allclose(np.asarray(cyl).reshape(-1, 9), cp.cylinder_params, rtol=1e-5, atol=1e-6)  # "%g": 6 digits


# ------------------------------------------------------------------ a13: optional landmark blocks (include_landmark_error)
def add_landmarks(sba, p, keep, n_pts=40, seed=0, weight=1.0, const_points=()):
    """Reprojection blocks of every point in every image (SIMPLE_PINHOLE, the cylinder scene's cameras) with 1 px noise,
    ScaledLoss weight = landmark_error_weight / number of 2-D features (geometric_semantic_bundle_adjustment.cc:758)."""
    capi = sba.capi
    rng = np.random.default_rng(seed + 50)
    q, t, cams = keep[0].reshape(-1, 4), keep[1].reshape(-1, 3), np.asarray(keep[2]).reshape(-1, 3)
    n_img = len(q)
    X = np.column_stack([rng.uniform(-1.5, 1.5, n_pts), rng.uniform(-1.0, 1.0, n_pts), rng.uniform(4, 7, n_pts)])
    li, lp, lxy = [], [], []
    for i in range(n_img):
        R = sba.scene.rotmat_from_quat((q[i] / np.linalg.norm(q[i]))[None])[0]
        Pc = X @ R.T + t[i]
        xy = cams[i, 0] * Pc[:, :2] / Pc[:, 2:3] + cams[i, 1:3] + rng.normal(0, 1.0, (n_pts, 2))
        li += [i] * n_pts; lp += list(range(n_pts)); lxy.append(xy)
    X += rng.normal(0, 0.05, X.shape)   # the points start off their true positions
    pc = np.zeros(n_pts, np.uint8); pc[list(const_points)] = 1
    extra = [np.ascontiguousarray(X), pc, np.asarray(li, np.int32), np.asarray(lp, np.int32), np.ascontiguousarray(np.vstack(lxy))]
    p.num_points, p.points, p.point_constant = n_pts, capi.ptr(extra[0], capi._dp), capi.ptr(extra[1], capi._u8p)
    p.num_landmarks, p.landmark_image, p.landmark_point = len(li), capi.ptr(extra[2], capi._ip), capi.ptr(extra[3], capi._ip)
    p.landmark_xy, p.landmark_weight = capi.ptr(extra[4], capi._dp), weight / len(li)
    keep.append(extra)
    return p, keep


def test_oracle_cylinder_solve_with_landmarks(sba, oracle):
    p, keep = cylinder_problem(sba, oracle, kind=0, pose_const=[1, 0, 0])
    p, keep = add_landmarks(sba, p, keep, const_points=(0, 1))
    X0 = keep[-1][0].copy()
    rc, s, tr = _run(oracle.oracle_cylinder_solve, p, _opts(sba.capi, oracle, 8), 0)
    assert rc == 0
    # 9 IoU blocks + 3 x 40 landmark blocks of two rows; 2 poses x 6 + 3 cylinders x 8 + 38 variable points x 3
    assert (s.num_residual_blocks, s.num_residuals_reduced) == (9 + 120, 9 + 240)
    assert s.num_effective_parameters_reduced == 12 + 24 + 38 * 3
    assert s.final_cost < s.initial_cost
    X1 = keep[-1][0]
    assert np.array_equal(X1[:2], X0[:2]) and not np.array_equal(X1[2:], X0[2:])   # constant points stay, the others move


@pytest.mark.gpu
@pytest.mark.parametrize("kind,pc,weight", [(0, [1, 0, 0], 1.0), (1, [0, 0, 0], 50.0)])
def test_cylinder_solve_with_landmarks_parity(sba, oracle, kind, pc, weight):
    o = _opts(sba.capi, oracle, 6)
    po, ko = add_landmarks(sba, *cylinder_problem(sba, oracle, kind=kind, pose_const=pc), weight=weight, const_points=(3,))
    pg, kg = add_landmarks(sba, *cylinder_problem(sba, oracle, kind=kind, pose_const=pc), weight=weight, const_points=(3,))
    rco, so, tro = _run(oracle.oracle_cylinder_solve, po, o, 0)
    rcg, sg, trg = _run(sba.capi.load().sba_cylinder_solve, pg, o, C.cast(None, sba.capi.ITERATION_CALLBACK), None)
    assert rco == 0 and rcg == 0
    assert (sg.num_residual_blocks, sg.num_residuals_reduced, sg.num_effective_parameters_reduced) == \
           (so.num_residual_blocks, so.num_residuals_reduced, so.num_effective_parameters_reduced)
    assert sg.initial_cost == pytest.approx(so.initial_cost, rel=1e-12)
    _same_trace(trg, tro, rel=1e-7)
    assert np.allclose(kg[-1][0], ko[-1][0], atol=1e-6) and np.allclose(kg[1], ko[1], atol=1e-6) and np.allclose(kg[6], ko[6], atol=1e-6)
    assert np.array_equal(kg[-1][0][3], ko[-1][0][3])
    # the Python wrapper carries the same blocks
    e = kg[-1]
    cp = sba.semantic_adjusters.CylinderProblem(*cylinder_problem(sba, oracle, kind=kind, pose_const=pc)[1][:4],
                                                cylinder_problem(sba, oracle, kind=kind)[1][6], cylinder_kind=kind, image_pose_constant=pc,
                                                points=add_landmarks(sba, *cylinder_problem(sba, oracle, kind=kind))[1][-1][0],
                                                point_constant=e[1], landmark_image=e[2], landmark_point=e[3], landmark_xy=e[4],
                                                landmark_error_weight=weight)
    s2, rows2 = cp.solve(sba.semantic_adjusters.semantic_solver_options(6))
    assert s2 is not None and s2.num_residual_blocks == sg.num_residual_blocks and s2.final_cost == pytest.approx(sg.final_cost, rel=1e-9)
