"""Python host side of the fork's two semantic adjusters: flat problems for `sba_semantic_solve` / `sba_cylinder_solve`
(include/sba_b200.h) -- the reference's SemanticBundleAdjuster (/root/reference/src/optim/semantic_bundle_adjustment.cc:430-544,
646-906) and GeometricSemanticBundleAdjuster (/root/reference/src/optim/geometric_semantic_bundle_adjustment.cc:481-604,
714-996).  Images are given in the iteration order of config.Images() (= residual order); poses (and cylinders) are
updated in place, as the reference mutates the Reconstruction.  Nothing here computes residuals on the CPU."""
import ctypes as C
import os

import numpy as np

from . import _capi as capi
from . import model_io


def semantic_solver_options(max_num_iterations=100):
    """SemanticBundleAdjustmentOptions' solver options (semantic_bundle_adjustment.h:115-131): tolerances 1e-8."""
    o = capi.default_solver_options()
    o.function_tolerance = o.gradient_tolerance = o.parameter_tolerance = 1e-8
    o.max_num_iterations = max_num_iterations
    o.update_state_every_iteration = 1
    return o


class StepDumper:
    """Per-iteration dump of the two semantic adjusters: SBACallbackFunctor (semantic_bundle_adjustment.cc:1090-1123) and the
    GSBA callback (geometric_semantic_bundle_adjustment.cc:1482-1558) write, after EVERY iteration (0 = the starting point),
        <optim_steps_folder>/step_<iteration>/{cameras,images,points3D}.bin      Reconstruction::Write
        <optim_steps_folder>/step_<iteration>/text/{cameras,images,points3D}.txt Reconstruction::WriteText
        <optim_steps_folder>/step_<iteration>/cylinders.txt                      exportCylindersToText (GSBA only)
    which needs solver_options.update_state_every_iteration (semantic_bundle_adjustment.h:129): the device solve then
    copies the current poses (and cylinders) into the problem's arrays before each callback.  Use as the `callback` of
    SemanticProblem.solve / CylinderProblem.solve; `image_ids[k]` is the Reconstruction image behind the problem's k-th pose.
    The reference's JPG mask renderings and CSV error export are not written (no image codec in this package)."""

    def __init__(self, problem, reconstruction, image_ids, optim_steps_folder, inner=None):
        self.problem, self.rec, self.image_ids, self.folder, self.inner = problem, reconstruction, list(image_ids), optim_steps_folder, inner
        self.steps = []

    def __call__(self, it):
        for k, image_id in enumerate(self.image_ids):
            im = self.rec.Image(image_id)
            im.Qvec()[:] = self.problem.image_qvec[k]
            im.Tvec()[:] = self.problem.image_tvec[k]
        step = os.path.join(self.folder, "step_%d" % it.iteration)
        model_io.write_model(self.rec, step, ".bin")
        model_io.write_model(self.rec, os.path.join(step, "text"), ".txt")
        if hasattr(self.problem, "cylinder_params") and getattr(self.problem, "cylinder_kind", 0) == 0:
            model_io.write_cylinders(os.path.join(step, "cylinders.txt"), self.problem.cylinder_params)
        self.steps.append(it.iteration)
        return int(self.inner(it) or 0) if self.inner is not None else 0


class _PoseProblem:
    def _poses(self, image_qvec, image_tvec, camera_params, image_pose_constant, image_tvec_constant_mask):
        self.image_qvec = capi.as_f64(image_qvec).reshape(-1, 4).copy()
        self.image_tvec = capi.as_f64(image_tvec).reshape(-1, 3).copy()
        n = len(self.image_qvec)
        self.camera_params = capi.as_f64(camera_params).reshape(n, 3)
        self.image_pose_constant = (np.zeros(n, np.uint8) if image_pose_constant is None
                                    else np.ascontiguousarray(image_pose_constant, dtype=np.uint8))
        self.image_tvec_constant_mask = (np.zeros(n, np.uint8) if image_tvec_constant_mask is None
                                         else np.ascontiguousarray(image_tvec_constant_mask, dtype=np.uint8))
        return n

    def _solve(self, fn, desc, options, callback, cap=256):
        o = options if options is not None else semantic_solver_options()
        s = capi.Summary()
        rows = (capi.IterationSummary * cap)()
        n = C.c_int32()
        if callback is not None:
            cb = capi.ITERATION_CALLBACK(lambda it, _u: int(callback(it.contents) or 0))
        else:
            cb = C.cast(None, capi.ITERATION_CALLBACK)
        rc = fn(C.byref(desc), C.byref(o), cb, None, C.byref(s), rows, cap, C.byref(n))
        if rc == capi.ERR_EMPTY_PROBLEM:  # Solve() returns false
            return None, []
        capi.check(rc)
        return s, [rows[i] for i in range(n.value)]


class SemanticProblem(_PoseProblem):
    """Pose-only semantic BA: one 0/1 label-mismatch residual per (ordered image pair, grid pixel with depth)."""

    def __init__(self, image_qvec, image_tvec, camera_params, depth_maps, semantic_maps, image_pose_constant=None,
                 image_tvec_constant_mask=None, pixel_step=10, depth_error_threshold=2.0, relative_step_size=1e-3):
        n = self._poses(image_qvec, image_tvec, camera_params, image_pose_constant, image_tvec_constant_mask)
        self.depth_maps = np.ascontiguousarray(depth_maps, dtype=np.float32)
        self.semantic_maps = np.ascontiguousarray(semantic_maps, dtype=np.float32)
        assert self.depth_maps.shape == self.semantic_maps.shape and self.depth_maps.shape[0] == n
        self.pixel_step, self.depth_error_threshold, self.relative_step_size = pixel_step, depth_error_threshold, relative_step_size

    def desc(self):
        p = capi.SemanticProblem()
        p.num_images, p.height, p.width = self.depth_maps.shape
        p.image_qvec, p.image_tvec = capi.ptr(self.image_qvec, capi._dp), capi.ptr(self.image_tvec, capi._dp)
        p.camera_params = capi.ptr(self.camera_params, capi._dp)
        p.depth_maps, p.semantic_maps = capi.ptr(self.depth_maps, capi._fp), capi.ptr(self.semantic_maps, capi._fp)
        p.image_pose_constant = capi.ptr(self.image_pose_constant, capi._u8p)
        p.image_tvec_constant_mask = capi.ptr(self.image_tvec_constant_mask, capi._u8p)
        p.pixel_step, p.depth_error_threshold, p.relative_step_size = self.pixel_step, self.depth_error_threshold, self.relative_step_size
        return p

    def solve(self, options=None, callback=None):
        """(summary, iteration table); (None, []) when the problem has no residual block.  Poses change in place."""
        return self._solve(capi.load().sba_semantic_solve, self.desc(), options, callback)


class CylinderProblem(_PoseProblem):
    """Geometric-semantic BA: one residual 1 - IoU per (image, cylinder); cylinder_kind 0 = (q, t, r, h), 1 = (t1, t2, r)."""

    def __init__(self, image_qvec, image_tvec, camera_params, masks, cylinder_params, cylinder_kind=0, refine_geometry=True,
                 image_pose_constant=None, image_tvec_constant_mask=None, relative_step_size=1e-3, points=None,
                 point_constant=None, landmark_image=None, landmark_point=None, landmark_xy=None, landmark_error_weight=1.0):
        n = self._poses(image_qvec, image_tvec, camera_params, image_pose_constant, image_tvec_constant_mask)
        self.masks = np.ascontiguousarray(masks, dtype=np.uint8)
        assert self.masks.shape[0] == n
        self.cylinder_params = np.zeros((len(cylinder_params), 9))
        cp = capi.as_f64(cylinder_params).reshape(len(cylinder_params), -1)
        self.cylinder_params[:, :cp.shape[1]] = cp
        self.cylinder_kind, self.refine_geometry, self.relative_step_size = cylinder_kind, refine_geometry, relative_step_size
        # optional landmark blocks (include_landmark_error): every listed 2-D feature of the config images
        self.points = None if points is None else capi.as_f64(points).reshape(-1, 3).copy()
        if self.points is not None:
            self.point_constant = (np.zeros(len(self.points), np.uint8) if point_constant is None
                                   else np.ascontiguousarray(point_constant, dtype=np.uint8))
            self.landmark_image, self.landmark_point = capi.as_i32(landmark_image), capi.as_i32(landmark_point)
            self.landmark_xy = capi.as_f64(landmark_xy).reshape(-1, 2)
            self.landmark_weight = landmark_error_weight / max(1, len(self.landmark_image))  # (:758: weight / total_num_2d_features)

    def desc(self):
        p = capi.CylinderProblem()
        p.num_images, p.height, p.width = self.masks.shape
        p.image_qvec, p.image_tvec = capi.ptr(self.image_qvec, capi._dp), capi.ptr(self.image_tvec, capi._dp)
        p.camera_params, p.masks = capi.ptr(self.camera_params, capi._dp), capi.ptr(self.masks, capi._u8p)
        p.image_pose_constant = capi.ptr(self.image_pose_constant, capi._u8p)
        p.image_tvec_constant_mask = capi.ptr(self.image_tvec_constant_mask, capi._u8p)
        p.num_cylinders, p.cylinder_kind = len(self.cylinder_params), self.cylinder_kind
        p.cylinder_params = capi.ptr(self.cylinder_params, capi._dp)
        p.refine_geometry, p.relative_step_size = 1 if self.refine_geometry else 0, self.relative_step_size
        if self.points is not None:
            p.num_points, p.points, p.point_constant = len(self.points), capi.ptr(self.points, capi._dp), capi.ptr(self.point_constant, capi._u8p)
            p.num_landmarks = len(self.landmark_image)
            p.landmark_image, p.landmark_point = capi.ptr(self.landmark_image, capi._ip), capi.ptr(self.landmark_point, capi._ip)
            p.landmark_xy, p.landmark_weight = capi.ptr(self.landmark_xy, capi._dp), self.landmark_weight
        return p

    def solve(self, options=None, callback=None):
        return self._solve(capi.load().sba_cylinder_solve, self.desc(), options, callback)
