"""Host-side mirror of the reference's BundleAdjuster front-end.

Same names, argument meaning and error behaviour as
  BundleAdjustmentOptions   /root/reference/src/optim/bundle_adjustment.h:49-100
  BundleAdjustmentConfig    /root/reference/src/optim/bundle_adjustment.h:103-167
  BundleAdjuster            /root/reference/src/optim/bundle_adjustment.h:171-203,
                            /root/reference/src/optim/bundle_adjustment.cc:252-530
over a small in-memory data model with the accessor contract of the reference's
Reconstruction / Image / Camera / Point2D / Point3D / Track
(src/base/{reconstruction,image,camera,point2d,point3d,track}.h).

Solve() flattens the reconstruction in the reference's residual order (SetUp,
bundle_adjustment.cc:326-342), hands the arrays to the CUDA library through the C ABI, and
scatters the optimised parameters back in place.  There is no CPU solve here.
"""
from collections import OrderedDict

import numpy as np

from . import _capi as capi
from .problem import FlatProblem, Solver

kInvalidPoint3DId = (1 << 64) - 1


class TrackElement:
    def __init__(self, image_id, point2D_idx):
        self.image_id, self.point2D_idx = image_id, point2D_idx


class Track:
    def __init__(self):
        self.elements = []

    def Length(self):
        return len(self.elements)

    def Elements(self):
        return self.elements

    def AddElement(self, image_id, point2D_idx):
        self.elements.append(TrackElement(image_id, point2D_idx))


class Point2D:
    def __init__(self, xy, point3D_id=kInvalidPoint3DId):
        self.xy = np.asarray(xy, dtype=np.float64)
        self.point3D_id = point3D_id

    def XY(self):
        return self.xy

    def HasPoint3D(self):
        return self.point3D_id != kInvalidPoint3DId

    def Point3DId(self):
        return self.point3D_id


class Point3D:
    def __init__(self, xyz, color=(0, 0, 0), error=-1.0):
        self.xyz = np.asarray(xyz, dtype=np.float64).copy()
        self.track = Track()
        self.color, self.error = tuple(int(c) for c in color), float(error)  # src/base/point3d.h (Color, Error)

    def XYZ(self):
        return self.xyz

    def Track(self):
        return self.track


class Camera:
    def __init__(self, camera_id, model, params, width=0, height=0):
        self.camera_id, self.width, self.height = camera_id, int(width), int(height)
        self.model_id = capi.CAMERA_MODELS[model] if isinstance(model, str) else int(model)
        if self.model_id not in capi.MODEL_NUM_PARAMS:
            raise ValueError("Camera model does not exist")  # std::domain_error, camera_models.h:140
        self.params = np.asarray(params, dtype=np.float64).copy()

    def ModelId(self):
        return self.model_id

    def Params(self):
        return self.params

    def NumParams(self):
        return len(self.params)

    def FocalLengthIdxs(self):
        return capi.MODEL_FOCAL_IDXS[self.model_id]

    def PrincipalPointIdxs(self):
        return capi.MODEL_PP_IDXS[self.model_id]

    def ExtraParamsIdxs(self):
        return capi.MODEL_EXTRA_IDXS[self.model_id]


class Image:
    def __init__(self, image_id, camera_id, qvec=(1, 0, 0, 0), tvec=(0, 0, 0), name=""):
        self.image_id, self.camera_id, self.name = image_id, camera_id, name
        self.qvec = np.asarray(qvec, dtype=np.float64).copy()
        self.tvec = np.asarray(tvec, dtype=np.float64).copy()
        self.points2D = []

    def CameraId(self):
        return self.camera_id

    def Qvec(self):
        return self.qvec

    def Tvec(self):
        return self.tvec

    def NormalizeQvec(self):  # src/base/pose.cc:82-91
        n = np.linalg.norm(self.qvec)
        if n == 0:
            self.qvec[:] = (1, 0, 0, 0)
        else:
            self.qvec /= n

    def Points2D(self):
        return self.points2D

    def Point2D(self, idx):
        return self.points2D[idx]


class Reconstruction:
    def __init__(self):
        self.cameras, self.images, self.points3D = OrderedDict(), OrderedDict(), OrderedDict()
        self._next_point3D_id = 1

    def AddCamera(self, camera):
        self.cameras[camera.camera_id] = camera

    def AddImage(self, image):
        self.images[image.image_id] = image

    def AddPoint3D(self, xyz, track_elements):
        pid = self._next_point3D_id
        self._next_point3D_id += 1
        p = Point3D(xyz)
        for image_id, idx in track_elements:
            p.track.AddElement(image_id, idx)
            self.images[image_id].points2D[idx].point3D_id = pid
        self.points3D[pid] = p
        return pid

    def AddPoint3DWithId(self, point3D_id, xyz, track_elements, color=(0, 0, 0), error=-1.0):
        """A point read from a model file (src/base/reconstruction.cc:1640-1680): its id and track come from the file; the
        images' Point2D::point3D_id stay as the image file stored them."""
        p = Point3D(xyz, color, error)
        for image_id, idx in track_elements:
            p.track.AddElement(image_id, idx)
        self.points3D[point3D_id] = p
        self._next_point3D_id = max(self._next_point3D_id, point3D_id + 1)
        return point3D_id

    def Camera(self, camera_id):
        return self.cameras[camera_id]

    def Image(self, image_id):
        return self.images[image_id]

    def Point3D(self, point3D_id):
        return self.points3D[point3D_id]


class SolverOptionsMirror:
    """The ceres::Solver::Options fields the reference sets (bundle_adjustment.h:80-93)."""

    def __init__(self):
        self.function_tolerance = 0.0
        self.gradient_tolerance = 0.0
        self.parameter_tolerance = 0.0
        self.minimizer_progress_to_stdout = False
        self.max_num_iterations = 100
        self.max_linear_solver_iterations = 200
        self.max_num_consecutive_invalid_steps = 10
        self.num_threads = -1  # meaningful for the CPU baseline only
        self.callbacks = []
        self.update_state_every_iteration = False
        self.linear_solver_type = None  # None: reference policy (bundle_adjustment.cc:276-286)


class BundleAdjustmentOptions:
    TRIVIAL, SOFT_L1, CAUCHY = capi.LOSS_TRIVIAL, capi.LOSS_SOFT_L1, capi.LOSS_CAUCHY

    def __init__(self):
        self.loss_function_type = self.TRIVIAL
        self.loss_function_scale = 1.0
        self.refine_focal_length = True
        self.refine_principal_point = False
        self.refine_extra_params = True
        self.refine_extrinsics = True
        self.print_summary = False
        self.min_num_residuals_for_multi_threading = 50000
        self.solver_options = SolverOptionsMirror()

    def Check(self):
        if not self.loss_function_scale >= 0:
            raise ValueError("loss_function_scale must be >= 0")  # CHECK_OPTION_GE, bundle_adjustment.cc:71
        return True


class BundleAdjustmentConfig:
    def __init__(self):
        self.constant_camera_ids, self.image_ids = set(), OrderedDict()
        self.variable_point3D_ids, self.constant_point3D_ids = OrderedDict(), OrderedDict()
        self.constant_poses, self.constant_tvecs = set(), {}

    def NumImages(self): return len(self.image_ids)
    def NumPoints(self): return len(self.variable_point3D_ids) + len(self.constant_point3D_ids)
    def NumConstantCameras(self): return len(self.constant_camera_ids)
    def NumConstantPoses(self): return len(self.constant_poses)
    def NumConstantTvecs(self): return len(self.constant_tvecs)
    def NumVariablePoints(self): return len(self.variable_point3D_ids)
    def NumConstantPoints(self): return len(self.constant_point3D_ids)

    def NumResiduals(self, reconstruction):  # bundle_adjustment.cc:103-139
        num_obs = 0
        for image_id in self.image_ids:
            num_obs += sum(1 for p in reconstruction.Image(image_id).Points2D() if p.HasPoint3D())
        for ids in (self.variable_point3D_ids, self.constant_point3D_ids):
            for pid in ids:
                for el in reconstruction.Point3D(pid).Track().Elements():
                    if el.image_id not in self.image_ids:
                        num_obs += 1
        return 2 * num_obs

    def AddImage(self, image_id): self.image_ids[image_id] = True
    def HasImage(self, image_id): return image_id in self.image_ids
    def RemoveImage(self, image_id): self.image_ids.pop(image_id, None)
    def SetConstantCamera(self, camera_id): self.constant_camera_ids.add(camera_id)
    def SetVariableCamera(self, camera_id): self.constant_camera_ids.discard(camera_id)
    def IsConstantCamera(self, camera_id): return camera_id in self.constant_camera_ids

    def SetConstantPose(self, image_id):
        assert self.HasImage(image_id) and not self.HasConstantTvec(image_id)
        self.constant_poses.add(image_id)

    def SetVariablePose(self, image_id): self.constant_poses.discard(image_id)
    def HasConstantPose(self, image_id): return image_id in self.constant_poses

    def SetConstantTvec(self, image_id, idxs):
        assert len(idxs) > 0 and len(idxs) <= 3 and self.HasImage(image_id) and not self.HasConstantPose(image_id)
        assert len(set(idxs)) == len(idxs), "Tvec indices must not contain duplicates"
        self.constant_tvecs[image_id] = list(idxs)

    def RemoveConstantTvec(self, image_id): self.constant_tvecs.pop(image_id, None)
    def HasConstantTvec(self, image_id): return image_id in self.constant_tvecs

    def AddVariablePoint(self, pid):
        assert not self.HasConstantPoint(pid)
        self.variable_point3D_ids[pid] = True

    def AddConstantPoint(self, pid):
        assert not self.HasVariablePoint(pid)
        self.constant_point3D_ids[pid] = True

    def HasPoint(self, pid): return self.HasVariablePoint(pid) or self.HasConstantPoint(pid)
    def HasVariablePoint(self, pid): return pid in self.variable_point3D_ids
    def HasConstantPoint(self, pid): return pid in self.constant_point3D_ids
    def RemoveVariablePoint(self, pid): self.variable_point3D_ids.pop(pid, None)
    def RemoveConstantPoint(self, pid): self.constant_point3D_ids.pop(pid, None)
    def Images(self): return list(self.image_ids)
    def VariablePoints(self): return list(self.variable_point3D_ids)
    def ConstantPoints(self): return list(self.constant_point3D_ids)
    def ConstantTvec(self, image_id): return self.constant_tvecs[image_id]


class BundleAdjuster:
    """BundleAdjuster(options, config).Solve(reconstruction) -> bool; .Summary()."""

    def __init__(self, options, config):
        import copy
        self.options_ = copy.deepcopy(options)  # copied by value, bundle_adjustment.cc:252-256
        self.config_ = copy.deepcopy(config)
        self.options_.Check()
        self.summary_ = None
        self.trace_ = []
        self._used = False
        self.flat_ = None

    # -- SetUp: bundle_adjustment.cc:326-530 ------------------------------------------------
    def Flatten(self, reconstruction):
        """Residual blocks in the reference order -> FlatProblem plus the id maps."""
        opt, cfg = self.options_, self.config_
        img_index, cam_index, pt_index = OrderedDict(), OrderedDict(), OrderedDict()
        obs_img, obs_pt, obs_xy = [], [], []
        pose_const, tvec_mask = [], []
        point3D_num_observations = OrderedDict()
        camera_ids = OrderedDict()

        def image_slot(image_id, constant):
            if image_id not in img_index:
                img_index[image_id] = len(img_index)
                pose_const.append(1 if constant else 0)
                tvec_mask.append(0)
            return img_index[image_id]

        def camera_slot(camera_id):
            if camera_id not in cam_index:
                cam_index[camera_id] = len(cam_index)
            return cam_index[camera_id]

        def point_slot(pid):
            if pid not in pt_index:
                pt_index[pid] = len(pt_index)
            return pt_index[pid]

        for image_id in cfg.Images():  # AddImageToProblem, :348-427
            image = reconstruction.Image(image_id)
            image.NormalizeQvec()
            constant_pose = (not opt.refine_extrinsics) or cfg.HasConstantPose(image_id)
            n_obs = 0
            for p2 in image.Points2D():
                if not p2.HasPoint3D():
                    continue
                n_obs += 1
                pid = p2.Point3DId()
                point3D_num_observations[pid] = point3D_num_observations.get(pid, 0) + 1
                obs_img.append(image_slot(image_id, constant_pose))
                obs_pt.append(point_slot(pid))
                obs_xy.append(p2.XY())
            if n_obs > 0:
                camera_ids[image.CameraId()] = True
                camera_slot(image.CameraId())
                if not constant_pose and cfg.HasConstantTvec(image_id):
                    tvec_mask[img_index[image_id]] = sum(1 << k for k in cfg.ConstantTvec(image_id))

        for pid in cfg.VariablePoints() + cfg.ConstantPoints():  # AddPointToProblem, :429-478
            p3 = reconstruction.Point3D(pid)
            point3D_num_observations.setdefault(pid, 0)
            if point3D_num_observations[pid] == p3.Track().Length():
                continue
            for el in p3.Track().Elements():
                if cfg.HasImage(el.image_id):
                    continue
                point3D_num_observations[pid] += 1
                image = reconstruction.Image(el.image_id)
                if image.CameraId() not in camera_ids:
                    camera_ids[image.CameraId()] = True
                    cfg.SetConstantCamera(image.CameraId())
                camera_slot(image.CameraId())
                obs_img.append(image_slot(el.image_id, True))
                obs_pt.append(point_slot(pid))
                obs_xy.append(image.Point2D(el.point2D_idx).XY())

        # ParameterizeCameras, :480-516
        constant_camera = not (opt.refine_focal_length or opt.refine_principal_point or opt.refine_extra_params)
        cam_mask, cam_model, cam_params = [], [], []
        for camera_id in cam_index:
            cam = reconstruction.Camera(camera_id)
            full = (1 << cam.NumParams()) - 1
            if constant_camera or cfg.IsConstantCamera(camera_id):
                m = full
            else:
                m = 0
                if not opt.refine_focal_length:
                    m |= sum(1 << j for j in cam.FocalLengthIdxs())
                if not opt.refine_principal_point:
                    m |= sum(1 << j for j in cam.PrincipalPointIdxs())
                if not opt.refine_extra_params:
                    m |= sum(1 << j for j in cam.ExtraParamsIdxs())
            cam_mask.append(m)
            cam_model.append(cam.ModelId())
            prm = np.zeros(capi.MAX_CAMERA_PARAMS); prm[:cam.NumParams()] = cam.Params()
            cam_params.append(prm)

        # ParameterizePoints, :518-530
        pt_const = np.zeros(len(pt_index), np.uint8)
        for pid, n in point3D_num_observations.items():
            if pid in pt_index and reconstruction.Point3D(pid).Track().Length() > n:
                pt_const[pt_index[pid]] = 1
        for pid in cfg.ConstantPoints():
            if pid in pt_index:
                pt_const[pt_index[pid]] = 1

        n_img = len(img_index)
        q = np.zeros((n_img, 4)); t = np.zeros((n_img, 3)); icam = np.zeros(n_img, np.int32)
        for image_id, i in img_index.items():
            im = reconstruction.Image(image_id)
            q[i], t[i], icam[i] = im.Qvec(), im.Tvec(), cam_index[im.CameraId()]
        X = np.zeros((len(pt_index), 3))
        for pid, j in pt_index.items():
            X[j] = reconstruction.Point3D(pid).XYZ()
        flat = FlatProblem(q, t, icam, np.asarray(cam_model, np.int32),
                           np.asarray(cam_params).reshape(len(cam_index), capi.MAX_CAMERA_PARAMS), X,
                           np.asarray(obs_img, np.int32), np.asarray(obs_pt, np.int32),
                           np.asarray(obs_xy, np.float64).reshape(-1, 2),
                           image_pose_constant=np.asarray(pose_const, np.uint8),
                           image_tvec_constant_mask=np.asarray(tvec_mask, np.uint8),
                           camera_param_constant_mask=np.asarray(cam_mask, np.uint16), point_constant=pt_const)
        self.flat_, self.maps_ = flat, (img_index, cam_index, pt_index)
        return flat

    def SolverOptions(self):
        so = self.options_.solver_options
        o = capi.default_solver_options()
        o.linear_solver_type = capi.SOLVER_AUTO if so.linear_solver_type is None else so.linear_solver_type
        o.num_config_images = self.config_.NumImages()
        o.max_num_iterations = so.max_num_iterations
        o.max_linear_solver_iterations = so.max_linear_solver_iterations
        o.max_num_consecutive_invalid_steps = so.max_num_consecutive_invalid_steps
        o.function_tolerance, o.gradient_tolerance = so.function_tolerance, so.gradient_tolerance
        o.parameter_tolerance = so.parameter_tolerance
        o.loss_function_type = self.options_.loss_function_type
        o.loss_function_scale = self.options_.loss_function_scale
        o.update_state_every_iteration = 1 if so.update_state_every_iteration else 0
        return o

    def Solve(self, reconstruction):
        if reconstruction is None:
            raise ValueError("reconstruction must not be None")  # CHECK_NOTNULL, :259
        if self._used:
            raise RuntimeError("Cannot use the same BundleAdjuster multiple times")  # CHECK, :260
        self._used = True
        flat = self.Flatten(reconstruction)
        if flat.num_observations == 0:
            return False  # :267-269
        solver = Solver(flat)
        try:
            callbacks = list(self.options_.solver_options.callbacks)

            def cb(it):
                if self.options_.solver_options.update_state_every_iteration and callbacks:
                    self._scatter(reconstruction, solver)
                for c in callbacks:
                    r = c(it)
                    if r:
                        return r
                return 0
            self.summary_ = solver.solve(self.SolverOptions(), cb if callbacks else None)
            self.trace_ = solver.trace()
            self._scatter(reconstruction, solver)
        finally:
            solver.close()
        if self.options_.print_summary:
            PrintSolverSummary(self.summary_)
        return True

    def _scatter(self, reconstruction, solver):
        q, t, cam, X = solver.parameters()
        img_index, cam_index, pt_index = self.maps_
        for image_id, i in img_index.items():
            im = reconstruction.Image(image_id)
            im.qvec[:] = q[i]; im.tvec[:] = t[i]
        for camera_id, c in cam_index.items():
            cm = reconstruction.Camera(camera_id)
            cm.params[:] = cam[c, :cm.NumParams()]
        for pid, j in pt_index.items():
            reconstruction.Point3D(pid).xyz[:] = X[j]

    def Summary(self):
        return self.summary_


def PrintSolverSummary(s):  # bundle_adjustment.cc:1142-1196
    term = {0: "Convergence", 1: "No convergence", 2: "Failure", 3: "User success", 4: "User failure"}
    print("    Residuals : %d" % s.num_residuals_reduced)
    print("   Parameters : %d" % s.num_effective_parameters_reduced)
    print("   Iterations : %d" % (s.num_successful_steps + s.num_unsuccessful_steps))
    print("         Time : %g [s]" % s.total_time_in_seconds)
    print(" Initial cost : %g [px]" % np.sqrt(s.initial_cost / max(s.num_residuals_reduced, 1)))
    print("   Final cost : %g [px]" % np.sqrt(s.final_cost / max(s.num_residuals_reduced, 1)))
    print("  Termination : %s" % term.get(s.termination_type, "Unknown"))
