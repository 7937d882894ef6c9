"""`RigBundleAdjuster` mirror over `sba_rig_solve`.

  RigBundleAdjuster            /root/reference/src/optim/bundle_adjustment.cc:789-1140, options .h:272-282
  CameraRig (what BA needs)    /root/reference/src/base/camera_rig.h:27-107, camera_rig.cc:86-110, 230-258
  pose helpers                 /root/reference/src/base/pose.cc:82-143, 183-196

Host side only: the rig poses are initialised by averaging (ComputeAbsolutePose), observations are gated by
max_reproj_error against the concatenated pose (:970-975) -- the per-observation error comes from the device pass
`sba_observation_errors` -- the residual blocks are flattened in the reference's order and `sba_rig_solve` runs the LM
solve on the device (dense small-problem path, see csrc/rig_solve.cu).  TearDown writes the concatenated poses back
into the images (:889-903).  Nothing here computes residuals or steps on the CPU.
"""
import copy
import ctypes as C
from collections import OrderedDict

import numpy as np

from . import _capi as capi
from .bundle_adjustment import PrintSolverSummary
from .filters import device_observation_errors
from .problem import FlatProblem


# ---- pose.cc helpers (host bookkeeping of a handful of poses) ---------------------------------------------------------
def NormalizeQuaternion(q):
    q = np.asarray(q, dtype=np.float64)
    n = np.linalg.norm(q)
    return np.array([1.0, 0.0, 0.0, 0.0]) if n == 0 else q / n


def _quat_mul(a, b):  # Eigen quaternion product a * b, (w, x, y, z)
    return np.array([a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3], a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2],
                     a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1], a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0]])


def InvertQuaternion(q):
    return np.array([q[0], -q[1], -q[2], -q[3]])


def ConcatenateQuaternions(q1, q2):  # pose.cc:97-108: quat2 * quat1
    return NormalizeQuaternion(_quat_mul(NormalizeQuaternion(q2), NormalizeQuaternion(q1)))


def QuaternionRotatePoint(q, p):  # pose.cc:110-116
    w, x, y, z = NormalizeQuaternion(q)
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y)],
                  [2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x)],
                  [2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)]])
    return R @ np.asarray(p, dtype=np.float64)


def ConcatenatePoses(q1, t1, q2, t2):  # pose.cc:183-190
    return ConcatenateQuaternions(q1, q2), np.asarray(t2, dtype=np.float64) + QuaternionRotatePoint(q2, t1)


def InvertPose(q, t):  # pose.cc:192-196
    iq = InvertQuaternion(q)
    return iq, -QuaternionRotatePoint(iq, t)


def AverageQuaternions(qvecs, weights):  # pose.cc:118-143: eigenvector of the largest eigenvalue of sum w q q^T
    if len(qvecs) == 1:
        return np.asarray(qvecs[0], dtype=np.float64).copy()
    A = np.zeros((4, 4))
    for q, w in zip(qvecs, weights):
        q = NormalizeQuaternion(q)
        A += w * np.outer(q, q)
    A /= float(sum(weights))
    _, vec = np.linalg.eigh(A)  # ascending eigenvalues, as Eigen::SelfAdjointEigenSolver
    q = vec[:, 3].copy()
    # an eigenvector's sign is the solver's choice: Eigen's returns the representative the reference's own test expects
    # (pose_test.cc:188-226: w > 0 in every case); LAPACK's may return the other one -- the same rotation
    k = int(np.flatnonzero(np.abs(q) > 0)[0])
    return -q if q[k] < 0 else q


class CameraRig:
    """The part of CameraRig the bundle adjuster uses."""

    def __init__(self):
        self.ref_camera_id = None
        self.rig_cameras = OrderedDict()  # camera_id -> [rel_qvec, rel_tvec]
        self.snapshots = []

    def AddCamera(self, camera_id, rel_qvec, rel_tvec):
        assert camera_id not in self.rig_cameras
        self.rig_cameras[camera_id] = [np.asarray(rel_qvec, dtype=np.float64).copy(), np.asarray(rel_tvec, dtype=np.float64).copy()]

    def AddSnapshot(self, image_ids):
        assert len(image_ids) > 0 and len(image_ids) <= len(self.rig_cameras)
        self.snapshots.append(list(image_ids))

    def SetRefCameraId(self, camera_id):
        assert camera_id in self.rig_cameras
        self.ref_camera_id = camera_id

    def RefCameraId(self): return self.ref_camera_id
    def HasCamera(self, camera_id): return camera_id in self.rig_cameras
    def GetCameraIds(self): return list(self.rig_cameras)
    def Snapshots(self): return self.snapshots
    def NumSnapshots(self): return len(self.snapshots)
    def NumCameras(self): return len(self.rig_cameras)
    def RelativeQvec(self, camera_id): return self.rig_cameras[camera_id][0]
    def RelativeTvec(self, camera_id): return self.rig_cameras[camera_id][1]

    def Check(self, reconstruction):  # camera_rig.cc:86-110
        if not self.HasCamera(self.ref_camera_id):
            raise ValueError("reference camera is not part of the rig")
        seen = set()
        for cid in self.rig_cameras:
            reconstruction.Camera(cid)
        for snap in self.snapshots:
            if not snap or len(snap) > self.NumCameras():
                raise ValueError("invalid snapshot size")
            has_ref = False
            for image_id in snap:
                if image_id in seen:
                    raise ValueError("image in several snapshots")
                seen.add(image_id)
                cam = reconstruction.Image(image_id).CameraId()
                if not self.HasCamera(cam):
                    raise ValueError("snapshot image's camera is not part of the rig")
                has_ref = has_ref or cam == self.ref_camera_id
            if not has_ref:
                raise ValueError("snapshot without the reference camera")

    def ComputeAbsolutePose(self, snapshot_idx, reconstruction):  # camera_rig.cc:230-258
        qs, t_sum = [], np.zeros(3)
        snap = self.snapshots[snapshot_idx]
        for image_id in snap:
            im = reconstruction.Image(image_id)
            inv_q, _ = InvertPose(self.RelativeQvec(im.CameraId()), self.RelativeTvec(im.CameraId()))
            qs.append(ConcatenateQuaternions(im.Qvec(), inv_q))
            t_sum += QuaternionRotatePoint(inv_q, np.asarray(im.Tvec()) - self.RelativeTvec(im.CameraId()))
        return AverageQuaternions(qs, [1.0] * len(snap)), t_sum / len(snap)


class RigBundleAdjusterOptions:  # bundle_adjustment.h:272-282
    def __init__(self, refine_relative_poses=True, max_reproj_error=1000.0):
        self.refine_relative_poses, self.max_reproj_error = refine_relative_poses, max_reproj_error


class RigBundleAdjuster:
    """RigBundleAdjuster(options, rig_options, config).Solve(reconstruction, camera_rigs) -> bool; .Summary()."""

    def __init__(self, options, rig_options, config):
        self.options_ = copy.deepcopy(options)
        self.rig_options_ = copy.deepcopy(rig_options)
        self.config_ = copy.deepcopy(config)
        self.options_.Check()
        self.summary_, self.trace_, self._used = None, [], False

    # ---- SetUp, :869-887 ----
    def Flatten(self, reconstruction, camera_rigs, errors_fn=device_observation_errors):
        opt, cfg, ro = self.options_, self.config_, self.rig_options_
        rig_cam_ids, image_rig = set(), {}
        for rig in camera_rigs:  # :800-817
            rig.Check(reconstruction)
            for cid in rig.GetCameraIds():
                if cid in rig_cam_ids:
                    raise ValueError("Camera must not be part of multiple camera rigs")
                rig_cam_ids.add(cid)
            for snap in rig.Snapshots():
                for image_id in snap:
                    if image_id in image_rig:
                        raise ValueError("Image must not be part of multiple camera rigs")
                    image_rig[image_id] = rig
        # pose blocks: slot 0 of the rig list is the constant identity ordinary images hang on
        rig_q, rig_t, rig_const = [np.array([1.0, 0.0, 0.0, 0.0])], [np.zeros(3)], [1]
        image_rig_slot = {}
        for rig in camera_rigs:  # ComputeCameraRigPoses, :1104-1129
            for s in range(rig.NumSnapshots()):
                q, t = rig.ComputeAbsolutePose(s, reconstruction)
                rig_q.append(q); rig_t.append(t); rig_const.append(0)
                for image_id in rig.Snapshots()[s]:
                    image_rig_slot[image_id] = len(rig_q) - 1
        rel_index, rel_q, rel_t, rel_const, rel_mask = OrderedDict(), [], [], [], []   # key: ("rig", id(rig), camera_id) | ("img", image_id)
        cam_index, pt_index = OrderedDict(), OrderedDict()
        b_rig, b_rel, b_cam, b_pt, b_xy = [], [], [], [], []
        num_obs_of_point, camera_ids = OrderedDict(), OrderedDict()

        def slot(d, key):
            if key not in d:
                d[key] = len(d)
            return d[key]

        def rel_slot(key, q, t, const, mask=0):
            if key not in rel_index:
                rel_index[key] = len(rel_index)
                rel_q.append(np.asarray(q, dtype=np.float64).copy()); rel_t.append(np.asarray(t, dtype=np.float64).copy())
                rel_const.append(1 if const else 0); rel_mask.append(mask)
            return rel_index[key]

        # the outlier gate of rig images (:943-975): squared reprojection error under the concatenated pose, one device pass
        gate = {}
        rig_images = [i for i in cfg.Images() if i in image_rig]
        if rig_images:
            gq, gt, gc, gm, gp, go_i, go_p, go_xy, gkeys, gX = [], [], [], [], [], [], [], [], [], []
            cam_l, pt_l = OrderedDict(), OrderedDict()
            for image_id in rig_images:
                im, rig = reconstruction.Image(image_id), image_rig[image_id]
                s = image_rig_slot[image_id]
                q, t = ConcatenatePoses(rig_q[s], rig_t[s], rig.RelativeQvec(im.CameraId()), rig.RelativeTvec(im.CameraId()))
                gq.append(q); gt.append(t); gc.append(slot(cam_l, im.CameraId()))
                for idx, p2 in enumerate(im.Points2D()):
                    if p2.HasPoint3D():
                        go_i.append(len(gq) - 1); go_p.append(slot(pt_l, p2.Point3DId())); go_xy.append(p2.XY()); gkeys.append((image_id, idx))
            if gkeys:
                params = np.zeros((len(cam_l), capi.MAX_CAMERA_PARAMS))
                for cid, c in cam_l.items():
                    params[c, :reconstruction.Camera(cid).NumParams()] = reconstruction.Camera(cid).Params()
                prob = FlatProblem(np.array(gq), np.array(gt), gc, [reconstruction.Camera(c).ModelId() for c in cam_l], params,
                                   np.array([reconstruction.Point3D(p).XYZ() for p in pt_l]), go_i, go_p, np.array(go_xy))
                err, _ = errors_fn(prob)
                gate = {k: e for k, e in zip(gkeys, err)}
        max_sq = ro.max_reproj_error * ro.max_reproj_error

        for image_id in cfg.Images():  # AddImageToProblem, :905-1055
            im = reconstruction.Image(image_id)
            constant_pose, constant_tvec = cfg.HasConstantPose(image_id), cfg.HasConstantTvec(image_id)
            rig = image_rig.get(image_id)
            if rig is not None:
                if constant_pose or constant_tvec:
                    raise ValueError("Images contained in a camera rig must not have constant pose / tvec")  # CHECK, :929-932
            else:
                im.NormalizeQvec()
            n_obs = 0
            for idx, p2 in enumerate(im.Points2D()):
                if not p2.HasPoint3D():
                    continue
                if rig is not None and gate[(image_id, idx)] > max_sq:
                    continue
                n_obs += 1
                pid = p2.Point3DId()
                num_obs_of_point[pid] = num_obs_of_point.get(pid, 0) + 1
                if rig is not None:
                    konst = (not ro.refine_relative_poses) or im.CameraId() == rig.RefCameraId()  # :1040-1046
                    r = rel_slot(("rig", id(rig), im.CameraId()), rig.RelativeQvec(im.CameraId()), rig.RelativeTvec(im.CameraId()), konst)
                    b_rig.append(image_rig_slot[image_id])
                else:
                    mask = sum(1 << k for k in cfg.ConstantTvec(image_id)) if (not constant_pose and constant_tvec) else 0
                    r = rel_slot(("img", image_id), im.Qvec(), im.Tvec(), constant_pose, mask)
                    b_rig.append(0)
                b_rel.append(r); b_cam.append(slot(cam_index, im.CameraId())); b_pt.append(slot(pt_index, pid)); b_xy.append(p2.XY())
            camera_ids[im.CameraId()] = True  # (:961: collected whether or not an observation was added)

        for pid in cfg.VariablePoints() + cfg.ConstantPoints():  # AddPointToProblem, :1057-1102
            p3 = reconstruction.Point3D(pid)
            num_obs_of_point.setdefault(pid, 0)
            if num_obs_of_point[pid] == p3.Track().Length():
                continue
            for el in p3.Track().Elements():
                if cfg.HasImage(el.image_id):
                    continue
                num_obs_of_point[pid] += 1
                im = reconstruction.Image(el.image_id)
                if im.CameraId() not in camera_ids:
                    camera_ids[im.CameraId()] = True
                    cfg.SetConstantCamera(im.CameraId())
                r = rel_slot(("img", el.image_id), im.Qvec(), im.Tvec(), True)
                b_rig.append(0); b_rel.append(r); b_cam.append(slot(cam_index, im.CameraId())); b_pt.append(slot(pt_index, pid))
                b_xy.append(im.Point2D(el.point2D_idx).XY())

        # ParameterizeCameras / ParameterizePoints as BundleAdjuster (:480-530)
        constant_camera = not (opt.refine_focal_length or opt.refine_principal_point or opt.refine_extra_params)
        cam_mask, cam_model, cam_params = [], [], []
        for camera_id in cam_index:
            cam = reconstruction.Camera(camera_id)
            if constant_camera or cfg.IsConstantCamera(camera_id):
                m = (1 << cam.NumParams()) - 1
            else:
                m = 0
                if not opt.refine_focal_length:
                    m |= sum(1 << j for j in cam.FocalLengthIdxs())
                if not opt.refine_principal_point:
                    m |= sum(1 << j for j in cam.PrincipalPointIdxs())
                if not opt.refine_extra_params:
                    m |= sum(1 << j for j in cam.ExtraParamsIdxs())
            cam_mask.append(m); cam_model.append(cam.ModelId())
            prm = np.zeros(capi.MAX_CAMERA_PARAMS); prm[:cam.NumParams()] = cam.Params()
            cam_params.append(prm)
        pt_const = np.zeros(len(pt_index), np.uint8)
        for pid, n in num_obs_of_point.items():
            if pid in pt_index and reconstruction.Point3D(pid).Track().Length() > n:
                pt_const[pt_index[pid]] = 1
        for pid in cfg.ConstantPoints():
            if pid in pt_index:
                pt_const[pt_index[pid]] = 1

        f = dict(rig_q=np.array(rig_q).reshape(-1, 4), rig_t=np.array(rig_t).reshape(-1, 3), rig_const=np.array(rig_const, np.uint8),
                 rel_q=np.array(rel_q).reshape(-1, 4), rel_t=np.array(rel_t).reshape(-1, 3), rel_const=np.array(rel_const, np.uint8),
                 rel_mask=np.array(rel_mask, np.uint8), cam_model=np.array(cam_model, np.int32),
                 cam_params=np.array(cam_params).reshape(-1, capi.MAX_CAMERA_PARAMS), cam_mask=np.array(cam_mask, np.uint16),
                 points=np.array([reconstruction.Point3D(p).XYZ() for p in pt_index]).reshape(-1, 3), pt_const=pt_const,
                 b_rig=np.array(b_rig, np.int32), b_rel=np.array(b_rel, np.int32), b_cam=np.array(b_cam, np.int32),
                 b_pt=np.array(b_pt, np.int32), b_xy=np.array(b_xy, np.float64).reshape(-1, 2))
        self.flat_, self.maps_ = f, (rel_index, cam_index, pt_index, image_rig, image_rig_slot)
        return f

    @staticmethod
    def desc(f):
        p = capi.RigProblem()
        p.num_rig_poses, p.rig_qvec, p.rig_tvec = len(f["rig_q"]), capi.ptr(f["rig_q"], capi._dp), capi.ptr(f["rig_t"], capi._dp)
        p.rig_pose_constant = capi.ptr(f["rig_const"], capi._u8p)
        p.num_rel_poses, p.rel_qvec, p.rel_tvec = len(f["rel_q"]), capi.ptr(f["rel_q"], capi._dp), capi.ptr(f["rel_t"], capi._dp)
        p.rel_pose_constant, p.rel_tvec_constant_mask = capi.ptr(f["rel_const"], capi._u8p), capi.ptr(f["rel_mask"], capi._u8p)
        p.num_cameras, p.camera_model = len(f["cam_model"]), capi.ptr(f["cam_model"], capi._ip)
        p.camera_params, p.camera_param_constant_mask = capi.ptr(f["cam_params"], capi._dp), capi.ptr(f["cam_mask"], capi._u16p)
        p.num_points, p.points, p.point_constant = len(f["points"]), capi.ptr(f["points"], capi._dp), capi.ptr(f["pt_const"], capi._u8p)
        p.num_blocks = len(f["b_rig"])
        p.block_rig, p.block_rel = capi.ptr(f["b_rig"], capi._ip), capi.ptr(f["b_rel"], capi._ip)
        p.block_camera, p.block_point = capi.ptr(f["b_cam"], capi._ip), capi.ptr(f["b_pt"], capi._ip)
        p.block_xy = capi.ptr(f["b_xy"], capi._dp)
        return p

    def SolverOptions(self):
        so = self.options_.solver_options
        o = capi.default_solver_options()
        o.linear_solver_type = capi.SOLVER_AUTO if so.linear_solver_type is None else so.linear_solver_type
        o.num_config_images = self.config_.NumImages()
        o.max_num_iterations = so.max_num_iterations
        o.max_num_consecutive_invalid_steps = so.max_num_consecutive_invalid_steps
        o.function_tolerance, o.gradient_tolerance, o.parameter_tolerance = so.function_tolerance, so.gradient_tolerance, so.parameter_tolerance
        o.loss_function_type, o.loss_function_scale = self.options_.loss_function_type, self.options_.loss_function_scale
        return o

    def Solve(self, reconstruction, camera_rigs, cap=512):
        if reconstruction is None or camera_rigs is None:
            raise ValueError("reconstruction and camera_rigs must not be None")  # CHECK_NOTNULL, :796-797
        if self._used:
            raise RuntimeError("Cannot use the same BundleAdjuster multiple times")
        self._used = True
        f = self.Flatten(reconstruction, camera_rigs)
        if len(f["b_rig"]) == 0:
            return False  # :824-826
        s = capi.Summary()
        rows = (capi.IterationSummary * cap)()
        n = C.c_int32()
        d = self.desc(f)
        o = self.SolverOptions()
        capi.check(capi.load().sba_rig_solve(C.byref(d), C.byref(o), C.cast(None, capi.ITERATION_CALLBACK), None, C.byref(s), rows, cap, C.byref(n)))
        self.summary_, self.trace_ = s, [rows[i] for i in range(n.value)]
        self.Scatter(reconstruction, f)
        if self.options_.print_summary:
            PrintSolverSummary(s)
        return True

    def Scatter(self, reconstruction, f):
        rel_index, cam_index, pt_index, image_rig, image_rig_slot = self.maps_
        for key, r in rel_index.items():
            if key[0] == "img":
                im = reconstruction.Image(key[1])
                im.qvec[:] = f["rel_q"][r]; im.tvec[:] = f["rel_t"][r]
        for rig in {id(r): r for r in image_rig.values()}.values():
            for cid in rig.GetCameraIds():
                key = ("rig", id(rig), cid)
                if key in rel_index:
                    rig.RelativeQvec(cid)[:] = f["rel_q"][rel_index[key]]; rig.RelativeTvec(cid)[:] = f["rel_t"][rel_index[key]]
        for camera_id, c in cam_index.items():
            cm = reconstruction.Camera(camera_id)
            cm.params[:] = f["cam_params"][c, :cm.NumParams()]
        for pid, j in pt_index.items():
            reconstruction.Point3D(pid).xyz[:] = f["points"][j]
        for image_id, rig in image_rig.items():  # TearDown, :889-903
            im = reconstruction.Image(image_id)
            s = image_rig_slot[image_id]
            q, t = ConcatenatePoses(f["rig_q"][s], f["rig_t"][s], rig.RelativeQvec(im.CameraId()), rig.RelativeTvec(im.CameraId()))
            im.qvec[:] = q; im.tvec[:] = t

    def Summary(self):
        return self.summary_
