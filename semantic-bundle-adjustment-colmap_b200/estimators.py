"""`RefineAbsolutePose` on the device (/root/reference/src/estimators/pose.cc:198-322,
options /root/reference/src/estimators/pose.h:81-105; called by the mapper when it registers an image,
/root/reference/src/sfm/incremental_mapper.cc:502).

The reference builds a one-image Ceres problem from the inlier 2D-3D correspondences -- the bundle-adjustment functor
(cost_functions.h:44-81) with every 3-D point constant, Cauchy loss, quaternion manifold, principal point always
fixed, focal length / extra parameters optional -- and solves it with DENSE_QR.  That is exactly an `sba_problem_desc`
with one image, one camera and constant points, so the existing device solve runs it (with no point blocks to
eliminate, DENSE_SCHUR factors the same normal equations DENSE_QR solves through J = QR; the step is the same to rounding).
The covariance output of the reference (ceres::Covariance) is not provided.  Nothing here computes residuals on the CPU.
"""
import numpy as np

from . import _capi as capi
from .problem import FlatProblem, Solver


class AbsolutePoseRefinementOptions:
    """pose.h:81-105"""

    def __init__(self, gradient_tolerance=1.0, max_num_iterations=100, loss_function_scale=1.0, refine_focal_length=True,
                 refine_extra_params=True, print_summary=True):
        self.gradient_tolerance, self.max_num_iterations, self.loss_function_scale = gradient_tolerance, max_num_iterations, loss_function_scale
        self.refine_focal_length, self.refine_extra_params, self.print_summary = refine_focal_length, refine_extra_params, print_summary

    def Check(self):
        if self.gradient_tolerance < 0.0 or self.max_num_iterations < 0 or self.loss_function_scale < 0.0:
            raise ValueError("AbsolutePoseRefinementOptions::Check failed")


def solver_options(options):
    """ceres::Solver::Options defaults with the three fields pose.cc:283-286 sets (DENSE_QR -> the dense direct path)."""
    o = capi.default_solver_options()
    o.function_tolerance, o.parameter_tolerance = 1e-6, 1e-8          # Ceres 2.1.0 defaults (the BA options override them, pose.cc does not)
    o.gradient_tolerance = options.gradient_tolerance
    o.max_num_iterations = options.max_num_iterations
    o.linear_solver_type = capi.DENSE_SCHUR
    o.loss_function_type, o.loss_function_scale = capi.LOSS_CAUCHY, options.loss_function_scale
    return o


def flatten(options, inlier_mask, points2D, points3D, qvec, tvec, camera):
    """The one-image problem in the reference's residual order (inlier correspondences in input order, pose.cc:219-243)."""
    inlier_mask = np.asarray(inlier_mask).astype(bool)
    points2D = np.asarray(points2D, dtype=np.float64).reshape(-1, 2)
    points3D = np.asarray(points3D, dtype=np.float64).reshape(-1, 3)
    if not (len(inlier_mask) == len(points2D) == len(points3D)):
        raise ValueError("inlier_mask, points2D and points3D differ in length")  # CHECK_EQ, pose.cc:204-205
    idx = np.nonzero(inlier_mask)[0]
    n = len(idx)
    # constant camera parameters: principal point always; focal length / extra parameters by option (pose.cc:249-281)
    const = list(camera.PrincipalPointIdxs())
    if not options.refine_focal_length:
        const += list(camera.FocalLengthIdxs())
    if not options.refine_extra_params:
        const += list(camera.ExtraParamsIdxs())
    num_params = camera.NumParams()
    if (not options.refine_focal_length and not options.refine_extra_params) or len(set(const)) == num_params:
        mask = (1 << num_params) - 1
    else:
        mask = 0
        for k in set(const):
            mask |= 1 << int(k)
    return FlatProblem(image_qvec=np.asarray(qvec, dtype=np.float64).reshape(1, 4), image_tvec=np.asarray(tvec, dtype=np.float64).reshape(1, 3),
                       image_camera=[0], camera_model=[camera.ModelId()], camera_params=np.asarray(camera.Params(), dtype=np.float64).reshape(1, -1),
                       points=points3D[idx].copy(), obs_image=np.zeros(n, np.int32), obs_point=np.arange(n, dtype=np.int32),
                       obs_xy=points2D[idx].copy(), point_constant=np.ones(n, np.uint8),
                       camera_param_constant_mask=np.array([mask], np.uint16))


def RefineAbsolutePose(options, inlier_mask, points2D, points3D, qvec, tvec, camera, summary_out=None):
    """In place on qvec (w, x, y, z), tvec and camera.Params(); returns Summary::IsSolutionUsable() as the reference does.
    summary_out: optional list that receives the sba Summary."""
    options.Check()
    prob = flatten(options, inlier_mask, points2D, points3D, qvec, tvec, camera)
    if prob.num_observations == 0:
        return True  # an empty ceres::Problem "converges" at once: the solution is usable (pose.cc:245, 321)
    s = Solver(prob)
    try:
        summ = s.solve(solver_options(options))
        q, t, cams, _ = s.parameters()
    finally:
        s.close()
    np.asarray(qvec)[...] = q.reshape(-1)[:4]
    np.asarray(tvec)[...] = t.reshape(-1)[:3]
    camera.Params()[:] = cams.reshape(-1)[:camera.NumParams()]
    if summary_out is not None:
        summary_out.append(summ)
    return summ.termination_type in (capi.CONVERGENCE, capi.NO_CONVERGENCE, capi.USER_SUCCESS)
