"""Flattened bundle-adjustment problem (the arrays of `sba_problem_desc`) and a thin
object wrapper over the C-ABI handle.

`FlatProblem` is what the host side produces from a Reconstruction in the reference's
residual order (/root/reference/src/optim/bundle_adjustment.cc:326-478); `Solver` owns one
`sba_handle` and forwards to the CUDA library.  Nothing here computes residuals on the CPU.
"""
import ctypes as C

import numpy as np

from . import _capi as capi


class FlatProblem:
    def __init__(self, image_qvec, image_tvec, image_camera, camera_model, camera_params, points,
                 obs_image, obs_point, obs_xy, image_pose_constant=None, image_tvec_constant_mask=None,
                 camera_param_constant_mask=None, point_constant=None, sem_maps=None, sem_weight=1.0):
        self.image_qvec = capi.as_f64(image_qvec).reshape(-1, 4)
        self.image_tvec = capi.as_f64(image_tvec).reshape(-1, 3)
        self.image_camera = capi.as_i32(image_camera)
        self.camera_model = capi.as_i32(camera_model)
        cp = capi.as_f64(camera_params)
        self.camera_params = np.zeros((len(self.camera_model), capi.MAX_CAMERA_PARAMS))
        if len(self.camera_model):
            cp = cp.reshape(len(self.camera_model), -1)
            self.camera_params[:, :cp.shape[1]] = cp
        self.points = capi.as_f64(points).reshape(-1, 3)
        self.obs_image = capi.as_i32(obs_image)
        self.obs_point = capi.as_i32(obs_point)
        self.obs_xy = capi.as_f64(obs_xy).reshape(-1, 2)
        n_img, n_cam, n_pt = len(self.image_camera), len(self.camera_model), len(self.points)

        def u8(a, n):
            return np.zeros(n, np.uint8) if a is None else np.ascontiguousarray(a, dtype=np.uint8)
        self.image_pose_constant = u8(image_pose_constant, n_img)
        self.image_tvec_constant_mask = u8(image_tvec_constant_mask, n_img)
        self.point_constant = u8(point_constant, n_pt)
        self.camera_param_constant_mask = (np.zeros(n_cam, np.uint16) if camera_param_constant_mask is None
                                           else np.ascontiguousarray(camera_param_constant_mask, dtype=np.uint16))
        self.sem_maps = None if sem_maps is None else np.ascontiguousarray(sem_maps, dtype=np.float32)
        self.sem_weight = float(sem_weight)

    @property
    def num_observations(self):
        return len(self.obs_image)

    def pin(self):
        """Move the input arrays into page-locked host memory (sba_host_alloc): sba_create then DMAs them straight
        from these buffers.  Needs a CUDA device."""
        arena = capi.PinnedArena()
        for name in ("image_qvec", "image_tvec", "camera_params", "points", "obs_image", "obs_point", "obs_xy", "sem_maps"):
            a = getattr(self, name)
            if a is not None:
                setattr(self, name, arena.copy(a))
        self._arena = arena
        return self

    def desc(self, shard_rank=0, shard_count=1, with_maps=True):
        d = capi.ProblemDesc()
        d.num_images, d.num_cameras, d.num_points = len(self.image_camera), len(self.camera_model), len(self.points)
        d.num_observations = len(self.obs_image)
        d.image_qvec = capi.ptr(self.image_qvec, capi._dp)
        d.image_tvec = capi.ptr(self.image_tvec, capi._dp)
        d.image_camera = capi.ptr(self.image_camera, capi._ip)
        d.image_pose_constant = capi.ptr(self.image_pose_constant, capi._u8p)
        d.image_tvec_constant_mask = capi.ptr(self.image_tvec_constant_mask, capi._u8p)
        d.camera_model = capi.ptr(self.camera_model, capi._ip)
        d.camera_params = capi.ptr(self.camera_params, capi._dp)
        d.camera_param_constant_mask = capi.ptr(self.camera_param_constant_mask, capi._u16p)
        d.points = capi.ptr(self.points, capi._dp)
        d.point_constant = capi.ptr(self.point_constant, capi._u8p)
        d.obs_image = capi.ptr(self.obs_image, capi._ip)
        d.obs_point = capi.ptr(self.obs_point, capi._ip)
        d.obs_xy = capi.ptr(self.obs_xy, capi._dp)
        if not with_maps:
            d.sem_weight = self.sem_weight  # the maps themselves are already on the device (ResidentMaps)
        elif self.sem_maps is not None:
            d.sem_maps = capi.ptr(self.sem_maps, capi._fp)
            d.sem_height, d.sem_width = self.sem_maps.shape[1], self.sem_maps.shape[2]
            d.sem_weight = self.sem_weight
        d.shard_rank, d.shard_count = shard_rank, shard_count
        d.global_num_observations = d.num_observations
        return d

    def counts(self):
        """Structural counts of the reduced program (host only; no device needed)."""
        s = capi.Summary()
        capi.check(capi.load().sba_problem_counts(C.byref(self.desc()), C.byref(s)))
        return s

    def shard_masks(self, nranks, rank):
        pk = np.zeros(len(self.points), np.uint8)
        ok = np.zeros(len(self.obs_image), np.uint8)
        capi.check(capi.load().sba_shard_points(C.byref(self.desc()), nranks, rank,
                                                capi.ptr(pk, capi._u8p), capi.ptr(ok, capi._u8p)))
        return pk, ok


class ResidentMaps:
    """Label maps kept on the device across problems (`sba_maps_create`): the reference's semantic adjusters read every
    image's map once and keep it for all their cost functors (ReadDepthAndSemanticMaps,
    /root/reference/src/optim/semantic_bundle_adjustment.cc:1020-1066; `semantic_maps_`, semantic_bundle_adjustment.h:260)."""

    def __init__(self, sem_maps):
        maps = np.ascontiguousarray(sem_maps, dtype=np.float32)
        if maps.ndim != 3:
            raise ValueError("sem_maps must be [num_images, height, width]")
        self.lib = capi.load()
        self._m = C.c_void_p()
        self.shape = maps.shape
        capi.check(self.lib.sba_maps_create(maps.shape[0], maps.shape[1], maps.shape[2], capi.ptr(maps, capi._fp), C.byref(self._m)))

    def close(self):
        if self._m:
            self.lib.sba_maps_destroy(self._m)
            self._m = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Solver:
    """One `sba_handle`: device-resident problem + LM solver.  `maps`: a ResidentMaps set to build the problem over
    (the problem's own `sem_maps`, if any, are then not uploaded again; its `sem_weight` applies)."""

    def __init__(self, problem, shard_rank=0, shard_count=1, maps=None):
        self.problem = problem
        self.lib = capi.load()
        self._h = C.c_void_p()
        self._desc = problem.desc(shard_rank, shard_count, with_maps=maps is None)
        if maps is None:
            capi.check(self.lib.sba_create(C.byref(self._desc), C.byref(self._h)))
        else:
            if not maps._m:
                raise ValueError("the ResidentMaps set has been closed")
            capi.check(self.lib.sba_create_with_maps(C.byref(self._desc), maps._m, C.byref(self._h)))

    def close(self):
        if self._h:
            self.lib.sba_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def init_comm(self, nranks, rank, unique_id):
        buf = (C.c_char * capi.COMM_ID_BYTES).from_buffer_copy(bytes(unique_id))
        capi.check(self.lib.sba_comm_init(self._h, nranks, rank, C.cast(buf, C.c_void_p)))

    def evaluate(self, loss_type=capi.LOSS_TRIVIAL, loss_scale=1.0, apply_loss=True, want=("r", "pose", "intr", "point")):
        n = self.problem.num_observations
        out = {}
        cost = C.c_double()
        bufs = {
            "r": np.zeros((n, 2)) if "r" in want else None,
            "pose": np.zeros((n, 2, 6)) if "pose" in want else None,
            "intr": np.zeros((n, 2, capi.MAX_CAMERA_PARAMS)) if "intr" in want else None,
            "point": np.zeros((n, 2, 3)) if "point" in want else None,
            "qvec": np.zeros((n, 2, 4)) if "qvec" in want else None,
            "sem": np.zeros(n) if "sem" in want else None,
        }
        capi.check(self.lib.sba_evaluate(self._h, loss_type, loss_scale, 1 if apply_loss else 0, C.byref(cost),
                                         *[capi.ptr(bufs[k], capi._dp) for k in ("r", "pose", "intr", "point", "qvec", "sem")]))
        out = {k: v for k, v in bufs.items() if v is not None}
        out["cost"] = cost.value
        if "sem_jac" in want:  # Jacobian of the third (semantic) row of the evaluation just made
            sp, si, sx = np.zeros((n, 6)), np.zeros((n, capi.MAX_CAMERA_PARAMS)), np.zeros((n, 3))
            capi.check(self.lib.sba_get_semantic_jacobian(self._h, capi.ptr(sp, capi._dp), capi.ptr(si, capi._dp),
                                                          capi.ptr(sx, capi._dp)))
            out["sem_pose"], out["sem_intr"], out["sem_point"] = sp, si, sx
        return out

    def cost_no_alloc(self, loss_type=capi.LOSS_TRIVIAL, loss_scale=1.0, device_out=None):
        """Cost at the current parameters through sba_cost_no_alloc (static device scratch, no allocation inside the
        call).  device_out: optional device address (int) of one double on the handle's GPU that also receives it."""
        cost = C.c_double()
        capi.check(self.lib.sba_cost_no_alloc(self._h, loss_type, loss_scale, C.byref(cost),
                                              C.c_void_p(device_out) if device_out else None))
        return cost.value

    def solve(self, options=None, callback=None):
        o = options if options is not None else capi.default_solver_options()
        s = capi.Summary()
        if callback is not None:
            def _cb(it_ptr, _user):
                return int(callback(it_ptr.contents) or 0)
            cb = capi.ITERATION_CALLBACK(_cb)
        else:
            cb = C.cast(None, capi.ITERATION_CALLBACK)
        capi.check(self.lib.sba_solve(self._h, C.byref(o), cb, None, C.byref(s)))
        return s

    def reset(self):
        capi.check(self.lib.sba_reset(self._h))

    def trace(self, cap=4096):
        rows = (capi.IterationSummary * cap)()
        n = C.c_int32()
        capi.check(self.lib.sba_get_trace(self._h, rows, cap, C.byref(n)))
        return [rows[i] for i in range(n.value)]

    def parameters(self):
        p = self.problem
        q = np.empty_like(p.image_qvec); t = np.empty_like(p.image_tvec)
        cam = np.empty_like(p.camera_params)
        # a shard returns its own range of points; the rest keeps the caller's values
        X = np.empty_like(p.points) if self._desc.shard_count <= 1 else p.points.copy()
        capi.check(self.lib.sba_get_parameters(self._h, capi.ptr(q, capi._dp), capi.ptr(t, capi._dp),
                                               capi.ptr(cam, capi._dp), capi.ptr(X, capi._dp)))
        return q, t, cam, X

    def bench_stage(self, stage, repeat=10):
        ms = C.c_double(); nbytes = C.c_int64()
        capi.check(self.lib.sba_bench_stage(self._h, stage.encode(), repeat, C.byref(ms), C.byref(nbytes)))
        return ms.value, nbytes.value

    def schur_product(self, x=None, use_blocks=False):
        """y = (S + D^2) x of the current linearisation; x=None returns the reduced dimension."""
        n = C.c_int32()
        if x is None:
            capi.check(self.lib.sba_schur_product(self._h, 0, None, None, C.byref(n)))
            return n.value
        x = capi.as_f64(x)
        y = np.zeros_like(x)
        capi.check(self.lib.sba_schur_product(self._h, 1 if use_blocks else 0, capi.ptr(x, capi._dp), capi.ptr(y, capi._dp), C.byref(n)))
        return y

    def launches(self):
        return int(self.lib.sba_kernel_launch_count(self._h))


def comm_unique_id():
    buf = (C.c_char * capi.COMM_ID_BYTES)()
    capi.check(capi.load().sba_comm_get_unique_id(C.cast(buf, C.c_void_p)))
    return bytes(buf)
