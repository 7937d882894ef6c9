"""ctypes binding of include/sba_b200.h (the C-ABI drop-in boundary).

The structures below mirror the header field for field; `load()` opens the in-tree
`libsba_b200.so` built by `csrc/Makefile` and fails loudly when it is missing -- there is
no CPU fallback anywhere in the product path (the CPU restatement under `oracle/` is
test infrastructure and is never imported from here).
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libsba_b200.so")

MAX_CAMERA_PARAMS = 12
COMM_ID_BYTES = 128

# enum sba_camera_model == CameraModel::kModelId (src/base/camera_models.h:189-349)
CAMERA_MODELS = {
    "SIMPLE_PINHOLE": 0, "PINHOLE": 1, "SIMPLE_RADIAL": 2, "RADIAL": 3, "OPENCV": 4,
    "OPENCV_FISHEYE": 5, "FULL_OPENCV": 6, "FOV": 7, "SIMPLE_RADIAL_FISHEYE": 8,
    "RADIAL_FISHEYE": 9, "THIN_PRISM_FISHEYE": 10,
}
MODEL_NUM_PARAMS = {0: 3, 1: 4, 2: 4, 3: 5, 4: 8, 5: 8, 6: 12, 7: 5, 8: 4, 9: 5, 10: 12}
# FocalLengthIdxs / PrincipalPointIdxs / ExtraParamsIdxs per model (camera_models.cc InitializeParamsInfo)
MODEL_FOCAL_IDXS = {0: [0], 1: [0, 1], 2: [0], 3: [0], 4: [0, 1], 5: [0, 1], 6: [0, 1], 7: [0, 1], 8: [0], 9: [0], 10: [0, 1]}
MODEL_PP_IDXS = {0: [1, 2], 1: [2, 3], 2: [1, 2], 3: [1, 2], 4: [2, 3], 5: [2, 3], 6: [2, 3], 7: [2, 3], 8: [1, 2], 9: [1, 2], 10: [2, 3]}
MODEL_EXTRA_IDXS = {0: [], 1: [], 2: [3], 3: [3, 4], 4: [4, 5, 6, 7], 5: [4, 5, 6, 7], 6: list(range(4, 12)), 7: [4], 8: [3], 9: [3, 4], 10: list(range(4, 12))}

LOSS_TRIVIAL, LOSS_SOFT_L1, LOSS_CAUCHY = 0, 1, 2
SOLVER_AUTO, DENSE_SCHUR, SPARSE_SCHUR, ITERATIVE_SCHUR = -1, 0, 1, 2
OK, ERR_INVALID_ARGUMENT, ERR_NO_DEVICE, ERR_CUDA, ERR_UNKNOWN_CAMERA_MODEL, ERR_EMPTY_PROBLEM, \
    ERR_ALREADY_SOLVED, ERR_COMM, ERR_INTERNAL = range(9)
CONVERGENCE, NO_CONVERGENCE, FAILURE, USER_SUCCESS, USER_FAILURE = range(5)

_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_u8p = C.POINTER(C.c_uint8)
_u16p = C.POINTER(C.c_uint16)
_fp = C.POINTER(C.c_float)


class ProblemDesc(C.Structure):
    _fields_ = [
        ("num_images", C.c_int32), ("num_cameras", C.c_int32), ("num_points", C.c_int32),
        ("num_observations", C.c_int64),
        ("image_qvec", _dp), ("image_tvec", _dp), ("image_camera", _ip),
        ("image_pose_constant", _u8p), ("image_tvec_constant_mask", _u8p),
        ("camera_model", _ip), ("camera_params", _dp), ("camera_param_constant_mask", _u16p),
        ("points", _dp), ("point_constant", _u8p),
        ("obs_image", _ip), ("obs_point", _ip), ("obs_xy", _dp),
        ("sem_maps", _fp), ("sem_height", C.c_int32), ("sem_width", C.c_int32), ("sem_weight", C.c_double),
        ("shard_rank", C.c_int32), ("shard_count", C.c_int32), ("global_num_observations", C.c_int64),
    ]


class SolverOptions(C.Structure):
    _fields_ = [
        ("linear_solver_type", C.c_int32), ("num_config_images", C.c_int32),
        ("max_num_iterations", C.c_int32), ("max_linear_solver_iterations", C.c_int32),
        ("min_linear_solver_iterations", C.c_int32), ("max_num_consecutive_invalid_steps", C.c_int32),
        ("function_tolerance", C.c_double), ("gradient_tolerance", C.c_double),
        ("parameter_tolerance", C.c_double),
        ("loss_function_type", C.c_int32), ("loss_function_scale", C.c_double),
        ("initial_trust_region_radius", C.c_double), ("max_trust_region_radius", C.c_double),
        ("min_trust_region_radius", C.c_double), ("min_relative_decrease", C.c_double),
        ("min_lm_diagonal", C.c_double), ("max_lm_diagonal", C.c_double), ("eta", C.c_double),
        ("jacobi_scaling", C.c_int32), ("update_state_every_iteration", C.c_int32),
    ]


class IterationSummary(C.Structure):
    _fields_ = [
        ("iteration", C.c_int32), ("step_is_valid", C.c_int32), ("step_is_successful", C.c_int32),
        ("linear_solver_iterations", C.c_int32),
        ("cost", C.c_double), ("cost_change", C.c_double), ("gradient_max_norm", C.c_double),
        ("step_norm", C.c_double), ("relative_decrease", C.c_double),
        ("trust_region_radius", C.c_double), ("eta", C.c_double),
    ]


class Summary(C.Structure):
    _fields_ = [
        ("termination_type", C.c_int32), ("linear_solver_type_used", C.c_int32),
        ("num_iterations", C.c_int32), ("num_successful_steps", C.c_int32),
        ("num_unsuccessful_steps", C.c_int32), ("num_residual_blocks", C.c_int32),
        ("num_residuals", C.c_int32), ("num_residuals_reduced", C.c_int32),
        ("num_parameter_blocks_reduced", C.c_int32), ("num_parameters_reduced", C.c_int32),
        ("num_effective_parameters_reduced", C.c_int32),
        ("num_linear_solver_iterations", C.c_int64),
        ("initial_cost", C.c_double), ("final_cost", C.c_double), ("fixed_cost", C.c_double),
        ("total_time_in_seconds", C.c_double),
        ("time_evaluate_s", C.c_double), ("time_eliminate_s", C.c_double),
        ("time_linear_solve_s", C.c_double), ("time_update_s", C.c_double), ("device_time_s", C.c_double),
    ]

    def as_dict(self):
        return {name: getattr(self, name) for name, _ in self._fields_}


class SemanticDesc(C.Structure):
    _fields_ = [
        ("num_images", C.c_int32), ("height", C.c_int32), ("width", C.c_int32),
        ("image_qvec", _dp), ("image_tvec", _dp), ("camera_params", _dp),
        ("depth_maps", _fp), ("semantic_maps", _fp),
        ("num_blocks", C.c_int64),
        ("block_image1", _ip), ("block_image2", _ip), ("block_x1", _ip), ("block_y1", _ip),
        ("depth_error_threshold", C.c_double), ("relative_step_size", C.c_double),
    ]


class SemanticProblem(C.Structure):
    _fields_ = [
        ("num_images", C.c_int32), ("height", C.c_int32), ("width", C.c_int32),
        ("image_qvec", _dp), ("image_tvec", _dp), ("camera_params", _dp),
        ("depth_maps", _fp), ("semantic_maps", _fp),
        ("image_pose_constant", _u8p), ("image_tvec_constant_mask", _u8p),
        ("pixel_step", C.c_int32), ("depth_error_threshold", C.c_double), ("relative_step_size", C.c_double),
    ]


class CylinderProblem(C.Structure):
    _fields_ = [
        ("num_images", C.c_int32), ("height", C.c_int32), ("width", C.c_int32),
        ("image_qvec", _dp), ("image_tvec", _dp), ("camera_params", _dp), ("masks", _u8p),
        ("image_pose_constant", _u8p), ("image_tvec_constant_mask", _u8p),
        ("num_cylinders", C.c_int32), ("cylinder_kind", C.c_int32), ("cylinder_params", _dp),
        ("refine_geometry", C.c_int32), ("relative_step_size", C.c_double),
        ("num_points", C.c_int32), ("points", _dp), ("point_constant", _u8p),
        ("num_landmarks", C.c_int64), ("landmark_image", _ip), ("landmark_point", _ip), ("landmark_xy", _dp),
        ("landmark_weight", C.c_double),
    ]


class CylinderDesc(C.Structure):
    _fields_ = [
        ("num_images", C.c_int32), ("height", C.c_int32), ("width", C.c_int32),
        ("image_qvec", _dp), ("image_tvec", _dp), ("camera_params", _dp), ("masks", _u8p),
        ("num_cylinders", C.c_int32),
        ("cyl_qvec", _dp), ("cyl_tvec", _dp), ("cyl_radius", _dp), ("cyl_height", _dp),
        ("num_blocks", C.c_int64), ("block_image", _ip), ("block_cylinder", _ip),
        ("relative_step_size", C.c_double),
    ]


class RigDesc(C.Structure):
    _fields_ = [
        ("num_rig_poses", C.c_int32), ("rig_qvec", _dp), ("rig_tvec", _dp),
        ("num_rel_poses", C.c_int32), ("rel_qvec", _dp), ("rel_tvec", _dp),
        ("num_cameras", C.c_int32), ("camera_model", _ip), ("camera_params", _dp),
        ("num_points", C.c_int32), ("points", _dp),
        ("num_blocks", C.c_int64), ("block_rig", _ip), ("block_rel", _ip), ("block_camera", _ip), ("block_point", _ip),
        ("block_xy", _dp),
    ]


RIG_JAC_COLS = 4 + 3 + 4 + 3 + 3 + 12


class RigProblem(C.Structure):
    _fields_ = [
        ("num_rig_poses", C.c_int32), ("rig_qvec", _dp), ("rig_tvec", _dp), ("rig_pose_constant", _u8p),
        ("num_rel_poses", C.c_int32), ("rel_qvec", _dp), ("rel_tvec", _dp), ("rel_pose_constant", _u8p),
        ("rel_tvec_constant_mask", _u8p),
        ("num_cameras", C.c_int32), ("camera_model", _ip), ("camera_params", _dp), ("camera_param_constant_mask", _u16p),
        ("num_points", C.c_int32), ("points", _dp), ("point_constant", _u8p),
        ("num_blocks", C.c_int64), ("block_rig", _ip), ("block_rel", _ip), ("block_camera", _ip), ("block_point", _ip),
        ("block_xy", _dp),
    ]

ITERATION_CALLBACK = C.CFUNCTYPE(C.c_int, C.POINTER(IterationSummary), C.c_void_p)

# every entry point include/sba_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "sba_default_solver_options": (None, [C.POINTER(SolverOptions)]),
    "sba_create": (C.c_int, [C.POINTER(ProblemDesc), C.POINTER(C.c_void_p)]),
    "sba_maps_create": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, _fp, C.POINTER(C.c_void_p)]),
    "sba_maps_destroy": (None, [C.c_void_p]),
    "sba_create_with_maps": (C.c_int, [C.POINTER(ProblemDesc), C.c_void_p, C.POINTER(C.c_void_p)]),
    "sba_problem_counts": (C.c_int, [C.POINTER(ProblemDesc), C.POINTER(Summary)]),
    "sba_destroy": (None, [C.c_void_p]),
    "sba_set_device": (C.c_int, [C.c_int32]),
    "sba_reset": (C.c_int, [C.c_void_p]),
    "sba_host_alloc": (C.c_void_p, [C.c_uint64]),
    "sba_host_free": (None, [C.c_void_p]),
    "sba_last_error": (C.c_char_p, []),
    "sba_version": (C.c_char_p, []),
    "sba_device_count": (C.c_int, []),
    "sba_evaluate": (C.c_int, [C.c_void_p, C.c_int32, C.c_double, C.c_int32, _dp, _dp, _dp, _dp, _dp, _dp, _dp]),
    "sba_get_semantic_jacobian": (C.c_int, [C.c_void_p, _dp, _dp, _dp]),
    "sba_cost_no_alloc": (C.c_int, [C.c_void_p, C.c_int32, C.c_double, _dp, C.c_void_p]),
    "sba_solve": (C.c_int, [C.c_void_p, C.POINTER(SolverOptions), ITERATION_CALLBACK, C.c_void_p, C.POINTER(Summary)]),
    "sba_get_trace": (C.c_int, [C.c_void_p, C.POINTER(IterationSummary), C.c_int32, _ip]),
    "sba_get_parameters": (C.c_int, [C.c_void_p, _dp, _dp, _dp, _dp]),
    "sba_bench_stage": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int32, _dp, C.POINTER(C.c_int64)]),
    "sba_release_cached_memory": (None, []),
    "sba_observation_errors": (C.c_int, [C.POINTER(ProblemDesc), _dp, _dp]),
    "sba_normalize": (C.c_int, [C.c_int32, _dp, _dp, _u8p, C.c_int64, _dp, C.c_double, C.c_double, C.c_double, C.c_int32, _dp, _dp]),
    "sba_schur_product": (C.c_int, [C.c_void_p, C.c_int32, _dp, _dp, _ip]),
    "sba_kernel_launch_count": (C.c_int64, [C.c_void_p]),
    "sba_comm_get_unique_id": (C.c_int, [C.c_void_p]),
    "sba_comm_init": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "sba_shard_points": (C.c_int, [C.POINTER(ProblemDesc), C.c_int32, C.c_int32, _u8p, _u8p]),
    "sba_semantic_evaluate": (C.c_int, [C.POINTER(SemanticDesc), _dp, _ip, _dp, _dp]),
    "sba_cylinder_evaluate": (C.c_int, [C.POINTER(CylinderDesc), _dp, _dp, _dp]),
    "sba_cylinder_quadrilateral": (C.c_int, [C.POINTER(CylinderDesc), _dp, _ip]),
    "sba_draw_quadrilateral": (C.c_int, [_dp, C.c_int32, C.c_int32, _u8p]),
    "sba_semantic_solve": (C.c_int, [C.POINTER(SemanticProblem), C.POINTER(SolverOptions), ITERATION_CALLBACK, C.c_void_p,
                                     C.POINTER(Summary), C.POINTER(IterationSummary), C.c_int32, _ip]),
    "sba_cylinder_solve": (C.c_int, [C.POINTER(CylinderProblem), C.POINTER(SolverOptions), ITERATION_CALLBACK, C.c_void_p,
                                     C.POINTER(Summary), C.POINTER(IterationSummary), C.c_int32, _ip]),
    "sba_rig_evaluate": (C.c_int, [C.POINTER(RigDesc), _dp, _dp, _dp]),
    "sba_rig_solve": (C.c_int, [C.POINTER(RigProblem), C.POINTER(SolverOptions), ITERATION_CALLBACK, C.c_void_p,
                                C.POINTER(Summary), C.POINTER(IterationSummary), C.c_int32, _ip]),
}

_lib = None


class SbaError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("sba_b200 error %d: %s" % (code, msg))
        self.code = code


def load():
    """Open libsba_b200.so and type every exported entry point. Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libsba_b200.so is not built (%s): run `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C %s/csrc`; there is no CPU fallback" % (LIB_PATH, HERE))
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise SbaError(rc, (load().sba_last_error() or b"").decode())


def ptr(a, typ):
    """numpy array (or None) -> typed ctypes pointer; the caller keeps `a` alive."""
    if a is None:
        return C.cast(None, typ)
    return a.ctypes.data_as(typ)


def as_f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def as_i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def default_solver_options():
    o = SolverOptions()
    load().sba_default_solver_options(C.byref(o))
    return o


class PinnedArena:
    """numpy arrays backed by page-locked host memory (sba_host_alloc): inputs kept here reach the device by DMA
    straight from the caller's buffer.  Keep the arena alive as long as the arrays are in use."""

    def __init__(self):
        self._blocks = []

    def copy(self, a):
        a = np.ascontiguousarray(a)
        if a.nbytes == 0:
            return a
        p = load().sba_host_alloc(a.nbytes)
        if not p:
            raise SbaError(3, (load().sba_last_error() or b"").decode())
        self._blocks.append(p)
        out = np.frombuffer((C.c_char * a.nbytes).from_address(p), dtype=a.dtype).reshape(a.shape)
        out[...] = a
        return out

    def close(self):
        for p in self._blocks:
            load().sba_host_free(p)
        self._blocks = []
