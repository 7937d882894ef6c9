"""ORACLE (test infrastructure only): ctypes loader for oracle/liboracle.so.

`build()` compiles the C++ restatement with the Makefile next to this file; `load()` types
the extern "C" surface of oracle_api.cpp.  Structures are shared with the product's header
(include/sba_b200.h) so that the oracle consumes exactly the arrays the CUDA library does.
"""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle.so")
_lib = None


def build(ref=False):
    subprocess.check_call(["make", "-s", "-C", HERE])
    if ref and os.path.isdir("/root/reference/lib/PBA") and os.path.exists(os.path.join(HERE, "pba_ref_shim.cpp")):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


def load(capi):
    """capi: the product's ctypes module (struct definitions only are used)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB):
        build()
    lib = C.CDLL(LIB)
    dp, ip = C.POINTER(C.c_double), C.POINTER(C.c_int32)
    PD = C.POINTER(capi.ProblemDesc)
    lib.oracle_num_threads.restype = C.c_int
    lib.oracle_world_to_image.argtypes = [C.c_int, dp, C.c_double, C.c_double, dp, dp]
    lib.oracle_image_to_world.argtypes = [C.c_int, dp, C.c_double, C.c_double, dp, dp]
    lib.oracle_model_num_params.argtypes = [C.c_int]
    lib.oracle_model_num_params.restype = C.c_int
    lib.oracle_eval_observation.argtypes = [C.c_int] + [dp] * 10
    lib.oracle_eval_rig_observation.argtypes = [C.c_int] + [dp] * 8
    lib.oracle_eval_rig_observation_jac.argtypes = [C.c_int] + [dp] * 9
    lib.oracle_observation_error.argtypes = [C.c_int] + [dp] * 7
    lib.oracle_observation_error.restype = None
    lib.oracle_ba_evaluate.argtypes = [PD, C.c_int, C.c_double, C.c_int, C.c_int, dp, dp, dp, dp, dp, dp]
    lib.oracle_ba_evaluate.restype = C.c_int
    lib.oracle_ba_evaluate_rows.argtypes = [PD, C.c_int, C.c_double, C.c_int, C.c_int] + [dp] * 10
    lib.oracle_ba_evaluate_rows.restype = C.c_int
    lib.oracle_default_solver_options.argtypes = [C.POINTER(capi.SolverOptions)]
    lib.oracle_default_solver_options.restype = None
    lib.oracle_ba_counts.argtypes = [PD, C.POINTER(capi.Summary)]
    lib.oracle_ba_solve.argtypes = [PD, C.POINTER(capi.SolverOptions), C.c_int, C.POINTER(capi.Summary),
                                    C.POINTER(capi.IterationSummary), C.c_int, ip, dp, dp, dp, dp]
    lib.oracle_ba_solve.restype = C.c_int
    lib.oracle_ba_time_linearize.argtypes = [PD, C.c_int, C.c_double, C.c_int, C.c_int]
    lib.oracle_ba_time_linearize.restype = C.c_double
    lib.oracle_semantic_evaluate.argtypes = [C.POINTER(capi.SemanticDesc), C.c_int, dp, ip, dp]
    lib.oracle_cylinder_evaluate.argtypes = [C.POINTER(capi.CylinderDesc), C.c_int, dp, dp]
    lib.oracle_semantic_solve.argtypes = [C.POINTER(capi.SemanticProblem), C.POINTER(capi.SolverOptions), C.c_int,
                                          C.POINTER(capi.Summary), C.POINTER(capi.IterationSummary), C.c_int, ip]
    lib.oracle_semantic_solve.restype = C.c_int
    lib.oracle_cylinder_solve.argtypes = [C.POINTER(capi.CylinderProblem), C.POINTER(capi.SolverOptions), C.c_int,
                                          C.POINTER(capi.Summary), C.POINTER(capi.IterationSummary), C.c_int, ip]
    lib.oracle_cylinder_solve.restype = C.c_int
    lib.oracle_rig_solve.argtypes = [C.POINTER(capi.RigProblem), C.POINTER(capi.SolverOptions), C.c_int,
                                     C.POINTER(capi.Summary), C.POINTER(capi.IterationSummary), C.c_int, ip]
    lib.oracle_rig_solve.restype = C.c_int
    lib.oracle_powell_solve.argtypes = [C.POINTER(capi.SolverOptions), C.POINTER(C.c_double), C.POINTER(capi.Summary),
                                        C.POINTER(capi.IterationSummary), C.c_int32, C.POINTER(C.c_int32)]
    lib.oracle_powell_solve.restype = C.c_int
    lib.oracle_helloworld_solve.argtypes = lib.oracle_powell_solve.argtypes
    lib.oracle_helloworld_solve.restype = C.c_int
    lib.oracle_draw_quadrilateral.argtypes = [dp, C.c_int, C.c_int, C.POINTER(C.c_uint8)]
    lib.oracle_cylinder_quadrilateral.argtypes = [dp, dp, dp, dp, dp, C.c_double, C.c_double, dp]
    lib.oracle_cylinder_quadrilateral.restype = C.c_int
    lib.oracle_ba_sem_rows.argtypes = [PD, dp]
    lib.oracle_ba_sem_rows.restype = C.c_int
    lib.oracle_bilinear.argtypes = [C.POINTER(C.c_float), C.c_int, C.c_int, C.c_double, C.c_double, ip]
    lib.oracle_bilinear.restype = C.c_double
    _lib = lib
    return lib
