"""oracle/_ref/libpba_ref.so is the REFERENCE's own lib/PBA CPU solver compiled where its sources lie (oracle/Makefile
`ref`).  A different LM than Ceres (multiplicative damping, float storage), so it pins nothing bit-wise -- but it is
reference-owned code, and on a SIMPLE_RADIAL scene it must land on the same minimum as the oracle's Ceres restatement."""
import ctypes as C
import os

import numpy as np
import pytest

from util import oracle_solve

REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libpba_ref.so")


@pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref not built (needs /root/reference; build() makes it)")
def test_pba_reference_agrees_with_oracle_minimum(sba, oracle):
    lib = C.CDLL(REF)
    lib.pba_ref_solve.argtypes = [C.POINTER(sba.capi.ProblemDesc), C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double),
                                  C.POINTER(C.c_double)]
    prob, _ = sba.scene.make_config("C1")
    d = prob.desc()
    out = np.zeros(5)
    assert lib.pba_ref_solve(C.byref(d), 25, 4, 1, out.ctypes.data_as(C.POINTER(C.c_double)), None) == 0
    o = sba.capi.default_solver_options()
    o.max_num_iterations = 25
    rc, s, _, _ = oracle_solve(oracle, sba.capi, prob, o)
    assert rc == 0
    assert out[0] == pytest.approx(s.initial_cost, rel=1e-4)   # same residuals at the start (float storage in PBA)
    assert out[1] == pytest.approx(s.final_cost, rel=2e-3)     # same minimum
    # outside PBA's accepted inputs (bundle_adjustment.cc:629-651): not SIMPLE_RADIAL
    prob2, _ = sba.scene.make_scene(n_cam=4, n_pt=50, mean_track=3.0, model="SIMPLE_PINHOLE", width=640, height=480, seed=5)
    d2 = prob2.desc()
    assert lib.pba_ref_solve(C.byref(d2), 5, 1, 1, out.ctypes.data_as(C.POINTER(C.c_double)), None) == 1
