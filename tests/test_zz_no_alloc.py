"""sba_cost_no_alloc: the cost evaluation whose scratch is a static device array (no cudaMalloc inside the call).

No counterpart in the reference.  The oracle of this test is the existing
entry point: sba_evaluate's `cost` on the same handle.  Two calls in a row must both agree (the static scratch keeps the
previous call's contents), also after the parameters changed, and with the three-row residual.  Collected last on
purpose (file name), so that a failure here cannot hide other tests under `pytest -x`.
"""
import ctypes as C
import importlib

import numpy as np
import pytest

REL_TOL = 1e-12  # two fixed-order fp64 reductions of the same terms (tests/util.py holds r / J parity to 1e-10)


@pytest.fixture(scope="module")
def m():
    return importlib.import_module("semantic-bundle-adjustment-colmap_b200")


def _check(m, s, loss_type, loss_scale):
    ref = s.evaluate(loss_type=loss_type, loss_scale=loss_scale, want=())["cost"]
    assert np.isfinite(ref) and ref > 0.0
    for _ in range(2):  # two calls in a row share the static scratch
        got = s.cost_no_alloc(loss_type=loss_type, loss_scale=loss_scale)
        assert abs(got - ref) <= REL_TOL * ref, (got, ref)
    return ref


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["C1", "C2"])  # two rows (SIMPLE_RADIAL) and three rows (SIMPLE_PINHOLE + maps)
def test_no_alloc_cost_matches_evaluate(m, name):
    prob, _ = m.scene.make_config(name)
    s = m.Solver(prob)
    try:
        c0 = _check(m, s, m.capi.LOSS_TRIVIAL, 1.0)
        _check(m, s, m.capi.LOSS_CAUCHY, 2.0)
        # move the parameters (a short solve), then again: nothing of the earlier calls may leak through the scratch
        o = m.capi.default_solver_options()
        o.max_num_iterations = 3
        s.solve(o)
        c1 = _check(m, s, m.capi.LOSS_TRIVIAL, 1.0)
        assert c1 < c0
    finally:
        s.close()


@pytest.mark.gpu
def test_no_alloc_cost_device_output_and_two_handles(m):
    import torch
    pa, _ = m.scene.make_config("C1")
    pb, _ = m.scene.make_scene(seed=5, **dict(m.scene.CONFIGS["C1"]))
    a, b = m.Solver(pa), m.Solver(pb)
    try:
        ra = a.evaluate(want=())["cost"]
        rb = b.evaluate(want=())["cost"]
        out = torch.zeros(1, dtype=torch.float64, device="cuda")
        for _ in range(2):  # alternate handles: every call starts from the other handle's leftovers
            ga = a.cost_no_alloc(device_out=out.data_ptr())
            assert abs(out.item() - ga) == 0.0
            gb = b.cost_no_alloc()
            assert abs(ga - ra) <= REL_TOL * ra and abs(gb - rb) <= REL_TOL * rb
    finally:
        a.close(); b.close()


@pytest.mark.gpu
def test_no_alloc_rejects_bad_arguments(m):
    prob, _ = m.scene.make_config("C1")
    s = m.Solver(prob)
    try:
        rc = s.lib.sba_cost_no_alloc(s._h, m.capi.LOSS_TRIVIAL, 1.0, None, None)
        assert rc == m.capi.ERR_INVALID_ARGUMENT
        assert b"no output" in s.lib.sba_last_error()
    finally:
        s.close()
