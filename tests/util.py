"""Shared helpers of the parity tests: run the oracle / the CUDA library on one FlatProblem."""
import ctypes as C

import numpy as np


def oracle_evaluate(ora, capi, prob, loss_type=0, loss_scale=1.0, apply_loss=True, threads=0, qvec=False):
    n = prob.num_observations
    out = dict(r=np.zeros((n, 2)), pose=np.zeros((n, 2, 6)), intr=np.zeros((n, 2, capi.MAX_CAMERA_PARAMS)),
               point=np.zeros((n, 2, 3)))
    if qvec:
        out["qvec"] = np.zeros((n, 2, 4))
    cost = C.c_double()
    d = prob.desc()
    rc = ora.oracle_ba_evaluate(C.byref(d), loss_type, loss_scale, 1 if apply_loss else 0, threads, C.byref(cost),
                                capi.ptr(out["r"], capi._dp), capi.ptr(out["pose"], capi._dp),
                                capi.ptr(out["intr"], capi._dp), capi.ptr(out["point"], capi._dp),
                                capi.ptr(out.get("qvec"), capi._dp))
    assert rc == 0
    out["cost"] = cost.value
    return out


def oracle_solve(ora, capi, prob, options, threads=0, cap=512):
    s = capi.Summary()
    tr = (capi.IterationSummary * cap)()
    n = C.c_int32()
    q = np.zeros_like(prob.image_qvec); t = np.zeros_like(prob.image_tvec)
    cam = np.zeros_like(prob.camera_params); X = np.zeros_like(prob.points)
    d = prob.desc()
    rc = ora.oracle_ba_solve(C.byref(d), C.byref(options), threads, C.byref(s), tr, cap, C.byref(n),
                             capi.ptr(q, capi._dp), capi.ptr(t, capi._dp), capi.ptr(cam, capi._dp), capi.ptr(X, capi._dp))
    return rc, s, [tr[i] for i in range(n.value)], (q, t, cam, X)


def rel_block_err(a, b):
    """max over observations of ||a_k - b_k||_F / max(||b_k||_F, tiny)."""
    a = a.reshape(len(a), -1); b = b.reshape(len(b), -1)
    nb = np.linalg.norm(b, axis=1)
    return float(np.max(np.linalg.norm(a - b, axis=1) / np.maximum(nb, 1e-300) * (nb > 0))) if len(a) else 0.0


def assert_eval_parity(ev, ref, prob, tol=1e-10):
    # residuals: relative to the magnitudes before the cancelling subtraction (parity definitions)
    scale = np.maximum(np.abs(prob.obs_xy), 1.0)
    assert np.all(np.abs(ev["r"] - ref["r"]) <= tol * np.maximum(scale, np.abs(ref["r"]))), \
        float(np.max(np.abs(ev["r"] - ref["r"]) / scale))
    for k in ("pose", "intr", "point"):
        assert (ev[k][ref[k] == 0] == 0).all(), "structurally zero entries must be exactly zero: " + k
        assert rel_block_err(ev[k], ref[k]) <= tol, (k, rel_block_err(ev[k], ref[k]))
    assert abs(ev["cost"] - ref["cost"]) <= 1e-12 * max(abs(ref["cost"]), 1.0)


def oracle_evaluate_rows(ora, capi, prob, loss_type=0, loss_scale=1.0, apply_loss=True, threads=0):
    """All three rows of every block (the problem carries semantic maps): the two-row dict of oracle_evaluate plus
    sem [n], sem_pose [n,6], sem_intr [n,12], sem_point [n,3]."""
    n = prob.num_observations
    out = dict(r=np.zeros((n, 2)), pose=np.zeros((n, 2, 6)), intr=np.zeros((n, 2, capi.MAX_CAMERA_PARAMS)),
               point=np.zeros((n, 2, 3)), sem=np.zeros(n), sem_pose=np.zeros((n, 6)),
               sem_intr=np.zeros((n, capi.MAX_CAMERA_PARAMS)), sem_point=np.zeros((n, 3)))
    cost = C.c_double()
    d = prob.desc()
    rc = ora.oracle_ba_evaluate_rows(C.byref(d), loss_type, loss_scale, 1 if apply_loss else 0, threads, C.byref(cost),
                                     *[capi.ptr(out.get(k), capi._dp) for k in
                                       ("r", "pose", "intr", "point", "qvec", "sem", "sem_pose", "sem_intr", "sem_point")])
    assert rc == 0
    out["cost"] = cost.value
    return out


def assert_sem_parity(ev, ref, tol=1e-10):
    """Third (semantic) row: residual relative to max(|r_s|, 1); Jacobian blocks in the Frobenius norm per block,
    structurally zero entries exactly zero."""
    assert np.all(np.abs(ev["sem"] - ref["sem"]) <= tol * np.maximum(np.abs(ref["sem"]), 1.0)), \
        float(np.max(np.abs(ev["sem"] - ref["sem"])))
    for k in ("sem_pose", "sem_intr", "sem_point"):
        assert (ev[k][ref[k] == 0] == 0).all(), "structurally zero entries must be exactly zero: " + k
        assert rel_block_err(ev[k], ref[k]) <= tol, (k, rel_block_err(ev[k], ref[k]))
