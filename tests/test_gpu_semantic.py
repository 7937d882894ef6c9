"""GPU parity of the fork's semantic residuals (reference rows a10, a12 and extension X1):
K-SEM-PAIR, K-CYL-IOU and K-SEM-BILIN against the CPU restatement.  Residuals are integer-derived
(0/1 mismatches, pixel counts): exact equality is required; central-difference Jacobians use the
same deltas, so they must agree to the last bit of the quotient as well."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def plane_scene(sba, n_img=4, H=120, W=160, seed=0):
    """Cameras looking at the world plane Z = 5 with a checker label pattern on it; depth maps are the
    exact ray/plane depths so the depth gate of semantic_cost_functions.h:184-196 passes."""
    rng = np.random.default_rng(seed)
    f, cx, cy = 1.2 * W, W / 2.0, H / 2.0
    w = rng.normal(0, 0.03, (n_img, 3))
    ang = np.linalg.norm(w, axis=1)
    q = np.concatenate([np.cos(ang / 2)[:, None], (np.sin(ang / 2) / ang)[:, None] * w], axis=1)
    t = rng.normal(0, 0.15, (n_img, 3))
    R = sba.scene.rotmat_from_quat(q)
    yy, xx = np.mgrid[0:H, 0:W]
    d = np.stack([(xx - cx) / f, (yy - cy) / f, np.ones_like(xx, float)], -1)  # rays in the camera frame
    depth = np.zeros((n_img, H, W), np.float32); label = np.zeros((n_img, H, W), np.float32)
    for i in range(n_img):
        # world = R^T (lam d - t); world_z = 5
        rt = R[i].T
        a = d @ rt[2]; b = rt[2] @ t[i]
        lam = (5.0 + b) / a
        Xw = (lam[..., None] * d - t[i]) @ rt.T
        depth[i] = lam
        label[i] = (np.floor(Xw[..., 0] * 6.0) + 3 * np.floor(Xw[..., 1] * 6.0)) % 8
    depth[:, :3, :] = 0.0  # zero-depth rows (skipped by the adjuster, semantic_bundle_adjustment.cc:809-814)
    cams = np.tile([f, cx, cy], (n_img, 1))
    return q, t, cams, depth, label


def pair_blocks(n_img, H, W, step, depth):
    """AddImagePairToProblem's enumeration (semantic_bundle_adjustment.cc:656-661, 796-814): all ordered
    pairs, y outer, x inner on a `step` grid, zero-depth pixels skipped."""
    i1, i2, x1, y1 = [], [], [], []
    for a in range(n_img):
        for b in range(n_img):
            if a == b:
                continue
            for y in range(0, H, step):
                for x in range(0, W, step):
                    if depth[a, y, x] < 1e-4:
                        continue
                    i1.append(a); i2.append(b); x1.append(x); y1.append(y)
    return [np.asarray(v, np.int32) for v in (i1, i2, x1, y1)]


def sem_desc(sba, q, t, cams, depth, label, blocks, thr=2.0, rel=1e-3):
    capi = sba.capi
    d = capi.SemanticDesc()
    keep = [np.ascontiguousarray(a) for a in (q, t, cams, depth, label)] + list(blocks)
    d.num_images, d.height, d.width = depth.shape
    d.image_qvec, d.image_tvec, d.camera_params = (capi.ptr(keep[k], capi._dp) for k in range(3))
    d.depth_maps, d.semantic_maps = capi.ptr(keep[3], capi._fp), capi.ptr(keep[4], capi._fp)
    d.num_blocks = len(blocks[0])
    d.block_image1, d.block_image2, d.block_x1, d.block_y1 = (capi.ptr(b, capi._ip) for b in blocks)
    d.depth_error_threshold, d.relative_step_size = thr, rel
    return d, keep


@pytest.mark.parametrize("with_jac", [False, True])
def test_semantic_pair_parity(sba, oracle, with_jac):
    q, t, cams, depth, label = plane_scene(sba)
    blocks = pair_blocks(4, 120, 160, 5, depth)
    d, keep = sem_desc(sba, q, t, cams, depth, label, blocks)
    n = d.num_blocks
    assert n > 5000
    rg, sg, jg = np.zeros(n), np.zeros(n, np.int32), np.zeros((n, 14))
    ro, so, jo = np.zeros(n), np.zeros(n, np.int32), np.zeros((n, 14))
    ms = C.c_double()
    capi = sba.capi
    capi.check(capi.load().sba_semantic_evaluate(C.byref(d), capi.ptr(rg, capi._dp), capi.ptr(sg, capi._ip),
                                                 capi.ptr(jg, capi._dp) if with_jac else None, C.byref(ms)))
    oracle.oracle_semantic_evaluate(C.byref(d), 0, capi.ptr(ro, capi._dp), capi.ptr(so, capi._ip),
                                    capi.ptr(jo, capi._dp) if with_jac else None)
    assert set(np.unique(so)) == {-2, -1, 10} or set(np.unique(so)) == {-1, 10}  # OOB, (bad depth), valid all occur
    assert 0.01 < ro.mean() < 0.9                                                  # both outcomes occur
    assert np.array_equal(sg, so) and np.array_equal(rg, ro)
    if with_jac:
        assert np.count_nonzero(jo) > 0
        assert np.array_equal(jg, jo)


def test_semantic_pair_bad_depth_and_threshold(sba, oracle):
    q, t, cams, depth, label = plane_scene(sba, seed=3)
    depth[1] += 3.0  # image 1's depth map disagrees with the geometry -> status -2 for pairs into image 1
    blocks = pair_blocks(4, 120, 160, 8, depth)
    d, keep = sem_desc(sba, q, t, cams, depth, label, blocks, thr=0.5)
    n = d.num_blocks
    rg, sg = np.zeros(n), np.zeros(n, np.int32)
    ro, so = np.zeros(n), np.zeros(n, np.int32)
    capi = sba.capi
    capi.check(capi.load().sba_semantic_evaluate(C.byref(d), capi.ptr(rg, capi._dp), capi.ptr(sg, capi._ip), None, None))
    oracle.oracle_semantic_evaluate(C.byref(d), 0, capi.ptr(ro, capi._dp), capi.ptr(so, capi._ip), None)
    assert (so == -2).sum() > 100
    assert np.array_equal(sg, so) and np.array_equal(rg, ro)


def cylinder_scene(sba, n_img=3, H=120, W=160, seed=0):
    rng = np.random.default_rng(seed)
    f, cx, cy = 1.2 * W, W / 2.0, H / 2.0
    q = np.tile([1.0, 0, 0, 0], (n_img, 1)) + rng.normal(0, 0.01, (n_img, 4))
    t = rng.normal(0, 0.1, (n_img, 3))
    cams = np.tile([f, cx, cy], (n_img, 1))
    # cylinders: axis roughly along world -y (image up), a few metres in front of the cameras
    n_cyl = 3
    ang = np.pi / 2 + rng.normal(0, 0.05, n_cyl)
    cq = np.stack([np.cos(ang / 2), np.sin(ang / 2), 0 * ang, 0 * ang], 1)   # rotate z-axis towards -y
    ct = np.stack([rng.uniform(-0.6, 0.6, n_cyl), rng.uniform(0.3, 0.6, n_cyl), rng.uniform(4, 6, n_cyl)], 1)
    cr = rng.uniform(0.08, 0.2, n_cyl); ch = rng.uniform(0.8, 1.5, n_cyl)
    return q, t, cams, cq, ct, cr, ch


def test_cylinder_iou_parity(sba, oracle):
    H, W = 120, 160
    q, t, cams, cq, ct, cr, ch = cylinder_scene(sba, H=H, W=W)
    capi = sba.capi
    # masks: the oracle's own rasterisation of slightly perturbed cylinders -> IoU strictly between 0 and 1
    masks = np.zeros((3, H, W), np.uint8)
    pts = np.zeros(8)
    for i in range(3):
        for c in range(3):
            ok = oracle.oracle_cylinder_quadrilateral(capi.ptr(q[i].copy(), capi._dp), capi.ptr((t[i] + 0.03).copy(), capi._dp),
                                                      capi.ptr(cams[i].copy(), capi._dp), capi.ptr(cq[c].copy(), capi._dp),
                                                      capi.ptr(ct[c].copy(), capi._dp), cr[c] * 1.2, ch[c], capi.ptr(pts, capi._dp))
            assert ok == 1
            m = np.zeros((H, W), np.uint8)
            oracle.oracle_draw_quadrilateral(capi.ptr(pts, capi._dp), H, W, m.ctypes.data_as(C.POINTER(C.c_uint8)))
            masks[i] |= m if c == i % 3 else 0
    bi = np.repeat(np.arange(3), 3).astype(np.int32); bc = np.tile(np.arange(3), 3).astype(np.int32)
    d = capi.CylinderDesc()
    keep = [np.ascontiguousarray(a) for a in (q, t, cams, masks, cq, ct, cr, ch, bi, bc)]
    d.num_images, d.height, d.width = 3, H, W
    d.image_qvec, d.image_tvec, d.camera_params = (capi.ptr(keep[k], capi._dp) for k in range(3))
    d.masks = capi.ptr(keep[3], capi._u8p)
    d.num_cylinders = 3
    d.cyl_qvec, d.cyl_tvec, d.cyl_radius, d.cyl_height = (capi.ptr(keep[k], capi._dp) for k in range(4, 8))
    d.num_blocks = 9
    d.block_image, d.block_cylinder = capi.ptr(keep[8], capi._ip), capi.ptr(keep[9], capi._ip)
    d.relative_step_size = 1e-3
    lib = capi.load()
    # (1) the quadrilateral: device vertices against the oracle's, block by block
    pg, okg = np.zeros((9, 8)), np.zeros(9, np.int32)
    capi.check(lib.sba_cylinder_quadrilateral(C.byref(d), capi.ptr(pg, capi._dp), capi.ptr(okg, capi._ip)))
    po = np.zeros((9, 8))
    for k in range(9):
        i, c = bi[k], bc[k]
        ok = oracle.oracle_cylinder_quadrilateral(capi.ptr(q[i].copy(), capi._dp), capi.ptr(t[i].copy(), capi._dp),
                                                  capi.ptr(cams[i].copy(), capi._dp), capi.ptr(cq[c].copy(), capi._dp),
                                                  capi.ptr(ct[c].copy(), capi._dp), cr[c], ch[c], capi.ptr(pts, capi._dp))
        assert ok == 1 and okg[k] == 1
        po[k] = pts
    assert np.abs(pg - po).max() <= 1e-11   # pixels; acos / sin / cos of the device and of libm differ by ulps
    # (2) the rasteriser: IDENTICAL vertices in, identical masks out (drawQuadrilateral is integer work on given doubles)
    for k in range(9):
        mg, mo = np.zeros((H, W), np.uint8), np.zeros((H, W), np.uint8)
        capi.check(lib.sba_draw_quadrilateral(capi.ptr(po[k].copy(), capi._dp), H, W, capi.ptr(mg, capi._u8p)))
        oracle.oracle_draw_quadrilateral(capi.ptr(po[k].copy(), capi._dp), H, W, mo.ctypes.data_as(C.POINTER(C.c_uint8)))
        assert mo.sum() > 50 and np.array_equal(mg, mo)
    # (3) the residual 1 - IoU and its central differences: pixel counts are integers, so blocks whose vertices agree
    # BITWISE must agree exactly; the others to the count one boundary pixel can move
    rg, jg, ro, jo = np.zeros(9), np.zeros((9, 16)), np.zeros(9), np.zeros((9, 16))
    capi.check(lib.sba_cylinder_evaluate(C.byref(d), capi.ptr(rg, capi._dp), capi.ptr(jg, capi._dp), None))
    oracle.oracle_cylinder_evaluate(C.byref(d), 0, capi.ptr(ro, capi._dp), capi.ptr(jo, capi._dp))
    assert ((ro > 0) & (ro < 1)).sum() >= 3 and np.count_nonzero(jo) > 0
    same = np.all(pg == po, axis=1)
    assert np.array_equal(rg[same], ro[same])
    assert np.array_equal(rg, ro), (rg - ro, np.abs(pg - po).max(axis=1))
    # central differences over 32 perturbed evaluations per block: the same deltas, the same counts
    assert np.array_equal(jg, jo), np.argwhere(jg != jo)


def test_semantic_bilinear_rows(sba, oracle):
    """Extension X1, raw row: weight * bilinear(label map, projection) per observation (the row inside the residual
    blocks, its Jacobian and the solve are covered by tests/test_semantic_row.py)."""
    prob, _ = sba.scene.make_scene(n_cam=5, n_pt=400, mean_track=4.0, model="SIMPLE_PINHOLE", width=160, height=120,
                                   refine_intr=False, seed=2, label_maps=True)
    prob.sem_weight = 0.5
    s = sba.Solver(prob)
    ev = s.evaluate(want=("r", "sem"), apply_loss=False)
    ref = np.zeros(prob.num_observations)
    d = prob.desc()
    assert oracle.oracle_ba_sem_rows(C.byref(d), sba.capi.ptr(ref, sba.capi._dp)) == 0
    assert np.count_nonzero(ref) > 0.5 * len(ref)
    assert np.allclose(ev["sem"], ref, rtol=1e-12, atol=1e-12)
    s.close()
