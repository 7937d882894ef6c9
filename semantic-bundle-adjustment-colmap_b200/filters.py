"""Mapper-side filters that run right after global bundle adjustment in incremental SfM
(/root/reference/src/controllers/incremental_mapper.cc:110-130, src/sfm/incremental_mapper.cc:658-663, 711) -- SURVEY
section 8, row f4.  The per-observation arithmetic (squared reprojection error, depth) is ONE device launch over all track
elements (`sba_observation_errors`, csrc/filter.cu); the bookkeeping below replays the reference's loops over the results:

  FilterObservationsWithNegativeDepth           /root/reference/src/base/reconstruction.cc:647-665
  FilterPoints3DWithLargeReprojectionError      /root/reference/src/base/reconstruction.cc:1472-1523
  DeleteObservation / DeletePoint3D             /root/reference/src/base/reconstruction.cc:236-277

Poses, intrinsics and points do not change while filtering, so evaluating every observation up front and consulting the
table inside the sequential loops gives exactly what the reference computes one element at a time.  `errors_fn` lets the
tests substitute the oracle for the device call; the product default needs a CUDA device.
"""
import ctypes as C

import numpy as np

from . import _capi as capi
from .bundle_adjustment import kInvalidPoint3DId
from .problem import FlatProblem


def flatten_observations(rec):
    """Every (image, point2D) with a 3D point, image-major in container order.  Returns (FlatProblem, keys) where
    keys[k] = (image_id, point2D_idx) of observation k."""
    img_index = {iid: i for i, iid in enumerate(rec.images)}
    cam_index = {cid: c for c, cid in enumerate(rec.cameras)}
    pt_index = {pid: q for q, pid in enumerate(rec.points3D)}
    obs_img, obs_pt, obs_xy, keys = [], [], [], []
    for iid, img in rec.images.items():
        for idx, p2 in enumerate(img.points2D):
            if p2.HasPoint3D() and p2.point3D_id in pt_index:
                obs_img.append(img_index[iid]); obs_pt.append(pt_index[p2.point3D_id]); obs_xy.append(p2.xy)
                keys.append((iid, idx))
    params = np.zeros((len(rec.cameras), capi.MAX_CAMERA_PARAMS))
    for c, cam in enumerate(rec.cameras.values()):
        params[c, :len(cam.params)] = cam.params
    prob = FlatProblem(np.array([im.qvec for im in rec.images.values()]).reshape(-1, 4),
                       np.array([im.tvec for im in rec.images.values()]).reshape(-1, 3),
                       [cam_index[im.camera_id] for im in rec.images.values()],
                       [cam.ModelId() for cam in rec.cameras.values()], params,
                       np.array([p.xyz for p in rec.points3D.values()]).reshape(-1, 3),
                       obs_img, obs_pt, np.asarray(obs_xy, dtype=np.float64).reshape(-1, 2))
    return prob, keys


def device_observation_errors(prob):
    """(squared reprojection error, depth) per observation of `prob` through the C ABI (device)."""
    n = prob.num_observations
    err, depth = np.zeros(n), np.zeros(n)
    d = prob.desc()
    capi.check(capi.load().sba_observation_errors(C.byref(d), capi.ptr(err, capi._dp), capi.ptr(depth, capi._dp)))
    return err, depth


def delete_point3D(rec, point3D_id):
    for el in rec.points3D[point3D_id].track.Elements():
        rec.images[el.image_id].points2D[el.point2D_idx].point3D_id = kInvalidPoint3DId
    del rec.points3D[point3D_id]


def delete_observation(rec, image_id, point2D_idx):
    pid = rec.images[image_id].points2D[point2D_idx].point3D_id
    p = rec.points3D[pid]
    if p.track.Length() <= 2:
        delete_point3D(rec, pid)
        return
    p.track.elements = [e for e in p.track.elements if not (e.image_id == image_id and e.point2D_idx == point2D_idx)]
    rec.images[image_id].points2D[point2D_idx].point3D_id = kInvalidPoint3DId


def filter_observations_with_negative_depth(rec, errors_fn=device_observation_errors):
    prob, keys = flatten_observations(rec)
    if not keys:
        return 0
    _, depth = errors_fn(prob)
    eps = np.finfo(np.float64).eps
    num_filtered = 0
    for (iid, idx), z in zip(keys, depth):  # keys are in the reference's loop order: images outer, Points2D inner
        if rec.images[iid].points2D[idx].HasPoint3D() and not (z >= eps):
            delete_observation(rec, iid, idx)
            num_filtered += 1
    return num_filtered


def filter_points3D_with_large_reprojection_error(rec, max_reproj_error, point3D_ids=None, errors_fn=device_observation_errors):
    prob, keys = flatten_observations(rec)
    err = dict(zip(keys, errors_fn(prob)[0].tolist())) if keys else {}
    max_sq = max_reproj_error * max_reproj_error
    num_filtered = 0
    for pid in (list(rec.points3D) if point3D_ids is None else list(point3D_ids)):
        if pid not in rec.points3D:
            continue
        p = rec.points3D[pid]
        length = p.track.Length()
        if length < 2:
            num_filtered += length
            delete_point3D(rec, pid)
            continue
        total, to_delete = 0.0, []
        for el in p.track.Elements():
            e = err[(el.image_id, el.point2D_idx)]
            if e > max_sq:
                to_delete.append(el)
            else:
                total += np.sqrt(e)
        if len(to_delete) >= length - 1:
            num_filtered += length
            delete_point3D(rec, pid)
        else:
            num_filtered += len(to_delete)
            for el in to_delete:
                delete_observation(rec, el.image_id, el.point2D_idx)
            p.error = total / p.track.Length()
    return num_filtered


def device_normalize(qvec, tvec, registered, points, extent, p0, p1, use_images):
    """sba_normalize on flat arrays (in place); returns (scale, translation)."""
    scale, tr = C.c_double(), np.zeros(3)
    capi.check(capi.load().sba_normalize(len(qvec), capi.ptr(qvec, capi._dp), capi.ptr(tvec, capi._dp), capi.ptr(registered, capi._u8p),
                                         len(points), capi.ptr(points, capi._dp), extent, p0, p1, 1 if use_images else 0,
                                         C.byref(scale), capi.ptr(tr, capi._dp)))
    return scale.value, tr


def normalize(rec, extent=10.0, p0=0.1, p1=0.9, use_images=True, registered_image_ids=None, normalize_fn=device_normalize):
    """Reconstruction::Normalize (/root/reference/src/base/reconstruction.cc:321-345), the mapper's step after global BA
    (/root/reference/src/sfm/incremental_mapper.cc:711).  The mirror Reconstruction has no registration flag: every image
    counts as registered unless `registered_image_ids` is given.  Returns (scale, translation) of the similarity applied
    ((1, 0) when fewer than two elements: nothing changes).  `normalize_fn` lets the tests substitute the oracle."""
    if not extent > 0:
        raise ValueError("extent must be positive")  # CHECK_GT, :323
    ids = list(rec.images)
    reg = np.ones(len(ids), np.uint8) if registered_image_ids is None else np.array([1 if i in set(registered_image_ids) else 0 for i in ids], np.uint8)
    q = np.array([rec.images[i].qvec for i in ids], dtype=np.float64).reshape(-1, 4)
    t = np.array([rec.images[i].tvec for i in ids], dtype=np.float64).reshape(-1, 3)
    pids = list(rec.points3D)
    X = np.array([rec.points3D[p].xyz for p in pids], dtype=np.float64).reshape(-1, 3)
    scale, tr = normalize_fn(q, t, reg, X, extent, p0, p1, use_images)
    for k, i in enumerate(ids):
        rec.images[i].qvec[:] = q[k]; rec.images[i].tvec[:] = t[k]
    for k, p in enumerate(pids):
        rec.points3D[p].xyz[:] = X[k]
    return scale, tr
