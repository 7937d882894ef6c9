"""Seeded synthetic bundle-adjustment scenes of the BASELINE.json shapes.

Modelled on the reference's own test generator
(/root/reference/src/optim/bundle_adjustment_test.cc:111-184: points U[-1,1]^3, f = 1.2 * width,
pixel noise U[-2,2]) but with partial visibility and look-at cameras on a trajectory, as
SURVEY.md section 8d specifies.  Residual order is the reference loop nest with a sorted
image set: images ascending, within an image observations by ascending point index
(bundle_adjustment.cc:330-332, 366-412).
"""
import numpy as np

from . import _capi as capi
from .problem import FlatProblem

# BASELINE.json configs: (cameras, points, mean observations per point, model, free intrinsics)
CONFIGS = {
    "C1": dict(n_cam=10, n_pt=2000, mean_track=7.5, model="SIMPLE_RADIAL", width=1000, height=1000, refine_intr=True),
    "C2": dict(n_cam=50, n_pt=20000, mean_track=7.5, model="SIMPLE_PINHOLE", width=640, height=480, refine_intr=False,
               semantic="distance"),
    "C3": dict(n_cam=500, n_pt=200000, mean_track=7.5, model="SIMPLE_RADIAL", width=640, height=480, refine_intr=True,
               semantic="distance"),
    "C4": dict(n_cam=1778, n_pt=993000, mean_track=5.035, model="SIMPLE_RADIAL", width=1000, height=1000, refine_intr=True),
    "C5": dict(n_cam=5000, n_pt=2000000, mean_track=7.5, model="SIMPLE_PINHOLE", width=640, height=480, refine_intr=False,
               semantic="distance"),
}


def quat_from_rotmat(R):
    """(n,3,3) rotation matrices -> (n,4) quaternions (w,x,y,z), w >= 0."""
    R = np.asarray(R)
    n = R.shape[0]
    q = np.zeros((n, 4))
    tr = R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2]
    for i in range(n):  # n = cameras (<= a few thousand)
        m = R[i]
        if tr[i] > 0:
            s = np.sqrt(tr[i] + 1.0) * 2
            q[i] = [0.25 * s, (m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s]
        elif m[0, 0] > m[1, 1] and m[0, 0] > m[2, 2]:
            s = np.sqrt(1.0 + m[0, 0] - m[1, 1] - m[2, 2]) * 2
            q[i] = [(m[2, 1] - m[1, 2]) / s, 0.25 * s, (m[0, 1] + m[1, 0]) / s, (m[0, 2] + m[2, 0]) / s]
        elif m[1, 1] > m[2, 2]:
            s = np.sqrt(1.0 + m[1, 1] - m[0, 0] - m[2, 2]) * 2
            q[i] = [(m[0, 2] - m[2, 0]) / s, (m[0, 1] + m[1, 0]) / s, 0.25 * s, (m[1, 2] + m[2, 1]) / s]
        else:
            s = np.sqrt(1.0 + m[2, 2] - m[0, 0] - m[1, 1]) * 2
            q[i] = [(m[1, 0] - m[0, 1]) / s, (m[0, 2] + m[2, 0]) / s, (m[1, 2] + m[2, 1]) / s, 0.25 * s]
    q[q[:, 0] < 0] *= -1
    return q / np.linalg.norm(q, axis=1, keepdims=True)


def rotmat_from_quat(q):
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = np.empty((len(q), 3, 3))
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - w * z); R[:, 0, 2] = 2 * (x * z + w * y)
    R[:, 1, 0] = 2 * (x * y + w * z); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - w * x)
    R[:, 2, 0] = 2 * (x * z - w * y); R[:, 2, 1] = 2 * (y * z + w * x); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


def quat_mul(a, b):
    w1, x1, y1, z1 = a.T
    w2, x2, y2, z2 = b.T
    return np.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                     w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2], axis=1)


def _lookat_cameras(n_cam, rng):
    """Cameras on a closed trajectory of radius ~10 around the unit cube, looking at the origin."""
    th = 2 * np.pi * np.arange(n_cam) / max(n_cam, 1)
    rad = 10.0 + rng.uniform(-2.0, 2.0, n_cam)
    C = np.stack([rad * np.cos(th), 1.5 * np.sin(3 * th) + rng.uniform(-0.5, 0.5, n_cam), rad * np.sin(th)], axis=1)
    z = -C / np.linalg.norm(C, axis=1, keepdims=True)
    up = np.tile(np.array([0.0, 1.0, 0.0]), (n_cam, 1))
    x = np.cross(up, z); x /= np.linalg.norm(x, axis=1, keepdims=True)
    y = np.cross(z, x)
    R = np.stack([x, y, z], axis=1)  # rows = camera axes: p_cam = R (X - C)
    # small random rotation about a random axis (|w| <= 0.1)
    w = rng.normal(size=(n_cam, 3)); w /= np.linalg.norm(w, axis=1, keepdims=True)
    ang = rng.uniform(0, 0.1, n_cam)
    dq = np.concatenate([np.cos(ang / 2)[:, None], np.sin(ang / 2)[:, None] * w], axis=1)
    q = quat_mul(dq, quat_from_rotmat(R))
    Rq = rotmat_from_quat(q)
    t = -np.einsum("nij,nj->ni", Rq, C)
    return q, t


def _project(model, params, q, t, X, obs_img, obs_pt):
    R = rotmat_from_quat(q)
    p = np.einsum("nij,nj->ni", R[obs_img], X[obs_pt]) + t[obs_img]
    u, v = p[:, 0] / p[:, 2], p[:, 1] / p[:, 2]
    prm = params[obs_img]
    if model == "SIMPLE_PINHOLE":
        return np.stack([prm[:, 0] * u + prm[:, 1], prm[:, 0] * v + prm[:, 2]], axis=1)
    if model == "SIMPLE_RADIAL":
        rad = prm[:, 3] * (u * u + v * v)
        return np.stack([prm[:, 0] * (u + u * rad) + prm[:, 1], prm[:, 0] * (v + v * rad) + prm[:, 2]], axis=1)
    raise ValueError("scene generator supports SIMPLE_PINHOLE / SIMPLE_RADIAL, got " + model)


def make_scene(n_cam, n_pt, mean_track=7.5, model="SIMPLE_RADIAL", width=1000, height=1000, refine_intr=True,
               seed=0, window=32, noise_px=2.0, perturb=True, label_maps=False, n_classes=8, semantic=None,
               sem_weight=0.1, perturb_scale=1.0):
    """Returns (FlatProblem, ground_truth dict). Gauge as the reference controller fixes it
    (src/controllers/bundle_adjustment.cc:95-96): pose 0 constant, tvec[0] of pose 1 constant."""
    rng = np.random.default_rng(seed)
    X = rng.uniform(-1.0, 1.0, (n_pt, 3))
    q_gt, t_gt = _lookat_cameras(n_cam, rng)
    f = 1.2 * width
    if model == "SIMPLE_RADIAL":
        params = np.tile(np.array([f, width / 2.0, height / 2.0, 0.0]), (n_cam, 1))
    else:
        params = np.tile(np.array([f, width / 2.0, height / 2.0]), (n_cam, 1))
    # visibility: track length 2 + Poisson(mean - 2), cameras drawn from a window on the trajectory
    win = int(min(window, n_cam))
    length = np.minimum(2 + rng.poisson(max(mean_track - 2.0, 0.0), n_pt), win).astype(np.int64)
    center = rng.integers(0, n_cam, n_pt)
    obs_img_parts, obs_pt_parts = [], []
    chunk = 1 << 18
    for s in range(0, n_pt, chunk):
        e = min(n_pt, s + chunk)
        keys = rng.random((e - s, win))
        order = np.argsort(keys, axis=1)  # random permutation of window offsets per point
        take = np.arange(win)[None, :] < length[s:e, None]
        cams = (center[s:e, None] + order - win // 2) % n_cam
        obs_img_parts.append(cams[take])
        obs_pt_parts.append(np.broadcast_to(np.arange(s, e)[:, None], cams.shape)[take])
    obs_img = np.concatenate(obs_img_parts).astype(np.int32)
    obs_pt = np.concatenate(obs_pt_parts).astype(np.int32)
    order = np.lexsort((obs_pt, obs_img))  # reference order: image-major, then point2D index
    obs_img, obs_pt = obs_img[order], obs_pt[order]
    xy = _project(model, params, q_gt, t_gt, X, obs_img, obs_pt)
    xy = xy + rng.uniform(-noise_px, noise_px, xy.shape)

    q0, t0, X0 = q_gt.copy(), t_gt.copy(), X.copy()
    if perturb:
        t0 = t0 + rng.normal(0, 0.01 * perturb_scale, t0.shape)
        w = rng.normal(0, 0.002 * perturb_scale, (n_cam, 3))
        ang = np.linalg.norm(w, axis=1)
        dq = np.concatenate([np.cos(ang / 2)[:, None], (np.sin(ang / 2) / np.maximum(ang, 1e-300))[:, None] * w], axis=1)
        q0 = quat_mul(dq, q0)
        q0 /= np.linalg.norm(q0, axis=1, keepdims=True)
        X0 = X0 + rng.normal(0, 0.01 * perturb_scale, X0.shape)
        q0[0], t0[0] = q_gt[0], t_gt[0]  # the gauge-fixing pose stays exact

    pose_const = np.zeros(n_cam, np.uint8); pose_const[0] = 1
    tvec_mask = np.zeros(n_cam, np.uint8)
    if n_cam > 1:
        tvec_mask[1] = 1
    mid = capi.CAMERA_MODELS[model]
    if refine_intr:  # default BA: focal + extra free, principal point fixed (bundle_adjustment.h:58-64)
        cmask = sum(1 << j for j in capi.MODEL_PP_IDXS[mid])
    else:
        cmask = (1 << capi.MODEL_NUM_PARAMS[mid]) - 1
    # semantic maps of extension X1: "labels" (piecewise constant class ids) or "distance" (truncated distance to the
    # nearest label boundary, the smooth map a bilinear residual row has a gradient on); label_maps=True == "labels"
    sem = None
    if semantic is None and label_maps:
        semantic = "labels"
    if semantic == "labels":
        sem = make_label_maps(n_cam, height, width, n_classes, seed + 1)
    elif semantic == "distance":
        sem = make_distance_maps(n_cam, height, width, n_classes, seed + 1)
    elif semantic is not None:
        raise ValueError("semantic must be None, 'labels' or 'distance'")
    prob = FlatProblem(q0, t0, np.arange(n_cam), np.full(n_cam, mid), params, X0, obs_img, obs_pt, xy,
                       image_pose_constant=pose_const, image_tvec_constant_mask=tvec_mask,
                       camera_param_constant_mask=np.full(n_cam, cmask, np.uint16), sem_maps=sem,
                       sem_weight=sem_weight)
    return prob, dict(q=q_gt, t=t_gt, X=X, params=params)


def make_config(name, seed=0, **over):
    cfg = dict(CONFIGS[name]); cfg.update(over)
    return make_scene(seed=seed, **cfg)


def make_label_maps(n_img, height, width, n_classes=8, seed=1, n_seeds=64):
    """Piecewise-constant Voronoi label maps (float32 class ids), one per image."""
    rng = np.random.default_rng(seed)
    maps = np.empty((n_img, height, width), np.float32)
    yy, xx = np.mgrid[0:height, 0:width]
    step = 8  # Voronoi on a coarse grid, then nearest-upsampled: cheap and still piecewise constant
    ys, xs = yy[::step, ::step], xx[::step, ::step]
    for i in range(n_img):
        sx = rng.uniform(0, width, n_seeds); sy = rng.uniform(0, height, n_seeds)
        lab = rng.integers(0, n_classes, n_seeds)
        d = (xs[..., None] - sx) ** 2 + (ys[..., None] - sy) ** 2
        coarse = lab[np.argmin(d, axis=-1)].astype(np.float32)
        maps[i] = np.kron(coarse, np.ones((step, step), np.float32))[:height, :width]
    return maps


def make_distance_maps(n_img, height, width, n_classes=8, seed=1, n_unique=16, truncate=24.0):
    """Distance maps (float32, pixels, truncated): Euclidean distance to the nearest boundary of a Voronoi label map.
    `n_unique` distinct maps are generated and repeated over the images (the content does not matter to the solver,
    every image still owns its own copy in memory)."""
    from scipy import ndimage
    nu = int(min(n_unique, n_img))
    labels = make_label_maps(nu, height, width, n_classes, seed)
    uniq = np.empty((nu, height, width), np.float32)
    for i in range(nu):
        lab = labels[i]
        edge = np.zeros((height, width), bool)
        edge[:, 1:] |= lab[:, 1:] != lab[:, :-1]
        edge[1:, :] |= lab[1:, :] != lab[:-1, :]
        uniq[i] = np.minimum(ndimage.distance_transform_edt(~edge), truncate).astype(np.float32)
    maps = np.empty((n_img, height, width), np.float32)
    for i in range(n_img):
        maps[i] = uniq[i % nu]
    return maps
