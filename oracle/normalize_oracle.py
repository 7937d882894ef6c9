"""TEST INFRASTRUCTURE ONLY (imported by tests/): numpy restatement of Reconstruction::Normalize.

  Normalize                     /root/reference/src/base/reconstruction.cc:321-345
  ComputeBoundsAndCentroid      /root/reference/src/base/reconstruction.cc:358-422
  Transform                     /root/reference/src/base/reconstruction.cc:424-431
  SimilarityTransform3::TransformPoint / ::TransformPose   /root/reference/src/base/similarity_transform.cc:211-240
  ProjectionCenterFromPose, QuaternionToRotationMatrix, RotationMatrixToQuaternion   /root/reference/src/base/pose.cc:63-91, 164-171
The quaternion <-> matrix conversions are Eigen's (Quaternion::toRotationMatrix, the matrix constructor), restated from
its published algorithm.  Pinned by the reference's own TestNormalize (/root/reference/src/base/reconstruction_test.cc:230-286:
seven images, four calls with different extents and percentile ranges, every resulting tvec asserted to 1e-6), replayed
statement by statement on this restatement and on the device path (tests/test_filters.py); beyond that test the device path
is held to this restatement at 1e-12 and to the properties the transform must have."""
import numpy as np


def quat_to_rot(q):
    q = np.asarray(q, dtype=np.float64)
    n = np.linalg.norm(q)
    w, x, y, z = (1.0, 0.0, 0.0, 0.0) if n == 0 else q / n
    tx, ty, tz = 2 * x, 2 * y, 2 * z
    twx, twy, twz, txx, txy, txz, tyy, tyz, tzz = tx * w, ty * w, tz * w, tx * x, ty * x, tz * x, ty * y, tz * y, tz * z
    return np.array([[1 - (tyy + tzz), txy - twz, txz + twy], [txy + twz, 1 - (txx + tzz), tyz - twx], [txz - twy, tyz + twx, 1 - (txx + tyy)]])


def rot_to_quat(m):
    q = np.zeros(4)
    t = m[0, 0] + m[1, 1] + m[2, 2]
    if t > 0:
        t = np.sqrt(t + 1.0)
        q[0] = 0.5 * t
        t = 0.5 / t
        q[1], q[2], q[3] = (m[2, 1] - m[1, 2]) * t, (m[0, 2] - m[2, 0]) * t, (m[1, 0] - m[0, 1]) * t
    else:
        i = 0
        if m[1, 1] > m[0, 0]:
            i = 1
        if m[2, 2] > m[i, i]:
            i = 2
        j = (i + 1) % 3
        k = (j + 1) % 3
        t = np.sqrt(m[i, i] - m[j, j] - m[k, k] + 1.0)
        q[1 + i] = 0.5 * t
        t = 0.5 / t
        q[0] = (m[k, j] - m[j, k]) * t
        q[1 + j] = (m[j, i] + m[i, j]) * t
        q[1 + k] = (m[k, i] + m[i, k]) * t
    return q


def bounds_and_centroid(coords, p0, p1):
    c = np.sort(np.asarray(coords, dtype=np.float64).astype(np.float32), axis=0)   # three independent per-axis sorts of floats
    n = len(c)
    P0 = int(p0 * (n - 1)) if n > 3 else 0
    P1 = int(p1 * (n - 1)) if n > 3 else n - 1
    mean = c[P0:P1 + 1].astype(np.float64).sum(axis=0) / (P1 - P0 + 1)
    return c[P0].astype(np.float64), c[P1].astype(np.float64), mean


def normalize(qvec, tvec, registered, points, extent, p0, p1, use_images):
    """In place, same signature as filters.device_normalize."""
    reg = np.nonzero(registered)[0]
    n_el = len(reg) if use_images else len(points)
    if n_el < 2:
        return 1.0, np.zeros(3)
    if use_images:
        coords = np.array([-(quat_to_rot(qvec[i]).T @ tvec[i]) for i in reg])
    else:
        coords = points
    lo, hi, mean = bounds_and_centroid(coords, p0, p1)
    old = np.linalg.norm(hi - lo)
    scale = 1.0 if old < np.finfo(np.float64).eps else extent / old
    tr = -scale * mean
    points[:] = scale * points + tr
    inv = 1.0 / scale
    it = -(inv * tr)
    for i in range(len(qvec)):
        R = quat_to_rot(qvec[i])
        nt = (R @ it + tvec[i]) * scale
        qvec[i] = rot_to_quat((R * inv) * scale)
        tvec[i] = nt
    return scale, tr
