"""Writes tests/golden/model_bin and tests/golden/model_txt with the REFERENCE's own model writer
(/root/reference/scripts/python/read_write_model.py, imported in place -- it exists only in the build container, so the
files it wrote are committed).  The model is tiny and synthetic: three cameras of different models, four images (one
without points), six points; values are not round so that formatting and byte layout are exercised."""
import os
import sys

import numpy as np

sys.path.insert(0, "/root/reference/scripts/python")
import read_write_model as rw  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def build():
    cameras = {
        1: rw.Camera(id=1, model="SIMPLE_PINHOLE", width=640, height=480, params=np.array([655.123, 320.5, 240.25])),
        2: rw.Camera(id=2, model="SIMPLE_RADIAL", width=1000, height=1000, params=np.array([1200.0 / 7, 500.0, 500.0, -0.0123456789])),
        5: rw.Camera(id=5, model="OPENCV", width=800, height=600,
                     params=np.array([651.123, 655.123, 386.123, 511.123, -0.471, 0.223, -0.001, 0.001])),
    }
    rng = np.random.default_rng(7)
    images = {}
    spec = {1: (1, "a/frame_000001.png", 4), 2: (2, "frame 2.jpg", 3), 3: (5, "c.png", 5), 7: (1, "empty.png", 0)}
    for iid, (cid, name, n) in spec.items():
        q = rng.normal(size=4); q /= np.linalg.norm(q)
        images[iid] = rw.Image(id=iid, qvec=q, tvec=rng.normal(size=3), camera_id=cid, name=name,
                               xys=rng.uniform(0, 600, (n, 2)), point3D_ids=np.full(n, -1, dtype=np.int64))
    tracks = {11: [(1, 0), (2, 1)], 12: [(1, 1), (3, 0)], 13: [(1, 3), (2, 0), (3, 4)], 20: [(2, 2), (3, 1)],
              21: [(3, 2), (1, 2)], 4000000000: [(3, 3), (2, 1)]}
    points = {}
    for pid, tr in tracks.items():
        for iid, idx in tr:
            images[iid].point3D_ids[idx] = pid
        points[pid] = rw.Point3D(id=pid, xyz=rng.normal(size=3) * 3, rgb=rng.integers(0, 256, 3).astype(np.uint8),
                                 error=float(rng.uniform(0.1, 2.0)), image_ids=np.array([t[0] for t in tr]),
                                 point2D_idxs=np.array([t[1] for t in tr]))
    return cameras, images, points


if __name__ == "__main__":
    cameras, images, points = build()
    for ext, sub in ((".bin", "model_bin"), (".txt", "model_txt")):
        out = os.path.join(HERE, sub)
        os.makedirs(out, exist_ok=True)
        rw.write_model(cameras, images, points, out, ext)
    back = rw.read_model(os.path.join(HERE, "model_bin"), ".bin")
    assert sorted(back[0]) == [1, 2, 5] and sorted(back[1]) == [1, 2, 3, 7] and len(back[2]) == 6
    print("wrote", os.listdir(os.path.join(HERE, "model_bin")), os.listdir(os.path.join(HERE, "model_txt")))
