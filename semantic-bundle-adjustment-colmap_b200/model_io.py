"""On-disk formats either side of the bundle-adjustment path: read model -> BA -> write model.

  * sparse model, binary and text   Reconstruction::Read/Write{Cameras,Images,Points3D}{Binary,Text}
                                    /root/reference/src/base/reconstruction.cc:1525-2100 (text headers and precision
                                    :1868-2010), mirrored by /root/reference/scripts/python/read_write_model.py:101-451
  * float32 TIFF maps               matrixFromTiff, /root/reference/src/util/matrix_vis.h:130-176 (32 bits per pixel only;
                                    FreeImage keeps scanlines bottom-up and the reference flips them back, so the matrix is
                                    the image in ordinary top-down row order)
  * cylinders.txt                   Cylinder::to_string / Cylinder(const std::string&),
                                    /root/reference/src/util/cylinder.h:289-329, 608-628: `q w x y z t x y z r R h H`

Host-side Python over the Reconstruction mirror of bundle_adjustment.py; no device work.  The golden files under
tests/golden/model_{bin,txt} were written by the reference's own read_write_model.py (tests/golden/make_model_golden.py).
"""
import os
import struct

import numpy as np

from . import _capi as capi
from .bundle_adjustment import Camera, Image, Point2D, Reconstruction, kInvalidPoint3DId

MODEL_NAMES = {v: k for k, v in capi.CAMERA_MODELS.items()}
_INVALID_U64 = 0xFFFFFFFFFFFFFFFF


def _g17(v):
    return "%.17g" % v  # std::ostream with precision(17)


# ---------------------------------------------------------------- cameras
def read_cameras_binary(path):
    cams = {}
    with open(path, "rb") as f:
        (n,) = struct.unpack("<Q", f.read(8))
        for _ in range(n):
            camera_id, model_id, width, height = struct.unpack("<iiQQ", f.read(24))
            if model_id not in capi.MODEL_NUM_PARAMS:
                raise ValueError("Camera model does not exist")
            k = capi.MODEL_NUM_PARAMS[model_id]
            params = struct.unpack("<%dd" % k, f.read(8 * k))
            cams[camera_id] = Camera(camera_id, model_id, params, width=width, height=height)
    return cams


def write_cameras_binary(cameras, path):
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(cameras)))
        for cid, c in cameras.items():
            f.write(struct.pack("<iiQQ", cid, c.ModelId(), c.width, c.height))
            f.write(struct.pack("<%dd" % len(c.params), *c.params))


def read_cameras_text(path):
    cams = {}
    with open(path) as f:
        for line in f:
            line = line.strip()
            if not line or line[0] == "#":
                continue
            e = line.split()
            cid, model, width, height = int(e[0]), e[1], int(e[2]), int(e[3])
            if model not in capi.CAMERA_MODELS:
                raise ValueError("Camera model does not exist")
            cams[cid] = Camera(cid, model, [float(x) for x in e[4:]], width=width, height=height)
    return cams


def write_cameras_text(cameras, path):
    with open(path, "w") as f:
        f.write("# Camera list with one line of data per camera:\n#   CAMERA_ID, MODEL, WIDTH, HEIGHT, PARAMS[]\n")
        f.write("# Number of cameras: %d\n" % len(cameras))
        for cid, c in cameras.items():
            f.write(" ".join([str(cid), MODEL_NAMES[c.ModelId()], str(c.width), str(c.height)] + [_g17(p) for p in c.params]) + "\n")


# ---------------------------------------------------------------- images
def read_images_binary(path):
    images = {}
    with open(path, "rb") as f:
        (n,) = struct.unpack("<Q", f.read(8))
        for _ in range(n):
            rec = struct.unpack("<I7dI", f.read(64))
            name = bytearray()
            while True:
                ch = f.read(1)
                if ch in (b"\x00", b""):
                    break
                name += ch
            img = Image(rec[0], rec[8], rec[1:5], rec[5:8], name=name.decode("utf-8"))
            (m,) = struct.unpack("<Q", f.read(8))
            raw = np.frombuffer(f.read(24 * m), dtype=np.dtype([("x", "<f8"), ("y", "<f8"), ("id", "<u8")]))
            for x, y, pid in zip(raw["x"].tolist(), raw["y"].tolist(), raw["id"].tolist()):
                img.points2D.append(Point2D((x, y), kInvalidPoint3DId if pid == _INVALID_U64 else pid))
            images[img.image_id] = img
    return images


def write_images_binary(images, path):
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(images)))
        for iid, img in images.items():
            f.write(struct.pack("<I7dI", iid, *img.qvec, *img.tvec, img.camera_id))
            f.write(img.name.encode("utf-8") + b"\x00")
            f.write(struct.pack("<Q", len(img.points2D)))
            for p in img.points2D:
                f.write(struct.pack("<ddQ", p.xy[0], p.xy[1], p.point3D_id if p.HasPoint3D() else _INVALID_U64))


def read_images_text(path):
    images = {}
    with open(path) as f:
        lines = [ln.rstrip("\n") for ln in f]
    k = 0
    while k < len(lines):
        line = lines[k].strip()
        k += 1
        if not line or line[0] == "#":
            continue
        e = line.split()
        img = Image(int(e[0]), int(e[8]), [float(x) for x in e[1:5]], [float(x) for x in e[5:8]], name=" ".join(e[9:]))
        pts = lines[k].split() if k < len(lines) else []  # the second line may be empty: an image without points
        k += 1
        for j in range(0, len(pts) - 2, 3):
            pid = int(pts[j + 2])
            img.points2D.append(Point2D((float(pts[j]), float(pts[j + 1])), kInvalidPoint3DId if pid == -1 else pid))
        images[img.image_id] = img
    return images


def write_images_text(images, path):
    n_obs = sum(sum(1 for p in img.points2D if p.HasPoint3D()) for img in images.values())
    with open(path, "w") as f:
        f.write("# Image list with two lines of data per image:\n#   IMAGE_ID, QW, QX, QY, QZ, TX, TY, TZ, CAMERA_ID, NAME\n"
                "#   POINTS2D[] as (X, Y, POINT3D_ID)\n")
        f.write("# Number of images: %d, mean observations per image: %s\n" % (len(images), _g17(n_obs / max(len(images), 1))))
        for iid, img in images.items():
            q = np.asarray(img.qvec, float)
            nq = np.linalg.norm(q)
            q = q / nq if nq > 0 else np.array([1.0, 0, 0, 0])  # NormalizeQuaternion at write, reconstruction.cc:1924
            f.write(" ".join([str(iid)] + [_g17(v) for v in q] + [_g17(v) for v in img.tvec] + [str(img.camera_id), img.name]) + "\n")
            f.write(" ".join("%s %s %d" % (_g17(p.xy[0]), _g17(p.xy[1]), p.point3D_id if p.HasPoint3D() else -1)
                             for p in img.points2D) + "\n")


# ---------------------------------------------------------------- points
def read_points3D_binary(path, rec):
    with open(path, "rb") as f:
        (n,) = struct.unpack("<Q", f.read(8))
        for _ in range(n):
            pid, x, y, z, r, g, b, err, m = struct.unpack("<Q3d3BdQ", f.read(51))
            track = np.frombuffer(f.read(8 * m), dtype="<u4").reshape(-1, 2)
            rec.AddPoint3DWithId(pid, (x, y, z), [(int(a), int(c)) for a, c in track], color=(r, g, b), error=err)


def write_points3D_binary(points3D, path):
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(points3D)))
        for pid, p in points3D.items():
            f.write(struct.pack("<Q3d3BdQ", pid, *p.xyz, *p.color, p.error, p.track.Length()))
            for el in p.track.Elements():
                f.write(struct.pack("<II", el.image_id, el.point2D_idx))


def read_points3D_text(path, rec):
    with open(path) as f:
        for line in f:
            line = line.strip()
            if not line or line[0] == "#":
                continue
            e = line.split()
            tr = [(int(e[j]), int(e[j + 1])) for j in range(8, len(e) - 1, 2)]
            rec.AddPoint3DWithId(int(e[0]), [float(x) for x in e[1:4]], tr, color=tuple(int(x) for x in e[4:7]), error=float(e[7]))


def write_points3D_text(points3D, path):
    n_el = sum(p.track.Length() for p in points3D.values())
    with open(path, "w") as f:
        f.write("# 3D point list with one line of data per point:\n#   POINT3D_ID, X, Y, Z, R, G, B, ERROR, TRACK[] as (IMAGE_ID, POINT2D_IDX)\n")
        f.write("# Number of points: %d, mean track length: %s\n" % (len(points3D), _g17(n_el / max(len(points3D), 1))))
        for pid, p in points3D.items():
            f.write(" ".join([str(pid)] + [_g17(v) for v in p.xyz] + [str(int(c)) for c in p.color] + [_g17(p.error)] +
                             ["%d %d" % (el.image_id, el.point2D_idx) for el in p.track.Elements()]) + "\n")


# ---------------------------------------------------------------- whole model
def detect_model_format(path):
    for ext in (".bin", ".txt"):  # Reconstruction::Read prefers binary, reconstruction.cc:733-747
        if all(os.path.isfile(os.path.join(path, n + ext)) for n in ("cameras", "images", "points3D")):
            return ext
    raise FileNotFoundError("cameras, images, points3D files do not exist at " + path)


def read_model(path, ext=None):
    """Reconstruction::Read (reconstruction.cc:733-747).  Point2D::point3D_id of the image files is kept as stored; the
    points' tracks are read from the points file (the two agree in a model the reference wrote)."""
    ext = ext or detect_model_format(path)
    rec = Reconstruction()
    binary = ext == ".bin"
    cams = (read_cameras_binary if binary else read_cameras_text)(os.path.join(path, "cameras" + ext))
    for c in cams.values():
        rec.AddCamera(c)
    imgs = (read_images_binary if binary else read_images_text)(os.path.join(path, "images" + ext))
    for im in imgs.values():
        rec.AddImage(im)
    (read_points3D_binary if binary else read_points3D_text)(os.path.join(path, "points3D" + ext), rec)
    return rec


def write_model(rec, path, ext=".bin"):
    """Reconstruction::Write{Binary,Text} (reconstruction.cc:749-771)."""
    os.makedirs(path, exist_ok=True)
    binary = ext == ".bin"
    (write_cameras_binary if binary else write_cameras_text)(rec.cameras, os.path.join(path, "cameras" + ext))
    (write_images_binary if binary else write_images_text)(rec.images, os.path.join(path, "images" + ext))
    (write_points3D_binary if binary else write_points3D_text)(rec.points3D, os.path.join(path, "points3D" + ext))


# ---------------------------------------------------------------- float32 TIFF maps
_TIFF_TYPES = {1: "B", 2: "c", 3: "H", 4: "I", 5: "II", 16: "Q"}


def read_tiff_f32(path):
    """Baseline TIFF, one 32-bit IEEE-float sample per pixel, uncompressed strips (what the reference's data sets hold and
    what its loader accepts: bpp must be 32, matrix_vis.h:146-150).  Returns float32 [height, width], top row first."""
    with open(path, "rb") as f:
        buf = f.read()
    if buf[:2] == b"II":
        en = "<"
    elif buf[:2] == b"MM":
        en = ">"
    else:
        raise ValueError("not a TIFF file: " + path)
    magic, ifd = struct.unpack(en + "HI", buf[2:8])
    if magic != 42:
        raise ValueError("not a classic TIFF file: " + path)
    (n_tags,) = struct.unpack(en + "H", buf[ifd:ifd + 2])
    tags = {}
    for k in range(n_tags):
        tag, typ, cnt, raw = struct.unpack(en + "HHI4s", buf[ifd + 2 + 12 * k: ifd + 14 + 12 * k])
        fmt = _TIFF_TYPES.get(typ)
        if fmt is None or typ == 5:
            continue
        size = struct.calcsize(fmt) * cnt
        data = raw[:size] if size <= 4 else buf[struct.unpack(en + "I", raw)[0]: struct.unpack(en + "I", raw)[0] + size]
        tags[tag] = struct.unpack(en + "%d%s" % (cnt, fmt), data)
    width, height = tags[256][0], tags[257][0]
    bps, spp = tags.get(258, (1,))[0], tags.get(277, (1,))[0]
    if bps * spp != 32:
        raise ValueError("Error loading depth map: %d bits per pixel, float32 maps need 32" % (bps * spp))
    if tags.get(259, (1,))[0] != 1:
        raise ValueError("compressed TIFF maps are not supported")
    if tags.get(339, (1,))[0] != 3:
        raise ValueError("TIFF map samples are not IEEE floats")
    offsets, counts = tags[273], tags.get(279) or (width * height * 4,)
    data = b"".join(buf[o:o + c] for o, c in zip(offsets, counts))
    return np.frombuffer(data, dtype=en + "f4", count=width * height).reshape(height, width).astype(np.float32)


def write_tiff_f32(path, matrix):
    """One uncompressed strip, little endian; the inverse of read_tiff_f32."""
    m = np.ascontiguousarray(matrix, dtype="<f4")
    h, w = m.shape
    entries = [(256, 4, 1, w), (257, 4, 1, h), (258, 3, 1, 32), (259, 3, 1, 1), (262, 3, 1, 1), (273, 4, 1, 8),
               (277, 3, 1, 1), (278, 4, 1, h), (279, 4, 1, 4 * w * h), (339, 3, 1, 3)]
    ifd_off = 8 + 4 * w * h
    with open(path, "wb") as f:
        f.write(b"II" + struct.pack("<HI", 42, ifd_off))
        f.write(m.tobytes())
        f.write(struct.pack("<H", len(entries)))
        for tag, typ, cnt, val in entries:
            f.write(struct.pack("<HHI", tag, typ, cnt) + (struct.pack("<HH", val, 0) if typ == 3 else struct.pack("<I", val)))
        f.write(struct.pack("<I", 0))


def map_paths(data_path, image_name):
    """<data_path>/{depth_tiff,semantic_tiff}/<image stem>_{depth,semantic}.tiff
    (/root/reference/src/optim/semantic_bundle_adjustment.cc:290-300)."""
    stem = os.path.splitext(os.path.basename(image_name))[0]
    return (os.path.join(data_path, "depth_tiff", stem + "_depth.tiff"),
            os.path.join(data_path, "semantic_tiff", stem + "_semantic.tiff"))


# ---------------------------------------------------------------- cylinders
def read_cylinders(path):
    """Returns an [n, 9] array: qvec(4) tvec(3) radius height.  A malformed line raises, as the reference's constructor
    does (cylinder.h:301-328)."""
    out = []
    with open(path) as f:
        for line in f:
            e = line.split()
            if len(e) != 13 or e[0] != "q" or e[5] != "t" or e[9] != "r" or e[11] != "h":
                raise RuntimeError("ERROR: creating Cylinder from string failed.")
            out.append([float(x) for x in e[1:5]] + [float(x) for x in e[6:9]] + [float(e[10]), float(e[12])])
    return np.asarray(out, dtype=np.float64).reshape(-1, 9)


def write_cylinders(path, cylinders):
    """std::ostream default formatting (6 significant digits), cylinder.h:290-298."""
    with open(path, "w") as f:
        for c in np.asarray(cylinders, dtype=np.float64).reshape(-1, 9):
            f.write("q %s t %s r %s h %s\n" % (" ".join("%g" % v for v in c[0:4]), " ".join("%g" % v for v in c[4:7]),
                                              "%g" % c[7], "%g" % c[8]))
