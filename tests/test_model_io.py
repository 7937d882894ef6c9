"""On-disk formats either side of the path.  The golden model under tests/golden/model_{bin,txt}
was written by the reference's own scripts/python/read_write_model.py (tests/golden/make_model_golden.py):
  * reading it gives the values the generator put in; binary and text agree;
  * writing it back is BYTE-IDENTICAL for the binary files (Reconstruction::Write*Binary layout,
    /root/reference/src/base/reconstruction.cc:2012-2100) and value-identical for text, with the C++ writer's header
    lines and 17 significant digits (:1868-2010);
  * float32 TIFF maps and cylinders.txt round-trip (src/util/matrix_vis.h:130-176, src/util/cylinder.h:289-329)."""
import filecmp
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _same(a, b):
    assert list(a.cameras) == list(b.cameras) and list(a.images) == list(b.images) and list(a.points3D) == list(b.points3D)
    for k in a.cameras:
        ca, cb = a.cameras[k], b.cameras[k]
        assert (ca.ModelId(), ca.width, ca.height) == (cb.ModelId(), cb.width, cb.height)
        assert np.array_equal(ca.params, cb.params)
    for k in a.images:
        ia, ib = a.images[k], b.images[k]
        assert (ia.camera_id, ia.name, len(ia.points2D)) == (ib.camera_id, ib.name, len(ib.points2D))
        assert np.array_equal(ia.qvec, ib.qvec) and np.array_equal(ia.tvec, ib.tvec)
        for pa, pb in zip(ia.points2D, ib.points2D):
            assert np.array_equal(pa.xy, pb.xy) and pa.point3D_id == pb.point3D_id
    for k in a.points3D:
        pa, pb = a.points3D[k], b.points3D[k]
        assert np.array_equal(pa.xyz, pb.xyz) and pa.color == pb.color and pa.error == pb.error
        assert [(e.image_id, e.point2D_idx) for e in pa.track.Elements()] == [(e.image_id, e.point2D_idx) for e in pb.track.Elements()]


def test_golden_model_reads_as_generated(sba):
    rec = sba.model_io.read_model(os.path.join(GOLD, "model_bin"))
    assert list(rec.cameras) == [1, 2, 5] and list(rec.images) == [1, 2, 3, 7]
    assert list(rec.points3D) == [11, 12, 13, 20, 21, 4000000000]          # 64-bit point ids
    assert rec.Camera(5).ModelId() == sba.capi.CAMERA_MODELS["OPENCV"] and rec.Camera(5).NumParams() == 8
    assert rec.Camera(1).Params().tolist() == [655.123, 320.5, 240.25] and (rec.Camera(1).width, rec.Camera(1).height) == (640, 480)
    assert rec.Camera(2).Params()[0] == 1200.0 / 7
    assert rec.Image(1).name == "a/frame_000001.png" and rec.Image(2).name == "frame 2.jpg" and rec.Image(7).Points2D() == []
    assert abs(np.linalg.norm(rec.Image(3).Qvec()) - 1) < 1e-15 and rec.Image(3).CameraId() == 5
    assert [p.Point3DId() for p in rec.Image(1).Points2D()] == [11, 12, 21, 13]
    assert not rec.Image(3).Point2D(3).HasPoint3D() or rec.Image(3).Point2D(3).Point3DId() == 4000000000
    assert [(e.image_id, e.point2D_idx) for e in rec.Point3D(13).Track().Elements()] == [(1, 3), (2, 0), (3, 4)]
    assert all(0 <= c <= 255 for c in rec.Point3D(11).color) and 0.1 <= rec.Point3D(11).error <= 2.0


def test_binary_and_text_models_agree(sba):
    _same(sba.model_io.read_model(os.path.join(GOLD, "model_bin")), sba.model_io.read_model(os.path.join(GOLD, "model_txt")))
    assert sba.model_io.detect_model_format(os.path.join(GOLD, "model_bin")) == ".bin"
    with pytest.raises(FileNotFoundError):
        sba.model_io.detect_model_format(GOLD)


def test_binary_writer_is_byte_identical_to_the_reference(sba, tmp_path):
    rec = sba.model_io.read_model(os.path.join(GOLD, "model_bin"))
    sba.model_io.write_model(rec, str(tmp_path), ".bin")
    for name in ("cameras.bin", "images.bin", "points3D.bin"):
        assert filecmp.cmp(os.path.join(GOLD, "model_bin", name), str(tmp_path / name), shallow=False), name


def test_text_writer_round_trips_with_the_cpp_headers(sba, tmp_path):
    rec = sba.model_io.read_model(os.path.join(GOLD, "model_txt"))
    sba.model_io.write_model(rec, str(tmp_path), ".txt")
    _same(rec, sba.model_io.read_model(str(tmp_path), ".txt"))               # 17 significant digits: lossless
    head = open(tmp_path / "cameras.txt").read().splitlines()[:3]
    assert head == ["# Camera list with one line of data per camera:", "#   CAMERA_ID, MODEL, WIDTH, HEIGHT, PARAMS[]",
                    "# Number of cameras: 3"]
    assert open(tmp_path / "images.txt").read().splitlines()[3] == "# Number of images: 4, mean observations per image: 3"
    assert open(tmp_path / "points3D.txt").read().splitlines()[2].startswith("# Number of points: 6, mean track length: 2.1666666666666665")
    lines = [ln for ln in open(tmp_path / "images.txt").read().split("\n")]
    assert lines[-2] == "" and lines[-3].endswith(" 1 empty.png")             # an image without points keeps its empty line


def test_unknown_camera_model_is_rejected(sba, tmp_path):
    (tmp_path / "cameras.txt").write_text("1 NOT_A_MODEL 10 10 1 2 3\n")
    with pytest.raises(ValueError):
        sba.model_io.read_cameras_text(str(tmp_path / "cameras.txt"))


def test_float32_tiff_maps(sba, tmp_path):
    rng = np.random.default_rng(3)
    m = rng.normal(size=(48, 64)).astype(np.float32)
    m[0, 0], m[47, 63] = np.float32(7.25), np.float32(-1.5)
    p = str(tmp_path / "x_depth.tiff")
    sba.model_io.write_tiff_f32(p, m)
    back = sba.model_io.read_tiff_f32(p)
    assert back.dtype == np.float32 and back.shape == (48, 64) and np.array_equal(back, m)   # top row first
    # big-endian, two strips: the same image as another writer would store it
    import struct
    body = m.astype(">f4").tobytes()
    half = 24 * 64 * 4
    tags = [(256, 3, 1, 64), (257, 3, 1, 48), (258, 3, 1, 32), (259, 3, 1, 1), (277, 3, 1, 1), (278, 3, 1, 24), (339, 3, 1, 3)]
    ifd = 8 + len(body)
    extra = ifd + 2 + 12 * (len(tags) + 2) + 4
    with open(tmp_path / "be.tiff", "wb") as f:
        f.write(b"MM" + struct.pack(">HI", 42, ifd) + body + struct.pack(">H", len(tags) + 2))
        for tag, typ, cnt, val in tags:
            f.write(struct.pack(">HHIHH", tag, typ, cnt, val, 0))
        f.write(struct.pack(">HHII", 273, 4, 2, extra) + struct.pack(">HHII", 279, 4, 2, extra + 8) + struct.pack(">I", 0))
        f.write(struct.pack(">II", 8, 8 + half) + struct.pack(">II", half, half))
    assert np.array_equal(sba.model_io.read_tiff_f32(str(tmp_path / "be.tiff")), m)
    sba.model_io.write_tiff_f32(p, np.zeros((2, 2)))
    raw = bytearray(open(p, "rb").read())
    raw[raw.index(struct.pack("<HHI", 258, 3, 1)) + 8] = 16                 # 16 bits per sample: the reference refuses (bpp != 32)
    open(p, "wb").write(raw)
    with pytest.raises(ValueError):
        sba.model_io.read_tiff_f32(p)
    assert sba.model_io.map_paths("/data", "sub/img_0007.png") == ("/data/depth_tiff/img_0007_depth.tiff", "/data/semantic_tiff/img_0007_semantic.tiff")


def test_cylinders_text(sba, tmp_path):
    c = np.array([[1, 0, 0, 0, 0.5, -2, 3.25, 0.4, 2.0], [0.5, 0.5, -0.5, 0.5, 0, 0, 1e-3, 1.25, 10]])
    p = str(tmp_path / "cylinders.txt")
    sba.model_io.write_cylinders(p, c)
    assert open(p).read().splitlines()[0] == "q 1 0 0 0 t 0.5 -2 3.25 r 0.4 h 2"
    assert np.array_equal(sba.model_io.read_cylinders(p), c)
    (tmp_path / "bad.txt").write_text("q 1 0 0 0 t 0 0 0 r 1 H 2\n")
    with pytest.raises(RuntimeError):
        sba.model_io.read_cylinders(str(tmp_path / "bad.txt"))


@pytest.mark.gpu
def test_read_model_adjust_write_model(sba, tmp_path):
    """The step either side of the path: a model on disk -> BundleAdjuster on the device -> a model on disk."""
    prob, _ = sba.scene.make_scene(n_cam=6, n_pt=300, mean_track=4.0, seed=12)
    rec = sba.Reconstruction()
    mid = int(prob.camera_model[0])
    for c in range(len(prob.camera_model)):
        rec.AddCamera(sba.Camera(c + 1, mid, prob.camera_params[c, :sba.capi.MODEL_NUM_PARAMS[mid]], width=1000, height=1000))
    for i in range(len(prob.image_camera)):
        rec.AddImage(sba.Image(i + 1, int(prob.image_camera[i]) + 1, prob.image_qvec[i], prob.image_tvec[i], name="im%d.png" % i))
    tracks = {}
    for k in range(prob.num_observations):
        img = rec.Image(int(prob.obs_image[k]) + 1)
        img.points2D.append(sba.Point2D(prob.obs_xy[k]))
        tracks.setdefault(int(prob.obs_point[k]), []).append((img.image_id, len(img.points2D) - 1))
    for q in range(len(prob.points)):
        rec.AddPoint3D(prob.points[q], tracks.get(q, []))
    sba.model_io.write_model(rec, str(tmp_path / "in"), ".bin")
    loaded = sba.model_io.read_model(str(tmp_path / "in"))
    cfg = sba.BundleAdjustmentConfig()
    for iid in loaded.images:
        cfg.AddImage(iid)
    cfg.SetConstantPose(1)
    cfg.SetConstantTvec(2, [0])
    opt = sba.BundleAdjustmentOptions()
    opt.solver_options.max_num_iterations = 10
    ba = sba.BundleAdjuster(opt, cfg)
    assert ba.Solve(loaded)
    assert ba.Summary().final_cost < 0.5 * ba.Summary().initial_cost
    sba.model_io.write_model(loaded, str(tmp_path / "out"), ".txt")
    again = sba.model_io.read_model(str(tmp_path / "out"))
    assert any(not np.array_equal(again.Point3D(k).XYZ(), rec.Point3D(k).XYZ()) for k in rec.points3D)
    assert np.array_equal(again.Image(1).Tvec(), rec.Image(1).Tvec())       # the constant pose came back untouched
