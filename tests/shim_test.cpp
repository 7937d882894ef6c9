// shim_test.cpp -- the scenarios of the reference's own BundleAdjuster test
//   /root/reference/src/optim/bundle_adjustment_test.cc:111-645
// replayed through the C++ shim (shim/sba_bundle_adjuster.h) over the mock Reconstruction.
//   ./shim_test counts   host only: flatten + sba_problem_counts (needs no device)
//   ./shim_test solve    also runs BundleAdjuster::Solve on cuda:0 and checks changed / unchanged blocks
// Built and run by tests/test_shim_cpp.py.  Exit code 0 == every check passed.
#include <cstdio>
#include <cstring>
#include <map>
#include <random>

#include "mock_reconstruction.h"
#include "sba_bundle_adjuster.h"
#include "sba_mapper_passes.h"
#include "sba_rig_bundle_adjuster.h"
#include "sba_semantic_adjusters.h"

using namespace sba_mock;
using sba_shim::BundleAdjustmentConfig;
using sba_shim::BundleAdjustmentOptions;
using BundleAdjuster = sba_shim::BundleAdjusterT<Reconstruction>;

static int g_failed = 0;
#define CHECK_THAT(cond)                                                        \
  do {                                                                          \
    if (!(cond)) {                                                              \
      ++g_failed;                                                               \
      std::printf("FAILED %s:%d  %s\n", __FILE__, __LINE__, #cond);             \
    }                                                                           \
  } while (0)

// GenerateReconstruction, bundle_adjustment_test.cc:123-184: SIMPLE_RADIAL f=1.2*1000 on a 1000x1000 image,
// identity rotation, t = (U[-1,1], U[-1,1], 10), points U[-1,1]^3 seen by every image, pixel noise U[-2,2].
static void Generate(size_t num_images, size_t num_points, Reconstruction* rec, unsigned seed = 0) {
  std::mt19937_64 gen(seed);
  std::uniform_real_distribution<double> unit(-1.0, 1.0), noise(-2.0, 2.0);
  std::vector<Vector3d> xyz(num_points);
  for (auto& p : xyz) p = {{unit(gen), unit(gen), unit(gen)}};
  for (size_t i = 0; i < num_images; ++i) {
    Camera cam;
    cam.SetCameraId(static_cast<camera_t>(i));
    cam.SetModelId(SBA_SIMPLE_RADIAL);
    cam.Params() = {1200.0, 500.0, 500.0, 0.0};
    rec->AddCamera(cam);
    Image im;
    im.SetImageId(static_cast<image_t>(i));
    im.SetCameraId(static_cast<camera_t>(i));
    im.Tvec() = {{unit(gen), unit(gen), 10.0}};
    std::vector<Vector2d> xy(num_points);
    for (size_t j = 0; j < num_points; ++j) {
      const double z = xyz[j][2] + im.Tvec()[2];
      xy[j] = {{1200.0 * (xyz[j][0] + im.Tvec()[0]) / z + 500.0 + noise(gen),
                1200.0 * (xyz[j][1] + im.Tvec()[1]) / z + 500.0 + noise(gen)}};
    }
    im.SetPoints2D(xy);
    rec->AddImage(im);
  }
  for (size_t j = 0; j < num_points; ++j) {
    Track tr;
    for (size_t i = 0; i < num_images; ++i) tr.AddElement(static_cast<image_t>(i), static_cast<uint32_t>(j));
    rec->AddPoint3D(xyz[j], tr);
  }
}

// The residual order of the reference (SetUp, bundle_adjustment.cc:326-478), restated with ordered maps and no shortcuts:
// images of the config in set order, their features with a 3-D point in feature order; then the config's variable and
// constant points, each adding its track elements from images outside the config.  Slots are handed out in first-seen
// order.  Flatten -- which maps point ids through a direct table or a hash map and walks the point container for the
// packing -- must produce the same arrays on random scenes: partial configs, features without points, deleted
// observations, points of the config, both id maps.
static void TestFlattenOrder() {
  for (int seed = 0; seed < 60; ++seed) {
    std::mt19937_64 rng(static_cast<unsigned>(seed));
    const int n_img = 3 + seed % 7, n_pt = 20 + (seed * 13) % 150;
    Reconstruction rec;
    for (camera_t c = 1; c <= 2; ++c) {
      Camera cam; cam.SetCameraId(c); cam.SetModelId(c == 1 ? SBA_SIMPLE_RADIAL : SBA_SIMPLE_PINHOLE);
      for (auto& v : cam.Params()) v = 100.0 + static_cast<double>(rng() % 100);
      rec.AddCamera(cam);
    }
    std::vector<std::vector<Vector2d>> xy(static_cast<size_t>(n_img));
    std::vector<Track> tracks(static_cast<size_t>(n_pt));
    for (auto& tr : tracks) {
      const int len = 2 + static_cast<int>(rng() % 3), first = static_cast<int>(rng() % static_cast<unsigned>(n_img));
      for (int k = 0; k < len && k < n_img; ++k) {
        const size_t i = static_cast<size_t>((first + k) % n_img);
        tr.AddElement(static_cast<image_t>(i + 1), static_cast<uint32_t>(xy[i].size()));
        xy[i].push_back({{static_cast<double>(rng() % 640), static_cast<double>(rng() % 480)}});
      }
    }
    for (int i = 0; i < n_img; ++i) {
      for (int e = 0; e < 3; ++e) xy[static_cast<size_t>(i)].push_back({{1.0, 2.0}});  // features without a 3-D point
      Image im; im.SetImageId(static_cast<image_t>(i + 1)); im.SetCameraId(static_cast<camera_t>(1 + i % 2)); im.SetPoints2D(xy[static_cast<size_t>(i)]);
      im.Qvec() = {{1.0, 0.0, 0.0, 0.0}}; im.Tvec() = {{0.1 * i, 0.2, 0.3}};
      rec.AddImage(im);
    }
    std::vector<point3D_t> ids;
    for (int p = 0; p < n_pt; ++p) ids.push_back(rec.AddPoint3D({{static_cast<double>(p), 1.0 + seed, 2.0}}, tracks[static_cast<size_t>(p)]));
    for (int k = 0; k < 3; ++k) {  // some features lose their point
      const image_t i = static_cast<image_t>(1 + rng() % static_cast<unsigned>(n_img));
      for (uint32_t j = 0; j < rec.Image(i).Points2D().size(); ++j)
        if (rec.Image(i).Point2D(j).HasPoint3D() && rec.Point3D(rec.Image(i).Point2D(j).Point3DId()).Track().Length() > 2) { rec.DeleteObservation(i, j); break; }
    }
    BundleAdjustmentConfig config;
    const int n_cfg = seed % 3 == 0 ? n_img : 1 + static_cast<int>(rng() % static_cast<unsigned>(n_img));  // global or local
    for (int i = 0; i < n_cfg; ++i) config.AddImage(static_cast<image_t>(1 + (i * 2) % n_img));
    for (int k = 0; k < 5; ++k) {
      const point3D_t id = ids[rng() % ids.size()];
      if (!config.HasPoint(id)) { if (k % 2) config.AddVariablePoint(id); else config.AddConstantPoint(id); }
    }
    // ---- the plain restatement ----
    std::vector<int32_t> want_img, want_pt;
    std::vector<double> want_xy;
    std::vector<image_t> img_ids;
    std::vector<point3D_t> pt_ids;
    std::map<image_t, int32_t> img_slot;
    std::map<point3D_t, int32_t> pt_slot;
    std::map<point3D_t, size_t> seen;
    auto slot_of = [](auto& m, auto& order, auto id) { auto it = m.find(id); if (it != m.end()) return it->second; const int32_t s = static_cast<int32_t>(order.size()); m[id] = s; order.push_back(id); return s; };
    for (image_t id : config.Images())
      for (const auto& p2 : rec.Image(id).Points2D()) {
        if (!p2.HasPoint3D()) continue;
        want_img.push_back(slot_of(img_slot, img_ids, id));
        want_pt.push_back(slot_of(pt_slot, pt_ids, p2.Point3DId()));
        want_xy.push_back(p2.XY()[0]); want_xy.push_back(p2.XY()[1]);
        seen[p2.Point3DId()] += 1;
      }
    auto add_point = [&](point3D_t pid) {
      const auto& p3 = rec.Point3D(pid);
      if (seen[pid] == p3.Track().Length()) return;
      for (const auto& el : p3.Track().Elements()) {
        if (config.HasImage(el.image_id)) continue;
        seen[pid] += 1;
        want_img.push_back(slot_of(img_slot, img_ids, el.image_id));
        want_pt.push_back(slot_of(pt_slot, pt_ids, pid));
        const auto& x = rec.Image(el.image_id).Point2D(el.point2D_idx).XY();
        want_xy.push_back(x[0]); want_xy.push_back(x[1]);
      }
    };
    for (point3D_t id : config.VariablePoints()) add_point(id);
    for (point3D_t id : config.ConstantPoints()) add_point(id);
    for (int use_hash = 0; use_hash < 2; ++use_hash) {
      sba_shim::PointSlots::ForceHash() = use_hash != 0;
      BundleAdjuster ba(BundleAdjustmentOptions(), config);
      const sba_shim::FlatProblem& f = ba.Flatten(&rec);
      CHECK_THAT(f.obs_image == want_img && f.obs_point == want_pt && f.obs_xy == want_xy);
      CHECK_THAT(f.image_ids == img_ids && f.point_ids == pt_ids);
      bool packed = f.points.size() == 3 * pt_ids.size() && f.point_constant.size() == pt_ids.size();
      for (size_t s = 0; packed && s < pt_ids.size(); ++s) {
        const auto& p3 = rec.Point3D(pt_ids[s]);
        packed = f.points[3 * s] == p3.XYZ()[0] && f.points[3 * s + 1] == p3.XYZ()[1] && f.points[3 * s + 2] == p3.XYZ()[2] &&
                 f.point_constant[s] == ((p3.Track().Length() > seen[pt_ids[s]] || config.HasConstantPoint(pt_ids[s])) ? 1 : 0);
      }
      CHECK_THAT(packed);
    }
    sba_shim::PointSlots::ForceHash() = false;
  }
}

static BundleAdjustmentConfig Views(size_t n, bool both_constant) {
  BundleAdjustmentConfig c;
  for (size_t i = 0; i < n; ++i) c.AddImage(static_cast<image_t>(i));
  c.SetConstantPose(0);
  if (both_constant) c.SetConstantPose(1); else c.SetConstantTvec(1, {0});
  return c;
}

struct Counts { int residuals, parameters; };
static Counts CountsOf(const BundleAdjustmentOptions& o, const BundleAdjustmentConfig& c, Reconstruction* rec,
                       sba_shim::FlatProblem* keep = nullptr) {
  BundleAdjuster ba(o, c);
  const sba_shim::FlatProblem& flat = ba.Flatten(rec);
  const sba_problem_desc d = flat.Desc();
  sba_summary s{};
  CHECK_THAT(sba_problem_counts(&d, &s) == SBA_OK);
  if (keep) *keep = flat;
  return {s.num_residuals_reduced, s.num_effective_parameters_reduced};
}

static void TestCounts() {
  BundleAdjustmentOptions opt;
  opt.print_summary = false;
  {  // TestConfigNumObservations :186-208
    Reconstruction rec; Generate(4, 100, &rec);
    BundleAdjustmentConfig c;
    c.AddImage(0); c.AddImage(1);
    CHECK_THAT(c.NumResiduals(rec) == 400);
    c.AddVariablePoint(1); CHECK_THAT(c.NumResiduals(rec) == 404);
    c.AddConstantPoint(2); CHECK_THAT(c.NumResiduals(rec) == 408);
    c.AddImage(2); CHECK_THAT(c.NumResiduals(rec) == 604);
    c.AddImage(3); CHECK_THAT(c.NumResiduals(rec) == 800);
  }
  {  // TestTwoView :210-246
    Reconstruction rec; Generate(2, 100, &rec);
    const Counts k = CountsOf(opt, Views(2, false), &rec);
    CHECK_THAT(k.residuals == 400 && k.parameters == 309);
  }
  {  // TestTwoViewConstantCamera :248-284
    Reconstruction rec; Generate(2, 100, &rec);
    BundleAdjustmentConfig c = Views(2, true);
    c.SetConstantCamera(0);
    const Counts k = CountsOf(opt, c, &rec);
    CHECK_THAT(k.residuals == 400 && k.parameters == 302);
  }
  {  // TestPartiallyContainedTracks :286-330
    Reconstruction rec; Generate(3, 100, &rec);
    rec.DeleteObservation(2, 0);
    sba_shim::FlatProblem flat;
    const Counts k = CountsOf(opt, Views(2, true), &rec, &flat);
    CHECK_THAT(k.residuals == 400 && k.parameters == 7);
    size_t n_const = 0;
    for (uint8_t v : flat.point_constant) n_const += v;
    CHECK_THAT(n_const == 99);
  }
  {  // TestPartiallyContainedTracksForceToOptimizePoint :332-388
    Reconstruction rec; Generate(3, 100, &rec);
    const point3D_t add_variable = rec.Image(2).Point2D(1).Point3DId();
    const point3D_t add_constant = rec.Image(2).Point2D(2).Point3DId();
    rec.DeleteObservation(2, 0);
    BundleAdjustmentConfig c = Views(2, true);
    c.AddVariablePoint(add_variable);
    c.AddConstantPoint(add_constant);
    sba_shim::FlatProblem flat;
    const Counts k = CountsOf(opt, c, &rec, &flat);
    CHECK_THAT(k.residuals == 402 && k.parameters == 10);
    // the camera of image 2 entered through AddPointToProblem and is forced constant (:456-459)
    CHECK_THAT(flat.camera_ids.size() == 3 && flat.cam_mask[2] == 0xF && flat.pose_constant[2] == 1);
  }
  {  // TestConstantPoints :390-435
    Reconstruction rec; Generate(2, 100, &rec);
    BundleAdjustmentConfig c = Views(2, true);
    c.AddConstantPoint(1); c.AddConstantPoint(2);
    const Counts k = CountsOf(opt, c, &rec);
    CHECK_THAT(k.residuals == 400 && k.parameters == 298);
  }
  {  // TestVariableImage :437-477
    Reconstruction rec; Generate(3, 100, &rec);
    const Counts k = CountsOf(opt, Views(3, false), &rec);
    CHECK_THAT(k.residuals == 600 && k.parameters == 317);
  }
  {  // TestConstantFocalLength :479-529, TestVariablePrincipalPoint :531-593, TestConstantExtraParam :595-645
    Reconstruction rec; Generate(2, 100, &rec);
    BundleAdjustmentOptions o = opt; o.refine_focal_length = false;
    CHECK_THAT(CountsOf(o, Views(2, false), &rec).parameters == 307);
    o = opt; o.refine_principal_point = true;
    CHECK_THAT(CountsOf(o, Views(2, false), &rec).parameters == 313);
    o = opt; o.refine_extra_params = false;
    CHECK_THAT(CountsOf(o, Views(2, false), &rec).parameters == 307);
  }
  {  // residual order == config.Images() iteration order, Points2D index inner (a7)
    Reconstruction rec; Generate(3, 5, &rec);
    BundleAdjustmentConfig c; c.AddImage(1); c.AddImage(0);
    BundleAdjuster ba(opt, c);
    const sba_shim::FlatProblem& flat = ba.Flatten(&rec);
    size_t k = 0;
    for (image_t id : c.Images())
      for (uint32_t j = 0; j < 5; ++j, ++k) {
        CHECK_THAT(flat.image_ids[static_cast<size_t>(flat.obs_image[k])] == id);
        CHECK_THAT(flat.point_ids[static_cast<size_t>(flat.obs_point[k])] == rec.Image(id).Point2D(j).Point3DId());
        CHECK_THAT(flat.obs_xy[2 * k] == rec.Image(id).Point2D(j).XY()[0]);
      }
    CHECK_THAT(k == flat.obs_image.size());
  }
  {  // error behaviour: Check(), empty problem -> false (:267-269), one-shot (:260), null (:259), unknown model
    BundleAdjustmentOptions bad = opt; bad.loss_function_scale = -1;
    bool threw = false;
    try { BundleAdjuster ba(bad, BundleAdjustmentConfig()); } catch (const std::invalid_argument&) { threw = true; }
    CHECK_THAT(threw);
    Reconstruction rec;
    BundleAdjuster ba(opt, BundleAdjustmentConfig());
    CHECK_THAT(ba.Solve(&rec) == false);
    threw = false;
    try { ba.Solve(&rec); } catch (const std::logic_error&) { threw = true; }
    CHECK_THAT(threw);
    threw = false;
    try { BundleAdjuster(opt, BundleAdjustmentConfig()).Solve(nullptr); } catch (const std::invalid_argument&) { threw = true; }
    CHECK_THAT(threw);
    threw = false;
    try { Camera c; c.SetModelId(99); } catch (const std::domain_error&) { threw = true; }
    CHECK_THAT(threw);
  }
}

struct StopAfterTwo : sba_shim::IterationCallback {
  int calls = 0;
  sba_shim::CallbackReturnType operator()(const sba_shim::IterationSummary& it) override {
    ++calls;
    return it.iteration >= 2 ? sba_shim::SOLVER_TERMINATE_SUCCESSFULLY : sba_shim::SOLVER_CONTINUE;
  }
};

static void TestSolve() {
  BundleAdjustmentOptions opt;
  opt.print_summary = false;
  {  // TestTwoView :210-246 end to end on the device
    Reconstruction rec; Generate(2, 100, &rec);
    const Reconstruction orig = rec;
    BundleAdjuster ba(opt, Views(2, false));
    CHECK_THAT(ba.Solve(&rec));
    const auto& s = ba.Summary();
    CHECK_THAT(s.num_residuals_reduced == 400 && s.num_effective_parameters_reduced == 309);
    CHECK_THAT(s.final_cost < s.initial_cost && s.IsSolutionUsable());
    for (camera_t c = 0; c < 2; ++c) {  // CheckVariableCamera :42-52: f, k change; principal point fixed
      CHECK_THAT(rec.Camera(c).Params()[0] != orig.Camera(c).Params()[0]);
      CHECK_THAT(rec.Camera(c).Params()[1] == orig.Camera(c).Params()[1]);
      CHECK_THAT(rec.Camera(c).Params()[2] == orig.Camera(c).Params()[2]);
      CHECK_THAT(rec.Camera(c).Params()[3] != orig.Camera(c).Params()[3]);
    }
    CHECK_THAT(rec.Image(0).Qvec() == orig.Image(0).Qvec() && rec.Image(0).Tvec() == orig.Image(0).Tvec());
    CHECK_THAT(rec.Image(1).Qvec() != orig.Image(1).Qvec() && rec.Image(1).Tvec() != orig.Image(1).Tvec());
    CHECK_THAT(rec.Image(1).Tvec()[0] == orig.Image(1).Tvec()[0]);
    for (const auto& kv : rec.Points3D()) CHECK_THAT(kv.second.XYZ() != orig.Point3D(kv.first).XYZ());
    std::printf("two_view: cost %.9g -> %.9g in %d iterations (solver %d)\n", s.initial_cost, s.final_cost,
                s.num_iterations, s.linear_solver_type_used);
  }
  {  // TestPartiallyContainedTracks :286-330: only the one fully contained point and the cameras move
    Reconstruction rec; Generate(3, 100, &rec);
    const point3D_t variable_id = rec.Image(2).Point2D(0).Point3DId();
    rec.DeleteObservation(2, 0);
    const Reconstruction orig = rec;
    BundleAdjuster ba(opt, Views(2, true));
    CHECK_THAT(ba.Solve(&rec));
    CHECK_THAT(ba.Summary().num_residuals_reduced == 400 && ba.Summary().num_effective_parameters_reduced == 7);
    for (image_t i = 0; i < 3; ++i) CHECK_THAT(rec.Image(i).Tvec() == orig.Image(i).Tvec());
    CHECK_THAT(rec.Camera(2).Params() == orig.Camera(2).Params());
    for (const auto& kv : rec.Points3D())
      CHECK_THAT((kv.second.XYZ() != orig.Point3D(kv.first).XYZ()) == (kv.first == variable_id));
  }
  {  // IterationCallback terminates the solve (src/controllers/bundle_adjustment.cc:43-61)
    Reconstruction rec; Generate(3, 100, &rec);
    StopAfterTwo cb;
    BundleAdjustmentOptions o = opt;
    o.solver_options.callbacks.push_back(&cb);
    o.solver_options.update_state_every_iteration = true;
    BundleAdjuster ba(o, Views(3, false));
    CHECK_THAT(ba.Solve(&rec));
    CHECK_THAT(ba.Summary().termination_type == SBA_USER_SUCCESS);
    CHECK_THAT(cb.calls == 3 && ba.Summary().iterations.size() == 3);
  }
}

// ---- the two semantic adjusters through their class shims (shim/sba_semantic_adjusters.h) ----------------------
using SemanticBundleAdjuster = sba_shim::SemanticBundleAdjusterT<Reconstruction>;
using GeometricSemanticBundleAdjuster = sba_shim::GeometricSemanticBundleAdjusterT<Reconstruction>;

// n cameras with identity rotation at t = (tx, ty, 0.5 i) looking at the world plane Z = 5 carrying a checker of
// labels; the depth of every pixel is then 5 + t_z, and the world point of pixel (x, y) is (u d - t_x, v d - t_y, 5).
static void SemanticScene(int n, int H, int W, Reconstruction* rec, std::vector<sba_shim::FloatMap>* depth,
                          std::vector<sba_shim::FloatMap>* label) {
  const double f = 1.2 * W, cx = W / 2.0, cy = H / 2.0;
  for (int i = 0; i < n; ++i) {
    Camera cam;
    cam.SetCameraId(static_cast<camera_t>(i)); cam.SetModelId(SBA_SIMPLE_PINHOLE); cam.Params() = {f, cx, cy};
    rec->AddCamera(cam);
    Image im;
    im.SetImageId(static_cast<image_t>(i)); im.SetCameraId(static_cast<camera_t>(i)); im.SetName("img" + std::to_string(i) + ".png");
    im.Tvec() = {{0.1 * i, -0.07 * i, 0.5 * i}};
    rec->AddImage(im);
    sba_shim::FloatMap d, l;
    d.height = l.height = H; d.width = l.width = W;
    d.data.assign(static_cast<size_t>(H) * W, static_cast<float>(5.0 + 0.5 * i)); l.data.resize(static_cast<size_t>(H) * W);
    for (int y = 0; y < H; ++y)
      for (int x = 0; x < W; ++x) {
        const double dd = 5.0 + 0.5 * i, X = (x - cx) / f * dd - 0.1 * i, Y = (y - cy) / f * dd + 0.07 * i;
        l.data[static_cast<size_t>(y) * W + x] = static_cast<float>((static_cast<long>(std::floor(X * 3.0)) + 3 * static_cast<long>(std::floor(Y * 3.0))) % 5 + 10);
      }
    depth->push_back(d); label->push_back(l);
  }
}

static BundleAdjustmentConfig SemanticViews(int n) {
  BundleAdjustmentConfig c;
  for (int i = 0; i < n; ++i) { c.AddImage(static_cast<image_t>(i)); c.SetConstantCamera(static_cast<camera_t>(i)); }
  c.SetConstantPose(0);
  return c;
}

// camera_rig_test.cc:288-330 (TestComputeAbsolutePose) on the mock's CameraRig -- the class the rig shim is tested with:
// two rig cameras at relative translations (0,1,2) and (3,4,5), images at (0,0,0) and (3,3,3) -> rig pose (1,0,0,0), (0,-1,-2).
static void TestMockCameraRig() {
  sba_mock::CameraRig rig;
  rig.AddCamera(0, {{1, 0, 0, 0}}, {{0, 1, 2}});
  rig.AddCamera(1, {{1, 0, 0, 0}}, {{3, 4, 5}});
  rig.AddSnapshot({0, 1});
  Reconstruction rec;
  for (camera_t c = 0; c < 2; ++c) { Camera cam; cam.SetCameraId(c); cam.SetModelId(SBA_PINHOLE); cam.Params() = {1, 1, 0.5, 0.5}; rec.AddCamera(cam); }
  Image a; a.SetImageId(0); a.SetCameraId(0); a.Tvec() = {{0, 0, 0}}; rec.AddImage(a);
  Image b; b.SetImageId(1); b.SetCameraId(1); b.Tvec() = {{3, 3, 3}}; rec.AddImage(b);
  rig.SetRefCameraId(0);
  rig.Check(rec);
  Vector4d q; Vector3d t;
  rig.ComputeAbsolutePose(0, rec, &q, &t);
  CHECK_THAT(std::fabs(std::fabs(q[0]) - 1.0) < 1e-12 && std::fabs(q[1]) < 1e-12 && std::fabs(q[2]) < 1e-12 && std::fabs(q[3]) < 1e-12);
  CHECK_THAT(t[0] == 0.0 && t[1] == -1.0 && t[2] == -2.0);
}

static void TestSemanticHost() {
  Reconstruction rec;
  std::vector<sba_shim::FloatMap> depth, label;
  SemanticScene(3, 48, 64, &rec, &depth, &label);
  auto throws = [&](const sba_shim::SemanticBundleAdjustmentOptions& o, const BundleAdjustmentConfig& c, Reconstruction* r) {
    try { SemanticBundleAdjuster ba(o, c); for (int i = 0; i < 3; ++i) ba.SetMaps(r->Image(i).Name(), depth[i], label[i]); ba.Solve(r); }
    catch (const std::runtime_error&) { return true; }
    return false;
  };
  sba_shim::SemanticBundleAdjustmentOptions o;
  CHECK_THAT(o.solver_options.function_tolerance == 1e-8 && o.solver_options.update_state_every_iteration);  // :116-127
  CHECK_THAT(o.error_computation_pixel_step == 10 && o.depth_error_threshold == 2 && o.numeric_relative_step_size == 1e-3);
  {  // Assert (semantic_bundle_adjustment.cc:604-644): intrinsics must be constant, SIMPLE_PINHOLE, refine_extrinsics
    BundleAdjustmentConfig c = SemanticViews(3);
    c.SetVariableCamera(1);
    CHECK_THAT(throws(o, c, &rec));
    Reconstruction other = rec;
    other.Camera(2).SetModelId(SBA_SIMPLE_RADIAL);
    CHECK_THAT(throws(o, SemanticViews(3), &other));
    sba_shim::SemanticBundleAdjustmentOptions o2 = o;
    o2.refine_extrinsics = false;
    CHECK_THAT(throws(o2, SemanticViews(3), &rec));
  }
  sba_shim::GeometricSemanticBundleAdjustmentOptions g;
  CHECK_THAT(g.trunk_semantic_class == 250. && g.refine_geometry && !g.include_landmark_error);
}

struct CountCalls : sba_shim::IterationCallback {
  int calls = 0;
  sba_shim::CallbackReturnType operator()(const sba_shim::IterationSummary&) override { ++calls; return sba_shim::SOLVER_CONTINUE; }
};

static void TestSemanticSolve() {
  {  // SemanticBundleAdjuster::Solve on the device: poses move, the constant one does not, the cost does not go up
    Reconstruction rec;
    std::vector<sba_shim::FloatMap> depth, label;
    SemanticScene(3, 96, 128, &rec, &depth, &label);
    rec.Image(1).Tvec()[0] += 0.04; rec.Image(2).Tvec()[1] -= 0.05;  // the error to remove
    const Reconstruction orig = rec;
    sba_shim::SemanticBundleAdjustmentOptions o;
    o.error_computation_pixel_step = 4;
    o.solver_options.max_num_iterations = 15;
    CountCalls cb;
    o.solver_options.callbacks.push_back(&cb);
    SemanticBundleAdjuster ba(o, SemanticViews(3));
    for (int i = 0; i < 3; ++i) ba.SetMaps(rec.Image(i).Name(), depth[i], label[i]);
    CHECK_THAT(ba.Solve(&rec));
    const auto& s = ba.Summary();
    CHECK_THAT(s.num_residual_blocks == 6 * 24 * 32 && s.num_effective_parameters_reduced == 12);
    CHECK_THAT(s.final_cost < s.initial_cost && s.num_successful_steps >= 1);
    CHECK_THAT(cb.calls == static_cast<int>(s.iterations.size()) && cb.calls >= 2);
    CHECK_THAT(rec.Image(0).Tvec() == orig.Image(0).Tvec() && rec.Image(0).Qvec() == orig.Image(0).Qvec());
    CHECK_THAT(rec.Image(1).Tvec() != orig.Image(1).Tvec());
    std::printf("semantic: cost %.9g -> %.9g in %d iterations\n", s.initial_cost, s.final_cost, s.num_iterations);
  }
  {  // GeometricSemanticBundleAdjuster::Solve: masks = the library's own rasterisation of the true cylinder
    Reconstruction rec;
    std::vector<sba_shim::FloatMap> depth, label;
    const int H = 96, W = 128;
    SemanticScene(3, H, W, &rec, &depth, &label);
    sba_shim::CylinderParams truth;
    const double ang = 1.5707963267948966;
    truth.v = {std::cos(ang / 2), std::sin(ang / 2), 0.0, 0.0, 0.1, 0.5, 5.0, 0.15, 1.2};  // axis towards -y, 5 m ahead
    std::vector<sba_shim::FloatMap> sem(3);
    for (int i = 0; i < 3; ++i) {
      sba_cylinder_desc d{};
      const double q[4] = {1, 0, 0, 0};
      const double* t = rec.Image(i).Tvec().data();
      const double cam[3] = {1.2 * W, W / 2.0, H / 2.0};
      const int32_t bi = 0, bc = 0;
      d.num_images = 1; d.height = H; d.width = W; d.image_qvec = q; d.image_tvec = t; d.camera_params = cam;
      d.num_cylinders = 1; d.cyl_qvec = truth.v.data(); d.cyl_tvec = truth.v.data() + 4; d.cyl_radius = truth.v.data() + 7;
      d.cyl_height = truth.v.data() + 8; d.num_blocks = 1; d.block_image = &bi; d.block_cylinder = &bc; d.relative_step_size = 1e-3;
      double pts[8]; int32_t ok = 0;
      CHECK_THAT(sba_cylinder_quadrilateral(&d, pts, &ok) == SBA_OK && ok == 1);
      std::vector<uint8_t> m(static_cast<size_t>(H) * W);
      CHECK_THAT(sba_draw_quadrilateral(pts, H, W, m.data()) == SBA_OK);
      sem[i].height = H; sem[i].width = W; sem[i].data.resize(m.size());
      for (size_t k = 0; k < m.size(); ++k) sem[i].data[k] = m[k] ? 250.f : 3.f;
    }
    sba_shim::GeometricSemanticBundleAdjustmentOptions o;
    o.solver_options.max_num_iterations = 10;
    GeometricSemanticBundleAdjuster ba(o, SemanticViews(3));
    for (int i = 0; i < 3; ++i) ba.SetSemanticMap(rec.Image(i).Name(), sem[i]);
    sba_shim::CylinderParams start = truth;
    start.v[4] += 0.05; start.v[7] *= 1.3;  // shifted, too thick
    ba.Cylinders().push_back(start);
    CHECK_THAT(ba.Solve(&rec));
    const auto& s = ba.Summary();
    CHECK_THAT(s.num_residual_blocks == 3 && s.num_effective_parameters_reduced == 2 * 6 + 8);
    CHECK_THAT(s.initial_cost > 0 && s.final_cost <= s.initial_cost && ba.Cylinders()[0].v[7] >= 0);
    std::printf("cylinder: cost %.9g -> %.9g in %d iterations\n", s.initial_cost, s.final_cost, s.num_iterations);
  }
}

// ---- mapper-side passes (shim/sba_mapper_passes.h) ------------------------------------------------------------------
static void TestMapperPasses() {
  {  // Normalize: the robust box of the projection centres gets extent 10 and its centroid the origin; projections stay
    Reconstruction rec; Generate(6, 50, &rec);
    const Reconstruction orig = rec;
    const double scale = sba_shim::Normalize(&rec);
    CHECK_THAT(scale > 0.0 && scale != 1.0);
    const auto& p0 = *orig.Points3D().begin();
    const Vector3d a = orig.Image(2).Tvec(), b = rec.Image(2).Tvec();
    const Vector3d X0 = p0.second.XYZ(), X1 = rec.Point3D(p0.first).XYZ();
    // identity rotations (Generate): camera frame = X + t; a similarity of the world scales it
    for (int j = 0; j < 3; ++j) CHECK_THAT(std::fabs((X1[j] + b[j]) - scale * (X0[j] + a[j])) < 1e-9 * (1.0 + std::fabs(scale * (X0[j] + a[j]))));
    Reconstruction one; Generate(1, 10, &one);
    const Reconstruction one_orig = one;
    CHECK_THAT(sba_shim::Normalize(&one) == 1.0 && one.Image(0).Tvec() == one_orig.Image(0).Tvec());
    bool threw = false;
    try { sba_shim::Normalize(&rec, 0.0); } catch (const std::invalid_argument&) { threw = true; }
    CHECK_THAT(threw);
  }
  {  // RefineAbsolutePose: image 1 of a generated scene, pose perturbed, points constant
    Reconstruction rec; Generate(2, 100, &rec);
    const Image& im = rec.Image(1);
    std::vector<Vector2d> xy; std::vector<Vector3d> X; std::vector<char> mask;
    for (const auto& p2 : im.Points2D()) { xy.push_back(p2.XY()); X.push_back(rec.Point3D(p2.Point3DId()).XYZ()); mask.push_back(1); }
    mask[3] = 0; xy[3][0] += 300.0;
    Vector4d q = im.Qvec(); Vector3d t = im.Tvec();
    const Vector3d t_true = t;
    t[0] += 0.05; t[1] -= 0.04; q[1] += 0.01;
    const Vector3d t_start = t;
    Camera cam = rec.Camera(im.CameraId());
    const std::vector<double> params0 = cam.Params();
    sba_shim::AbsolutePoseRefinementOptions o;
    o.print_summary = false;
    sba_summary s{};
    CHECK_THAT(sba_shim::RefineAbsolutePose(o, mask, xy, X, &q, &t, &cam, &s));
    CHECK_THAT(s.num_residuals_reduced == 198 && s.num_effective_parameters_reduced == 6 + 2);
    CHECK_THAT(s.final_cost < 0.5 * s.initial_cost);  // (the observations carry U[-2, 2] px of noise: the floor is ~56)
    double d0 = 0.0, d1 = 0.0;
    for (int j = 0; j < 2; ++j) { d0 += (t_start[j] - t_true[j]) * (t_start[j] - t_true[j]); d1 += (t[j] - t_true[j]) * (t[j] - t_true[j]); }
    CHECK_THAT(d1 < d0);
    CHECK_THAT(cam.Params()[1] == params0[1] && cam.Params()[2] == params0[2]);  // principal point never refined
    std::printf("refine_absolute_pose: cost %.6g -> %.6g in %d iterations\n", s.initial_cost, s.final_cost, s.num_iterations);
    std::vector<char> none(mask.size(), 0);
    CHECK_THAT(sba_shim::RefineAbsolutePose(o, none, xy, X, &q, &t, &cam));
  }
}

// ---- RigBundleAdjuster (shim/sba_rig_bundle_adjuster.h): the reference's own rig tests, bundle_adjustment_test.cc:755-969 ----
using RigBundleAdjuster = sba_shim::RigBundleAdjusterT<Reconstruction, sba_mock::CameraRig>;
static void RigScene(size_t n_img, const std::vector<std::vector<image_t>>& snaps, Reconstruction* rec, std::vector<sba_mock::CameraRig>* rigs,
                     BundleAdjustmentConfig* config) {
  Generate(n_img, 100, rec);
  if (n_img > 2) { rec->Image(2).SetCameraId(0); rec->Image(3).SetCameraId(1); }
  rigs->emplace_back();
  (*rigs)[0].AddCamera(0, Vector4d{1, 0, 0, 0}, Vector3d{0, 0, 0});
  (*rigs)[0].AddCamera(1, Vector4d{1, 0, 0, 0}, Vector3d{0, 0, 0});
  for (const auto& s : snaps) (*rigs)[0].AddSnapshot(s);
  (*rigs)[0].SetRefCameraId(0);
  for (image_t i = 0; i < n_img; ++i) config->AddImage(i);
}
static void TestRigSolve() {
  BundleAdjustmentOptions opt;
  opt.print_summary = false;
  struct Case { size_t n; std::vector<std::vector<image_t>> snaps; bool refine; int res, eff; };
  const std::vector<Case> cases = {{2, {{0, 1}}, true, 400, 316}, {4, {{0, 1}, {2, 3}}, true, 800, 322},
                                   {4, {{0, 1}, {2, 3}}, false, 800, 316}, {4, {{0, 1}, {2}}, true, 800, 328}};
  for (const Case& c : cases) {
    Reconstruction rec; std::vector<sba_mock::CameraRig> rigs; BundleAdjustmentConfig config;
    RigScene(c.n, c.snaps, &rec, &rigs, &config);
    const Reconstruction orig = rec;
    const auto orig_rigs = rigs;
    sba_shim::RigBundleAdjusterOptions ro;
    ro.refine_relative_poses = c.refine;
    RigBundleAdjuster ba(opt, ro, config);
    CHECK_THAT(ba.Solve(&rec, &rigs));
    const auto& s = ba.Summary();
    CHECK_THAT(s.num_residuals_reduced == c.res && s.num_effective_parameters_reduced == c.eff);
    CHECK_THAT(s.final_cost < s.initial_cost);
    // CheckVariableCameraRig / CheckConstantCameraRig: the reference camera's relative pose never moves, the other
    // one only with refine_relative_poses
    CHECK_THAT(rigs[0].RelativeTvec(0) == orig_rigs[0].RelativeTvec(0) && rigs[0].RelativeQvec(0) == orig_rigs[0].RelativeQvec(0));
    CHECK_THAT((rigs[0].RelativeTvec(1) != orig_rigs[0].RelativeTvec(1)) == c.refine);
    CHECK_THAT(rec.Image(0).Tvec() != orig.Image(0).Tvec());
    for (const auto& kv : rec.Points3D()) CHECK_THAT(kv.second.XYZ() != orig.Point3D(kv.first).XYZ());
    std::printf("rig %zu views%s: cost %.6g -> %.6g, %d / %d\n", c.n, c.refine ? "" : " (constant relative poses)", s.initial_cost,
                s.final_cost, s.num_residuals_reduced, s.num_effective_parameters_reduced);
  }
}

int main(int argc, char** argv) {
  const bool solve = argc > 1 && std::strcmp(argv[1], "solve") == 0;
  TestCounts();
  TestFlattenOrder();
  TestMockCameraRig();
  TestSemanticHost();
  if (solve) {
    if (sba_device_count() < 1) {
      std::printf("no CUDA device: the shim has no CPU solve\n");
      return 3;
    }
    TestSolve();
    TestSemanticSolve();
    TestMapperPasses();
    TestRigSolve();
  }
  std::printf("%s (%d failed)\n", g_failed ? "FAILED" : "OK", g_failed);
  return g_failed ? 1 : 0;
}
