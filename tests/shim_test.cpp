// shim_test.cpp -- the scenarios of the reference's own BundleAdjuster test
//   /root/reference/src/optim/bundle_adjustment_test.cc:111-645
// replayed through the C++ shim (shim/sba_bundle_adjuster.h) over the mock Reconstruction.
//   ./shim_test counts   host only: flatten + sba_problem_counts (needs no device)
//   ./shim_test solve    also runs BundleAdjuster::Solve on cuda:0 and checks changed / unchanged blocks
// Built and run by tests/test_shim_cpp.py.  Exit code 0 == every check passed.
#include <cstdio>
#include <cstring>
#include <random>

#include "mock_reconstruction.h"
#include "sba_bundle_adjuster.h"

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

int main(int argc, char** argv) {
  const bool solve = argc > 1 && std::strcmp(argv[1], "solve") == 0;
  TestCounts();
  if (solve) {
    if (sba_device_count() < 1) {
      std::printf("no CUDA device: the shim has no CPU solve\n");
      return 3;
    }
    TestSolve();
  }
  std::printf("%s (%d failed)\n", g_failed ? "FAILED" : "OK", g_failed);
  return g_failed ? 1 : 0;
}
