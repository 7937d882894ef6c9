// Host side of the drop-in at the headline size: how long the shim takes to walk a Reconstruction of C4's shape
// (1778 images / ~1 M points / ~5 M observations, every image in the config) into the arrays of sba_problem_desc, and to
// store the result back.  No device needed: only Flatten and the scatter loops run.
//   g++ -O2 -std=c++17 -Iinclude -Isemantic-bundle-adjustment-colmap_b200/shim tools/shim_flatten_bench.cpp
//       -Lsemantic-bundle-adjustment-colmap_b200 -lsba_b200 -Wl,-rpath,$PWD/semantic-bundle-adjustment-colmap_b200 -o /tmp/flatten_bench
#include <chrono>
#include <cstdio>
#include <algorithm>
#include <random>
#include <string>
#include <vector>

#include "mock_reconstruction.h"
#include "sba_bundle_adjuster.h"

using namespace sba_mock;
static double Now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main(int argc, char** argv) {
  const int n_img = argc > 1 ? atoi(argv[1]) : 1778;
  const int n_pt = argc > 2 ? atoi(argv[2]) : 993000;
  const double mean_track = 5.03;
  std::mt19937_64 rng(7);
  Reconstruction rec;
  Camera cam; cam.SetCameraId(1); cam.SetModelId(2); cam.Params() = {1000, 512, 384, 0.01};
  rec.AddCamera(cam);
  // observations per image first (Points2D are fixed-size per image), then the tracks
  std::vector<std::vector<Vector2d>> xy(n_img);
  std::vector<Track> tracks(n_pt);
  std::uniform_int_distribution<int> pick(0, n_img - 1);
  std::poisson_distribution<int> extra(mean_track - 2.0);
  size_t n_obs = 0;
  for (int p = 0; p < n_pt; ++p) {
    const int len = 2 + extra(rng);
    int first = pick(rng);
    for (int k = 0; k < len; ++k) {
      const int i = (first + k * 7) % n_img;  // distinct images for len < n_img / 7
      tracks[p].AddElement((image_t)(i + 1), (uint32_t)xy[i].size());
      xy[i].push_back({{(double)(rng() % 1024), (double)(rng() % 768)}});
      ++n_obs;
    }
  }
  for (int i = 0; i < n_img; ++i) {
    Image im; im.SetImageId((image_t)(i + 1)); im.SetCameraId(1); im.SetPoints2D(xy[i]);
    rec.AddImage(im);
  }
  for (int p = 0; p < n_pt; ++p) rec.AddPoint3D({{(double)p, 1.0, 2.0}}, tracks[p]);
  sba_shim::BundleAdjustmentConfig config;
  for (int i = 0; i < n_img; ++i) config.AddImage((image_t)(i + 1));
  config.SetConstantPose(1);
  config.SetConstantTvec(2, {0});
  sba_shim::BundleAdjustmentOptions options;
  const bool json = argc > 3 && std::string(argv[3]) == "--json";
  std::vector<double> ms;
  for (int rep = 0; rep < 3; ++rep) {
    sba_shim::BundleAdjusterT<Reconstruction> ba(options, config);
    const double t0 = Now();
    const auto& flat = ba.Flatten(&rec);
    const double t1 = Now();
    ms.push_back((t1 - t0) * 1e3);
    if (json) {
      if (rep == 2) {
        std::sort(ms.begin(), ms.end());
        std::printf("{\"flatten_ms\": %.2f, \"flatten_ms_min\": %.2f, \"flatten_ms_max\": %.2f, \"observations\": %zu, \"points\": %zu, \"images\": %zu}\n",
                    ms[1], ms[0], ms[2], flat.obs_image.size(), flat.point_ids.size(), flat.image_ids.size());
      }
      continue;
    }
    std::printf("Flatten: %zu observations, %zu points, %zu images in %.1f ms\n", flat.obs_image.size(), flat.point_ids.size(),
                flat.image_ids.size(), (t1 - t0) * 1e3);
    std::printf("         observations %.1f ms, extra track elements %.1f ms, parameter packing %.1f ms\n", ba.FlattenPhaseMs()[0],
                ba.FlattenPhaseMs()[1], ba.FlattenPhaseMs()[2]);
  }
  return 0;
}
