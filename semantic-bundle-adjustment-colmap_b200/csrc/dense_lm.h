// Levenberg-Marquardt on DENSE normal equations for the two semantic adjusters of the fork, whose problems have no
// point blocks to eliminate:
//   SemanticBundleAdjuster            /root/reference/src/optim/semantic_bundle_adjustment.cc:430-544, 646-906
//   GeometricSemanticBundleAdjuster   /root/reference/src/optim/geometric_semantic_bundle_adjustment.cc:481-604, 714-996
// Both build one-residual blocks with central numeric differentiation (Ceres NumericDiffCostFunction<CENTRAL>,
// SURVEY.md App. B.6) over camera poses (and cylinders), and call ceres::Solve with the solver policy of BundleAdjuster.
// With no e-blocks Ceres' Schur solvers reduce to a normal-equations solve on all blocks (App. B.1), so the step is
//   (S H S + D^2) y = S g,   delta = -S y,     H = J^T J, g = J^T r (loss-corrected, tangent space), S = Jacobi scaling
// with the same trust-region logic as csrc/lm.cu (App. B.3/B.4): radius, rho, accept / reject, diagonal reuse.
//
// Structure shared by both problems: ENTITIES (an image pose: quaternion + tvec; a cylinder: its parameter blocks) own
// a run of tangent columns; a residual GROUP is all residual blocks of one ordered entity pair (image 1 -> image 2;
// image x cylinder), or of one entity when the other side is constant.  The problem-specific kernels reduce a group's
// blocks to one small record (J^T J, J^T r, cost in the group's local tangent layout [entity 1 | entity 2]); the assembly
// below gathers the records into the dense H and g in a fixed order (no atomics, bit-reproducible).
#pragma once
#include <functional>
#include <vector>

#include "solver.h"

namespace sba {

constexpr int kGroupDim = 16;                                   // max local tangent dims of a group
constexpr int kGroupRec = kGroupDim * kGroupDim + kGroupDim + 2;  // J^T J (full 16 x 16), J^T r, cost, pad

struct DenseGroups {
  int n_ent = 0, n_groups = 0, n = 0;         // entities, groups, tangent dimension
  std::vector<int> ent_off, ent_dim;          // tangent run of every entity (dim 0: constant)
  std::vector<int> g_e1, g_e2;                // entities of a group (e2 = -1: none)
  DevBuf<int> d_ent_off, d_ent_dim, d_g_e1, d_g_e2;
  DevBuf<int> d_table;                        // [n_ent * n_ent] group of the ordered pair (e1, e2), [e * n_ent + e] unused; -1 none
  DevBuf<int> d_single;                       // [n_ent] group with e2 = -1 of entity e, or -1
  DevBuf<double> d_rec;                       // [n_groups * kGroupRec] filled by the problem's reduction kernel
  int Upload(cudaStream_t st);
};

// H (n x n row-major, lower part), g (n), cost <- group records
int AssembleDense(const DenseGroups& G, double* d_H, double* d_g, double* d_cost, cudaStream_t st, int64_t* launches);

struct DenseLMCallbacks {
  // current point: fills G.d_rec (all groups) -- called after every accepted step and once at the start
  std::function<int()> linearize;
  // candidate = Plus(current, delta) (delta: tangent, device); returns its cost and the squared norms the parameter
  // tolerance test needs (ambient step, ambient current point, variable blocks only)
  std::function<int(const double* d_delta, double* cost, double* step2, double* x2)> candidate;
  std::function<void()> accept;               // candidate becomes current
  sba_iteration_callback cb = nullptr;        // solver_options.callbacks
  void* user = nullptr;
  std::function<void()> before_callback;      // update_state_every_iteration: make the current point visible to cb
  // optional: H (n x n row-major, both triangles), g (n), cost of the current point, in place of the group records
  // (problems whose residual blocks span more than two entities: the rig solve); G then only carries n
  std::function<int(double* d_H, double* d_g, double* d_cost)> assemble;
};

// Runs the LM loop; summary counts (num_residuals ...) are the caller's.  trace receives the iteration table.
int DenseLMSolve(DenseGroups& G, const sba_solver_options& o, const DenseLMCallbacks& F, cudaStream_t st, sba_summary* sum,
                 std::vector<sba_iteration_summary>* trace);

// device helpers shared by the two problems ------------------------------------------------------------------------
#ifdef __CUDACC__
// tangent row of a quaternion block: J_local(3) = J_ambient(4) * PlusJacobian(q)   (QuaternionManifold, App. A.1)
__device__ __forceinline__ void QuatTangent(const double* q, const double* ja, double* jl) {
  const double q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3];
  jl[0] = -ja[0] * q1 + ja[1] * q0 - ja[2] * q3 + ja[3] * q2;
  jl[1] = -ja[0] * q2 + ja[1] * q3 + ja[2] * q0 - ja[3] * q1;
  jl[2] = -ja[0] * q3 - ja[1] * q2 + ja[2] * q1 + ja[3] * q0;
}
// QuaternionManifold::Plus: [cos|d|, sin|d|/|d| d] * x
__device__ __forceinline__ void QuatPlus(const double* x, const double* d, double* out) {
  const double n = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
  if (n == 0.0) { out[0] = x[0]; out[1] = x[1]; out[2] = x[2]; out[3] = x[3]; return; }
  const double s = sin(n) / n;
  const double z0 = cos(n), z1 = s * d[0], z2 = s * d[1], z3 = s * d[2];
  out[0] = z0 * x[0] - z1 * x[1] - z2 * x[2] - z3 * x[3];
  out[1] = z0 * x[1] + z1 * x[0] + z2 * x[3] - z3 * x[2];
  out[2] = z0 * x[2] - z1 * x[3] + z2 * x[0] + z3 * x[1];
  out[3] = z0 * x[3] + z1 * x[2] - z2 * x[1] + z3 * x[0];
}
#endif

}  // namespace sba
