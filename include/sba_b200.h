/*
 * sba_b200.h -- C ABI of the B200-native bundle-adjustment hot path.
 *
 * This is the drop-in boundary for the path the reference runs inside
 *   BundleAdjuster::Solve            /root/reference/src/optim/bundle_adjustment.cc:258-320
 * i.e. what the reference hands to Ceres: a set of per-observation reprojection
 * residual blocks (src/base/cost_functions.h:44-152) over pose / intrinsics /
 * point parameter blocks, manifolds (src/base/cost_functions.h:270-287), a loss
 * (src/optim/bundle_adjustment.cc:53-68) and `ceres::Solve`
 * (src/optim/bundle_adjustment.cc:306).
 *
 * The host side (the C++ shim in semantic-bundle-adjustment-colmap_b200/shim/,
 * mirroring BundleAdjuster / BundleAdjustmentConfig / BundleAdjustmentOptions)
 * flattens a Reconstruction in the reference's residual order
 * (src/optim/bundle_adjustment.cc:326-478) into the plain arrays below and calls
 * these entry points.  Plain pointers and sizes only; no torch / Eigen / Ceres
 * types cross this boundary.
 *
 * All functions return 0 on success and a non-zero sba_status otherwise; they
 * never abort.  sba_last_error() returns a thread-local message for the last
 * failure.  There is NO CPU fallback: every compute entry point fails with
 * SBA_ERR_NO_DEVICE when no CUDA device is usable.
 */
#ifndef SBA_B200_H_
#define SBA_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SBA_MAX_CAMERA_PARAMS 12 /* FULL_OPENCV / THIN_PRISM_FISHEYE, camera_models.h:280,349 */
#define SBA_POSE_TANGENT_DIM 6   /* quaternion manifold (3) + tvec (3) */

/* Camera model ids == CameraModel::kModelId, src/base/camera_models.h:189-349 */
enum sba_camera_model {
  SBA_SIMPLE_PINHOLE = 0,
  SBA_PINHOLE = 1,
  SBA_SIMPLE_RADIAL = 2,
  SBA_RADIAL = 3,
  SBA_OPENCV = 4,
  SBA_OPENCV_FISHEYE = 5,
  SBA_FULL_OPENCV = 6,
  SBA_FOV = 7,
  SBA_SIMPLE_RADIAL_FISHEYE = 8,
  SBA_RADIAL_FISHEYE = 9,
  SBA_THIN_PRISM_FISHEYE = 10,
  SBA_NUM_CAMERA_MODELS = 11
};

/* BundleAdjustmentOptions::LossFunctionType, src/optim/bundle_adjustment.h:51 */
enum sba_loss_type { SBA_LOSS_TRIVIAL = 0, SBA_LOSS_SOFT_L1 = 1, SBA_LOSS_CAUCHY = 2 };

/* Linear solver policy of BundleAdjuster::Solve, src/optim/bundle_adjustment.cc:271-286.
 * SPARSE_SCHUR sizes run the same device dense Cholesky as DENSE_SCHUR. */
enum sba_linear_solver {
  SBA_SOLVER_AUTO = -1, /* reference policy from num_config_images: <=50 dense, <=1000 sparse, else iterative */
  SBA_DENSE_SCHUR = 0,
  SBA_SPARSE_SCHUR = 1,
  SBA_ITERATIVE_SCHUR = 2
};

enum sba_status {
  SBA_OK = 0,
  SBA_ERR_INVALID_ARGUMENT = 1,
  SBA_ERR_NO_DEVICE = 2,
  SBA_ERR_CUDA = 3,
  SBA_ERR_UNKNOWN_CAMERA_MODEL = 4, /* std::domain_error in the reference, camera_models.h:140 */
  SBA_ERR_EMPTY_PROBLEM = 5,        /* Solve() returns false, bundle_adjustment.cc:267-269 */
  SBA_ERR_ALREADY_SOLVED = 6,       /* CHECK(!problem_), bundle_adjustment.cc:260 */
  SBA_ERR_COMM = 7,
  SBA_ERR_INTERNAL = 8
};

/* ceres::TerminationType values the reference prints (bundle_adjustment.cc:1142-1196) */
enum sba_termination {
  SBA_CONVERGENCE = 0,
  SBA_NO_CONVERGENCE = 1,
  SBA_FAILURE = 2,
  SBA_USER_SUCCESS = 3,
  SBA_USER_FAILURE = 4
};

/*
 * Flattened problem.  Everything is indexed densely; ids of the Reconstruction
 * are mapped by the shim.  Observation k is residual block k of the reference
 * (AddImageToProblem / AddPointToProblem order, bundle_adjustment.cc:348-478).
 *
 *  images   : qvec (w,x,y,z) and tvec as in Image::Qvec()/Tvec() (src/base/image.h:229-230)
 *  cameras  : params as in Camera::Params() padded to SBA_MAX_CAMERA_PARAMS per camera
 *  points   : Point3D::XYZ()
 *
 * Constness (what the reference expresses through SetParameterBlockConstant /
 * SubsetManifold, bundle_adjustment.cc:417-425, 480-530):
 *  image_pose_constant[i]      != 0 : constant-pose cost function (cost_functions.h:90-152)
 *  image_tvec_constant_mask[i] bit k: tvec[k] held constant (SetConstantTvec)
 *  camera_param_constant_mask[c] bit k: params[k] held constant; all params
 *                                 constant == SetParameterBlockConstant
 *  point_constant[p]           != 0 : point block constant
 */
typedef struct sba_problem_desc {
  int32_t num_images;
  int32_t num_cameras;
  int32_t num_points;
  int64_t num_observations;

  const double* image_qvec;                /* [num_images*4] */
  const double* image_tvec;                /* [num_images*3] */
  const int32_t* image_camera;             /* [num_images]   index into cameras */
  const uint8_t* image_pose_constant;      /* [num_images]   */
  const uint8_t* image_tvec_constant_mask; /* [num_images]   */

  const int32_t* camera_model;                /* [num_cameras] sba_camera_model */
  const double* camera_params;                /* [num_cameras*SBA_MAX_CAMERA_PARAMS] */
  const uint16_t* camera_param_constant_mask; /* [num_cameras] */

  const double* points;          /* [num_points*3] */
  const uint8_t* point_constant; /* [num_points]   */

  const int32_t* obs_image; /* [num_observations] */
  const int32_t* obs_point; /* [num_observations] */
  const double* obs_xy;     /* [num_observations*2] Point2D::XY() */

  /* ---- extension X1 (not in the reference) ----
   * Optional per-image semantic maps (label-probability or distance maps) sampled bilinearly at the
   * projection (x, y) of the point (convention of Bitmap::InterpolateBilinear,
   * src/util/bitmap.cc:238-269).  When given, EVERY residual block has three rows,
   *   r = (x - obs_x, y - obs_y, sem_weight * map_img(x, y)),
   * the third with the analytic Jacobian sem_weight * (d map/dx, d map/dy) * d(x,y)/d(pose, point, intrinsics)
   * (d map/d. = differences of the four texels of the cell); a rejected sample (neighbour outside the map)
   * gives r_s = 0 and a zero row.  The loss acts on the squared norm of all three rows.  sba_evaluate,
   * sba_solve and the structural counts (num_residuals = 3 per block) all include the row.
   * NULL / 0 keeps the reference's two-row blocks. */
  const float* sem_maps;         /* [num_images*sem_height*sem_width] row-major, top-down */
  int32_t sem_height;
  int32_t sem_width;
  double sem_weight;

  /* ---- sharding (multi-GPU) ----
   * A shard holds ALL images/cameras and a subset of points with all their
   * observations.  For an unsharded problem: shard_rank=0, shard_count=1. */
  int32_t shard_rank;
  int32_t shard_count;
  int64_t global_num_observations; /* 0 == num_observations */
} sba_problem_desc;

/* Subset of ceres::Solver::Options the reference sets or relies on
 * (src/optim/bundle_adjustment.h:78-93; Ceres 2.1.0 defaults otherwise). */
typedef struct sba_solver_options {
  int32_t linear_solver_type;    /* sba_linear_solver */
  int32_t num_config_images;     /* config_.NumImages() for SBA_SOLVER_AUTO; 0 == desc.num_images */
  int32_t max_num_iterations;                /* 100 */
  int32_t max_linear_solver_iterations;      /* 200 */
  int32_t min_linear_solver_iterations;      /* 0   */
  int32_t max_num_consecutive_invalid_steps; /* 10  */
  double function_tolerance;                 /* 0.0 */
  double gradient_tolerance;                 /* 0.0 */
  double parameter_tolerance;                /* 0.0 */
  int32_t loss_function_type;                /* sba_loss_type */
  double loss_function_scale;                /* 1.0 */
  double initial_trust_region_radius;        /* 1e4  */
  double max_trust_region_radius;            /* 1e16 */
  double min_trust_region_radius;            /* 1e-32 */
  double min_relative_decrease;              /* 1e-3 */
  double min_lm_diagonal;                    /* 1e-6 */
  double max_lm_diagonal;                    /* 1e32 */
  double eta;                                /* 1e-1 */
  int32_t jacobi_scaling;                    /* 1 */
  int32_t update_state_every_iteration;      /* 0; when 1 the callback may read parameters */
} sba_solver_options;

/* One row of the Ceres iteration table (IterationSummary). */
typedef struct sba_iteration_summary {
  int32_t iteration;
  int32_t step_is_valid;
  int32_t step_is_successful;
  int32_t linear_solver_iterations;
  double cost;
  double cost_change;
  double gradient_max_norm;
  double step_norm;
  double relative_decrease;
  double trust_region_radius;
  double eta;
} sba_iteration_summary;

/* Fields of ceres::Solver::Summary the reference's callers read
 * (bundle_adjustment.cc:1142-1196, incremental_mapper.cc:635-636). */
typedef struct sba_summary {
  int32_t termination_type; /* sba_termination */
  int32_t linear_solver_type_used;
  int32_t num_iterations; /* rows in the trace, including iteration 0 */
  int32_t num_successful_steps;
  int32_t num_unsuccessful_steps;
  int32_t num_residual_blocks;
  int32_t num_residuals;
  int32_t num_residuals_reduced;
  int32_t num_parameter_blocks_reduced;
  int32_t num_parameters_reduced;
  int32_t num_effective_parameters_reduced;
  int64_t num_linear_solver_iterations;
  double initial_cost;
  double final_cost;
  double fixed_cost;
  double total_time_in_seconds;
  /* device-side stage times (CUDA events), seconds */
  double time_evaluate_s;
  double time_eliminate_s;
  double time_linear_solve_s;
  double time_update_s;
  double device_time_s; /* first kernel to last kernel of the solve, CUDA events on the solver's stream */
} sba_summary;

typedef struct sba_handle sba_handle;

/* Return 0 to continue, 1 == SOLVER_TERMINATE_SUCCESSFULLY, 2 == SOLVER_ABORT
 * (ceres::IterationCallback, used at src/controllers/bundle_adjustment.cc:43-61). */
typedef int (*sba_iteration_callback)(const sba_iteration_summary* it, void* user);

/* --- lifecycle ---------------------------------------------------------- */
void sba_default_solver_options(sba_solver_options* o);
int sba_create(const sba_problem_desc* desc, sba_handle** out);
/* Label maps that stay on the device across problems.  The reference's semantic adjusters read every image's label
 * map once (SemanticBundleAdjuster::ReadDepthAndSemanticMaps, src/optim/semantic_bundle_adjustment.cc:1020-1066), keep
 * them in a member (semantic_maps_, semantic_bundle_adjustment.h:260) and hand pointers to every cost functor
 * (:728-729, :823); the maps of an image do not change between the mapper's adjustments (AdjustGlobalBundle,
 * src/sfm/incremental_mapper.cc:668-711).  sba_maps_create copies `maps` ([num_images*height*width] row-major,
 * top-down, as sba_problem_desc::sem_maps) to the current device once; sba_create_with_maps builds a problem over
 * them without moving them again (desc->sem_maps must be NULL, desc->num_images must equal the set's;
 * desc->sem_weight applies; sem_height / sem_width of the descriptor are ignored).  The set is reference-counted:
 * sba_maps_destroy may be called while handles created over it are alive, the device memory goes with the last of
 * them.  Results are bit-identical to sba_create with the same maps in the descriptor. */
typedef struct sba_maps sba_maps;
int sba_maps_create(int32_t num_images, int32_t height, int32_t width, const float* maps, sba_maps** out);
void sba_maps_destroy(sba_maps* maps);
int sba_create_with_maps(const sba_problem_desc* desc, const sba_maps* maps, sba_handle** out);
/* Host-only: the structural counts Ceres reports for this problem (Solver::Summary
 * num_residuals_reduced, num_effective_parameters_reduced, ... as pinned by
 * src/optim/bundle_adjustment_test.cc:186-645).  Needs no device. */
int sba_problem_counts(const sba_problem_desc* desc, sba_summary* counts);
void sba_destroy(sba_handle* h);
/* Select the CUDA device used by subsequent sba_create calls of this thread (one process per GPU). */
int sba_set_device(int32_t device);
/* Restore the parameters the handle was created with (kept resident on the device) and make the
 * handle solvable again.  The reference's BundleAdjuster is one-shot (bundle_adjustment.cc:260);
 * this exists so that a benchmark can time repeated solves with all inputs already in HBM. */
int sba_reset(sba_handle* h);
/* Page-locked host memory for the caller's input arrays (cudaHostAlloc).  Arrays that live here are copied to the
 * device by DMA straight from the caller's buffer; pageable arrays are staged through the library's own pinned
 * buffers first.  Purely an optimisation: every entry point accepts either kind.  NULL when no device / no memory. */
void* sba_host_alloc(uint64_t bytes);
void sba_host_free(void* ptr);
const char* sba_last_error(void);
const char* sba_version(void);
int sba_device_count(void);

/* --- evaluation (parity surface; Problem::Evaluate equivalent) ----------
 * All outputs are host pointers, nullable, in REFERENCE residual order:
 *   residuals   [n_obs*2]   loss-corrected
 *   jac_pose    [n_obs*2*6] row-major 2x6: d r / d (quat tangent(3), tvec(3)); zero columns when constant
 *   jac_intr    [n_obs*2*SBA_MAX_CAMERA_PARAMS] d r / d camera params (ambient; constant params zeroed)
 *   jac_point   [n_obs*2*3]
 *   jac_qvec_ambient [n_obs*2*4] raw autodiff-equivalent block w.r.t. (w,x,y,z) before the manifold
 *   sem_residuals [n_obs] third residual row (extension X1; only when maps were given), loss-corrected like `residuals`
 * apply_loss: 0 = raw functor outputs (CostFunction::Evaluate), 1 = loss-corrected (Problem::Evaluate). */
int sba_evaluate(sba_handle* h, int32_t loss_function_type, double loss_function_scale,
                 int32_t apply_loss, double* cost, double* residuals, double* jac_pose,
                 double* jac_intr, double* jac_point, double* jac_qvec_ambient,
                 double* sem_residuals);

/* Jacobian of the third (semantic) residual row as left by the last sba_evaluate / linearisation on this handle,
 * REFERENCE residual order, same column conventions as sba_evaluate: jac_pose [n_obs*6] (quat tangent, tvec),
 * jac_intr [n_obs*SBA_MAX_CAMERA_PARAMS] (ambient, constant params zero), jac_point [n_obs*3].  Any may be NULL. */
int sba_get_semantic_jacobian(sba_handle* h, double* jac_pose, double* jac_intr, double* jac_point);

/* --- cost evaluation that allocates nothing -----------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Returns 0.5 * sum rho(|r|^2) over all residual blocks at the handle's CURRENT parameters -- the value sba_evaluate
 * returns in `cost` -- through a variant of the residual-only kernel whose scratch (per-CTA partial sums, arrival
 * ticket, totals) is a file-scope `static __device__` array of SBA_NO_ALLOC_SCRATCH_BYTES instead of cudaMalloc'ed
 * memory.  The call makes no cudaMalloc / cudaFree / allocating library call; inputs are the handle's device-resident
 * problem, outputs are cost_host (host pointer) and / or cost_device (device pointer on the handle's GPU; the value is
 * written by a kernel on the handle's stream); either may be NULL.  The call returns after a stream synchronisation.
 * Limits: the launch grid min(ceil(n_obs / 128), 16 * SMs) must not exceed SBA_NO_ALLOC_MAX_CTAS
 * (SBA_ERR_INVALID_ARGUMENT otherwise; 2368 CTAs on a 148-SM B200 for any n_obs); sharded handles are rejected.
 * Every GPU that loads the library holds the array until the process exits.  ALL calls share it and it keeps the last
 * call's contents (the entry point clears what must start at zero): calls are serialised inside the library, and a
 * caller that bypasses that by overlapping two such launches on one GPU would race. */
#define SBA_NO_ALLOC_MAX_CTAS 4096
#define SBA_NO_ALLOC_SCRATCH_BYTES ((32 + 2 * SBA_NO_ALLOC_MAX_CTAS) * 8) /* 65 792 B */
int sba_cost_no_alloc(sba_handle* h, int32_t loss_function_type, double loss_function_scale,
                      double* cost_host, double* cost_device);

/* --- solve -------------------------------------------------------------- */
int sba_solve(sba_handle* h, const sba_solver_options* opts, sba_iteration_callback cb,
              void* user, sba_summary* summary);
/* Copy the per-iteration trace of the last solve (up to cap rows); returns rows written via *n. */
int sba_get_trace(sba_handle* h, sba_iteration_summary* rows, int32_t cap, int32_t* n);
/* D2H of the current parameters, layouts as in sba_problem_desc; any pointer may be NULL. */
int sba_get_parameters(sba_handle* h, double* image_qvec, double* image_tvec,
                       double* camera_params, double* points);

/* --- hot-path pieces exposed for measurement (bench.py) -----------------
 * Each runs `repeat` launches of the named stage on device-resident data and
 * returns the mean CUDA-event time per launch in milliseconds. */
int sba_bench_stage(sba_handle* h, const char* stage, int32_t repeat, double* ms_per_launch,
                    int64_t* algorithmic_bytes);
/* Mapper-side per-observation pass: for every observation of `desc`, in the caller's order,
 *   sq_error[k] = CalculateSquaredReprojectionError(xy, X, qvec, tvec, camera)   src/base/projection.cc:111-128
 *                 (DBL_MAX when the point is not in front of the camera), and
 *   depth[k]    = third row of the projection matrix times the point, the quantity HasPointPositiveDepth tests
 *                 (src/base/projection.cc:191-195).
 * Replaces the per-track-element loops of Reconstruction::FilterPoints3DWithLargeReprojectionError and
 * ::FilterObservationsWithNegativeDepth (src/base/reconstruction.cc:1472-1523, 647-665).  Either output may be NULL.
 * Only the geometry fields of `desc` are read (no constancy masks, no maps). */
int sba_observation_errors(const sba_problem_desc* desc, double* sq_error, double* depth);
/* Reconstruction::Normalize(extent, p0, p1, use_images) (src/base/reconstruction.cc:321-345; the mapper calls it after
 * every global BA, src/sfm/incremental_mapper.cc:711).  In place on the caller's host arrays: all image poses
 * (image_registered[i] != 0 marks reg_image_ids_; NULL = all registered) and all points.  Bounds and centroid come from
 * the float-cast, per-axis sorted projection centres of the registered images (use_images) or points
 * (ComputeBoundsAndCentroid, :358-422); the similarity (scale, identity rotation, -scale * centroid) is applied as
 * Reconstruction::Transform does (:424-431; SimilarityTransform3::TransformPoint / ::TransformPose,
 * src/base/similarity_transform.cc:211-240) -- quaternions come back normalised, through the rotation matrix.
 * Fewer than two elements: nothing changes (:325-328).  scale_out / translation_out[3] (nullable): the transform applied. */
int sba_normalize(int32_t num_images, double* image_qvec, double* image_tvec, const uint8_t* image_registered,
                  int64_t num_points, double* points, double extent, double p0, double p1, int32_t use_images,
                  double* scale_out, double* translation_out);
/* Device blocks released by destroyed handles are kept (per device, up to a third of its memory) for the next handle
 * of the same shape; this hands them all back to the driver's pool.  Current device; no handle may be in a call. */
void sba_release_cached_memory(void);
/* Parity surface for the reduced camera system: y = (S + D^2) x of the current linearisation (Jacobi-scaled, LM
 * diagonal of the default options), through the implicit products (use_blocks = 0; what Ceres' ITERATIVE_SCHUR does,
 * SURVEY.md App. B.5) or the explicit block-sparse S (use_blocks = 1; SBA_ERR_INVALID_ARGUMENT when the problem does
 * not use it).  *n receives the reduced dimension; x / y may be NULL to query it. */
int sba_schur_product(sba_handle* h, int32_t use_blocks, const double* x, double* y, int32_t* n);
/* Number of kernels this library launched on h's stream since creation. */
int64_t sba_kernel_launch_count(sba_handle* h);

/* --- multi-GPU (one process per GPU; NCCL over NVLink, and for collectives
 * of <= 1 MB this library's own all-reduce kernel over CUDA-IPC peer memory,
 * set up inside sba_comm_init; SBA_NO_P2P_ALLREDUCE=1 keeps all on NCCL) --- */
#define SBA_COMM_ID_BYTES 128
int sba_comm_get_unique_id(void* id128);
int sba_comm_init(sba_handle* h, int32_t nranks, int32_t rank, const void* id128);

/* Host-side sharding of a problem by point ownership (pure CPU).
 * Fills obs_keep[num_observations] (1 = this rank owns the observation) and
 * point_keep[num_points]; contiguous point ranges balanced by observation count. */
int sba_shard_points(const sba_problem_desc* desc, int32_t nranks, int32_t rank,
                     uint8_t* point_keep, uint8_t* obs_keep);

/* --- semantic pair residual (reference a10/a11) -------------------------
 * src/base/semantic_cost_functions.h:88-208 evaluated for a batch of residual
 * blocks, with Ceres-style CENTRAL numeric differentiation (App. B.6). */
typedef struct sba_semantic_desc {
  int32_t num_images;
  int32_t height, width;
  const double* image_qvec;     /* [num_images*4] */
  const double* image_tvec;     /* [num_images*3] */
  const double* camera_params;  /* [num_images*3] SIMPLE_PINHOLE f,cx,cy of each image's camera */
  const float* depth_maps;      /* [num_images*H*W] ROW-major (y*W+x) */
  const float* semantic_maps;   /* [num_images*H*W] ROW-major */
  int64_t num_blocks;
  const int32_t* block_image1;  /* [num_blocks] */
  const int32_t* block_image2;  /* [num_blocks] */
  const int32_t* block_x1;      /* [num_blocks] */
  const int32_t* block_y1;      /* [num_blocks] */
  double depth_error_threshold; /* 2.0, semantic_bundle_adjustment.h:78 */
  double relative_step_size;    /* 1e-3, semantic_bundle_adjustment.h:89 */
} sba_semantic_desc;

/* residuals[num_blocks]; status[num_blocks] (-1 OOB, -2 invalid depth, 10 valid);
 * jac[num_blocks*14]: ambient central-difference d r/d(q1(4),t1(3),q2(4),t2(3)); nullable. */
int sba_semantic_evaluate(const sba_semantic_desc* d, double* residuals, int32_t* status,
                          double* jac, double* ms_device);

/* --- SemanticBundleAdjuster as a solve (reference a11) -------------------
 * src/optim/semantic_bundle_adjustment.cc:430-544 (Solve), :646-668 (SetUp: all ordered image pairs), :699-906
 * (AddImagePairToProblem: grid of error_computation_pixel_step, zero-depth skip, constant-first / constant-second /
 * both-variable functor), :670-693 (manifolds).  Pose-only problem: one residual block per (ordered pair, grid pixel with
 * depth), central numeric differentiation, LM on the dense normal equations of the pose blocks (no point blocks to
 * eliminate: SURVEY.md App. B.1).  Poses are updated IN PLACE, as the reference mutates the Reconstruction; the
 * quaternions of all images are normalised first (:729-730).  Intrinsics are constant SIMPLE_PINHOLE (:604-644).
 * Returns SBA_ERR_EMPTY_PROBLEM when no residual block remains (Solve returns false, :484-486). */
typedef struct sba_semantic_problem {
  int32_t num_images;            /* config_.Images() in iteration order */
  int32_t height, width;
  double* image_qvec;            /* [num_images*4] in / out */
  double* image_tvec;            /* [num_images*3] in / out */
  const double* camera_params;   /* [num_images*3] f, cx, cy of each image's camera */
  const float* depth_maps;       /* [num_images*H*W] ROW-major (y*W+x) */
  const float* semantic_maps;    /* [num_images*H*W] ROW-major */
  const uint8_t* image_pose_constant;      /* [num_images] !refine_extrinsics || HasConstantPose; NULL = none */
  const uint8_t* image_tvec_constant_mask; /* [num_images] bit k: tvec[k] constant (SetConstantTvec); NULL = none */
  int32_t pixel_step;            /* error_computation_pixel_step, 10 (semantic_bundle_adjustment.h:83) */
  double depth_error_threshold;  /* 2.0 (:78) */
  double relative_step_size;     /* numeric_relative_step_size, 1e-3 (:89) */
} sba_semantic_problem;

/* opts: loss, iteration limits, tolerances as for sba_solve (linear_solver_type only names the policy: every choice is a
 * dense Cholesky of the pose system here).  trace (nullable) receives up to trace_cap rows of the iteration table. */
int sba_semantic_solve(const sba_semantic_problem* problem, const sba_solver_options* opts, sba_iteration_callback cb,
                       void* user, sba_summary* summary, sba_iteration_summary* trace, int32_t trace_cap, int32_t* trace_n);

/* --- GeometricSemanticBundleAdjuster as a solve (reference a12 + a13) ----
 * src/optim/geometric_semantic_bundle_adjustment.cc:481-604 (Solve), :714-726 (ScaledLoss(nullptr, 1 / N_images) on every
 * IoU block), :835-996 (per image x cylinder one of three functors: constant camera / both variable / constant
 * geometry, for Cylinder and for CylinderBy2Points -- the five functors of geometric_semantic_cost_functions.h:33-348),
 * :1126-1213 (QuaternionManifold on the cylinder quaternion, lower bound 0 on the radius), :670-693 pose manifolds.
 * One residual 1 - IoU per (image, cylinder), central numeric differentiation, dense LM over poses and cylinders.
 * The radius bound is enforced by clamping in Plus.  Only the TRIVIAL loss is legal (:699-706).
 * Optional landmark blocks (include_landmark_error, :729-794, 1361-1469): num_landmarks > 0 adds one ordinary
 * reprojection block per listed observation (BundleAdjustment[ConstantPose]CostFunction<SimplePinholeCameraModel>,
 * intrinsics constant) under ScaledLoss(nullptr, landmark_error_weight / total number of 2-D features) (:758-765), and
 * the 3-D points become parameter blocks (constant where point_constant says so, :1448-1469).  The step is then the
 * dense normal-equations solve over poses, cylinders AND points (small problems: 2 * num_landmarks * tangent
 * dimension <= 2^28).  Poses, cylinders and points are updated in place. */
typedef struct sba_cylinder_problem {
  int32_t num_images;
  int32_t height, width;
  double* image_qvec;            /* [num_images*4] in / out (normalised first, :849) */
  double* image_tvec;            /* [num_images*3] in / out */
  const double* camera_params;   /* [num_images*3] */
  const uint8_t* masks;          /* [num_images*H*W] ROW-major, non-zero = label == trunk_semantic_class (:1328-1333) */
  const uint8_t* image_pose_constant;
  const uint8_t* image_tvec_constant_mask;
  int32_t num_cylinders;
  int32_t cylinder_kind;         /* 0: Cylinder (qvec, tvec, radius, height); 1: CylinderBy2Points (tvec_1, tvec_2, radius) */
  double* cylinder_params;       /* [num_cylinders*9] in / out; kind 0: q(4) t(3) r h; kind 1: t1(3) t2(3) r, two unused */
  int32_t refine_geometry;       /* options_.refine_geometry (geometric_semantic_bundle_adjustment.h:78) */
  double relative_step_size;     /* 1e-3 (:111) */
  /* landmark blocks (all zero / NULL: none) */
  int32_t num_points;
  double* points;                /* [num_points*3] in / out */
  const uint8_t* point_constant; /* [num_points] nullable */
  int64_t num_landmarks;
  const int32_t* landmark_image; /* [num_landmarks] */
  const int32_t* landmark_point; /* [num_landmarks] */
  const double* landmark_xy;     /* [num_landmarks*2] */
  double landmark_weight;        /* options_.landmark_error_weight / total_num_2d_features (:758) */
} sba_cylinder_problem;

int sba_cylinder_solve(const sba_cylinder_problem* problem, const sba_solver_options* opts, sba_iteration_callback cb,
                       void* user, sba_summary* summary, sba_iteration_summary* trace, int32_t trace_cap, int32_t* trace_n);

/* --- cylinder IoU residual (reference a12) ------------------------------
 * 1 - IoU of the projected cylinder quadrilateral against a boolean mask,
 * src/util/cylinder.h:27-121, 352-540. */
typedef struct sba_cylinder_desc {
  int32_t num_images;
  int32_t height, width;
  const double* image_qvec;    /* [num_images*4] */
  const double* image_tvec;    /* [num_images*3] */
  const double* camera_params; /* [num_images*3] */
  const uint8_t* masks;        /* [num_images*H*W] ROW-major, non-zero = trunk */
  int32_t num_cylinders;
  const double* cyl_qvec;      /* [num_cylinders*4] */
  const double* cyl_tvec;      /* [num_cylinders*3] */
  const double* cyl_radius;    /* [num_cylinders]   */
  const double* cyl_height;    /* [num_cylinders]   */
  int64_t num_blocks;
  const int32_t* block_image;    /* [num_blocks] */
  const int32_t* block_cylinder; /* [num_blocks] */
  double relative_step_size;
} sba_cylinder_desc;

/* residuals[num_blocks] = 1 - IoU; jac[num_blocks*16] central differences over
 * (cam q(4), cam t(3), cyl q(4), cyl t(3), radius, height); nullable. */
int sba_cylinder_evaluate(const sba_cylinder_desc* d, double* residuals, double* jac,
                          double* ms_device);

/* Parity surface of the two halves of the cylinder residual:
 *   sba_cylinder_quadrilateral: Cylinder::ProjectToQuadrilateral (src/util/cylinder.h:352-474) per block -> pts8
 *     [num_blocks*8] (x, y of the four vertices) and ok [num_blocks] (0 where the reference throws);
 *   sba_draw_quadrilateral: drawQuadrilateral (src/util/cylinder.h:27-121) of four given vertices into a row-major
 *     height x width mask (1 inside) -- identical vertices must give identical masks on both sides. */
int sba_cylinder_quadrilateral(const sba_cylinder_desc* d, double* pts8, int32_t* ok);
int sba_draw_quadrilateral(const double* pts8, int32_t height, int32_t width, uint8_t* mask);

/* --- camera-rig reprojection functor (reference a3) ----------------------
 * RigBundleAdjustmentCostFunction, src/base/cost_functions.h:160-214: composed pose
 * q = rel_q (x) rig_q, t = R(rel_q) rig_t + rel_t, then the reprojection residual.  Evaluates a batch of
 * residual blocks as RigBundleAdjuster adds them (src/optim/bundle_adjustment.cc:1017-1031). */
typedef struct sba_rig_desc {
  int32_t num_rig_poses;        /* rig snapshots */
  const double* rig_qvec;       /* [num_rig_poses*4] */
  const double* rig_tvec;       /* [num_rig_poses*3] */
  int32_t num_rel_poses;        /* cameras of the rig */
  const double* rel_qvec;       /* [num_rel_poses*4] CameraRig::RelativeQvec */
  const double* rel_tvec;       /* [num_rel_poses*3] CameraRig::RelativeTvec */
  int32_t num_cameras;
  const int32_t* camera_model;  /* [num_cameras] sba_camera_model */
  const double* camera_params;  /* [num_cameras*SBA_MAX_CAMERA_PARAMS] */
  int32_t num_points;
  const double* points;         /* [num_points*3] */
  int64_t num_blocks;
  const int32_t* block_rig;     /* [num_blocks] */
  const int32_t* block_rel;     /* [num_blocks] */
  const int32_t* block_camera;  /* [num_blocks] */
  const int32_t* block_point;   /* [num_blocks] */
  const double* block_xy;       /* [num_blocks*2] observed Point2D::XY() */
} sba_rig_desc;

#define SBA_RIG_JAC_COLS (4 + 3 + 4 + 3 + 3 + SBA_MAX_CAMERA_PARAMS)
/* residuals[num_blocks*2]; jac[num_blocks*2*SBA_RIG_JAC_COLS] (nullable): per residual row the ambient blocks
 * d r / d (rig_q(4), rig_t(3), rel_q(4), rel_t(3), point(3), camera params(12)), as CostFunction::Evaluate returns. */
int sba_rig_evaluate(const sba_rig_desc* d, double* residuals, double* jac, double* ms_device);

/* --- RigBundleAdjuster as a solve ------------
 * RigBundleAdjuster::Solve, src/optim/bundle_adjustment.cc:789-1140.  The flat problem holds what SetUp adds to Ceres:
 * one residual block per kept observation with the parameter blocks (rig pose of the snapshot, relative pose of the
 * camera in the rig, point, intrinsics) (:1017-1031).  An image OUTSIDE every rig (:981-1014) is a block whose
 * "relative pose" is the image's own pose and whose rig pose is a constant identity (1,0,0,0 | 0,0,0): the composed
 * pose and its Jacobian are then exactly those of the ordinary functor.  Constancy:
 *   rig_pose_constant[i]       1: q and t constant (the identity above)
 *   rel_pose_constant[i]       1: q and t constant (reference camera of a rig, :1040-1046; refine_relative_poses = false;
 *                              images with constant pose)
 *   rel_tvec_constant_mask[i]  bit j: tvec[j] constant (SubsetManifold, :1049-1053)
 *   camera_param_constant_mask, point_constant: as in sba_problem_desc
 * Blocks that no residual block names are left untouched.  All parameter arrays are IN/OUT.
 * SMALL-PROBLEM PATH: the LM step solves the dense normal equations over ALL blocks, points included (what the Schur
 * solvers compute after eliminating the points, to rounding); J is held dense and problems with
 * 2 * num_blocks * tangent_dimension > 2^28 are refused with SBA_ERR_INVALID_ARGUMENT.
 * Returns SBA_ERR_EMPTY_PROBLEM when there is no residual block (Solve returns false, :824-826). */
typedef struct sba_rig_problem {
  int32_t num_rig_poses;
  double* rig_qvec;                   /* [num_rig_poses*4] */
  double* rig_tvec;                   /* [num_rig_poses*3] */
  const uint8_t* rig_pose_constant;   /* [num_rig_poses] nullable */
  int32_t num_rel_poses;
  double* rel_qvec;                   /* [num_rel_poses*4] */
  double* rel_tvec;                   /* [num_rel_poses*3] */
  const uint8_t* rel_pose_constant;       /* [num_rel_poses] nullable */
  const uint8_t* rel_tvec_constant_mask;  /* [num_rel_poses] nullable */
  int32_t num_cameras;
  const int32_t* camera_model;        /* [num_cameras] */
  double* camera_params;              /* [num_cameras*SBA_MAX_CAMERA_PARAMS] */
  const uint16_t* camera_param_constant_mask; /* [num_cameras] nullable */
  int32_t num_points;
  double* points;                     /* [num_points*3] */
  const uint8_t* point_constant;      /* [num_points] nullable */
  int64_t num_blocks;
  const int32_t* block_rig;           /* [num_blocks] */
  const int32_t* block_rel;           /* [num_blocks] */
  const int32_t* block_camera;        /* [num_blocks] */
  const int32_t* block_point;         /* [num_blocks] */
  const double* block_xy;             /* [num_blocks*2] */
} sba_rig_problem;
int sba_rig_solve(sba_rig_problem* problem, const sba_solver_options* opts, sba_iteration_callback cb, void* user,
                  sba_summary* summary, sba_iteration_summary* trace, int32_t trace_cap, int32_t* trace_n);

#ifdef __cplusplus
}
#endif
#endif /* SBA_B200_H_ */
