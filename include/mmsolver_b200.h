/*
 * mmsolver_b200.h -- C-ABI of the B200-native mmSolver solve hot path.
 *
 * This is the drop-in boundary for ONE path of the reference
 * (bpatchasaheb/mayaMatchMoveSolver): the Levenberg-Marquardt solve loop
 *   solveFrames()            src/mmSolver/adjust/adjust_base.cpp:713
 *   solveFunc()              src/mmSolver/adjust/adjust_solveFunc.cpp:529
 *   measureErrors()          src/mmSolver/adjust/adjust_measureErrors.cpp:523
 *   setParameters()          src/mmSolver/adjust/adjust_setParameters.cpp:253
 *   solve_3d_cminpack_lmder  src/mmSolver/adjust/adjust_cminpack_lmder.cpp:64
 *
 * Plain pointers and sizes only; no torch / C++ types cross this ABI.
 * Every function returns an mmb_status (0 = ok, <0 = error) and never throws.
 * The library owns device memory inside the opaque mmb_problem handle; the
 * caller owns every host array it passes in or receives results into.  No
 * host pointer is retained after a call returns.
 *
 * The scene description mirrors what the reference hands to its solve
 * callback in SolverData (adjust_data.h:188-261) when sceneGraphMode ==
 * kMMSceneGraph: the flattened mmscenegraph FlatScene
 * (lib/rust/mmscenegraph/src/scene/flat.rs:41-71) + AttrDataBlock
 * (src/attr/datablock.rs:76-111) + the lens model tables
 * (mayahelper/maya_lens_model_utils.cpp:859) + the solver attribute list.
 */
#ifndef MMSOLVER_B200_H
#define MMSOLVER_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MMB_ABI_VERSION 2

/* ---- status codes ------------------------------------------------------ */
typedef int32_t mmb_status;
#define MMB_OK 0
#define MMB_ERR_INVALID_ARGUMENT (-1)
#define MMB_ERR_CUDA (-2)
#define MMB_ERR_NO_DEVICE (-3)
#define MMB_ERR_NOT_ENOUGH_OBJECTS (-4) /* adjust_base.cpp:844-862 */
#define MMB_ERR_MORE_PARAMS_THAN_ERRORS (-5) /* adjust_base.cpp:864-882 */
#define MMB_ERR_UNSUPPORTED (-6)
#define MMB_ERR_INTERNAL (-7)
#define MMB_ERR_COMM (-8) /* NCCL missing / collective failed */

/* ---- attribute references ----------------------------------------------
 * Every scene value that the reference keeps in its AttrDataBlock is named by
 * an int32 "attr ref":  ref >= 0  -> static_values[ref]
 *                        ref <  0  -> anim_values[(-ref-1) * num_frames + f]
 * (datablock.rs:76-111, animdense.rs:45-53).
 */
typedef int32_t mmb_attr_ref;
#define MMB_ANIM_REF(a) (-(int32_t)(a)-1)
#define MMB_ATTR_NONE INT32_MIN /* unused slot (e.g. lens coeff) -> 0.0 */

/* rotate orders: lib/rust/mmscenegraph/src/math/rotate/euler.rs */
enum { MMB_ROO_XYZ = 0, MMB_ROO_YZX = 1, MMB_ROO_ZXY = 2,
       MMB_ROO_XZY = 3, MMB_ROO_YXZ = 4, MMB_ROO_ZYX = 5 };
/* film fit: lib/rust/mmscenegraph/src/math/camera.rs:32-37 */
enum { MMB_FILM_FIT_FILL = 0, MMB_FILM_FIT_HORIZONTAL = 1,
       MMB_FILM_FIT_VERTICAL = 2, MMB_FILM_FIT_OVERSCAN = 3 };
/* lens model types: lib/cppbind/mmlens (LensModelType) */
enum { MMB_LENS_PASSTHROUGH = 0,
       MMB_LENS_3DE_CLASSIC = 1,                 /* 5 coeffs  */
       MMB_LENS_3DE_RADIAL_STD_DEG4 = 2,         /* 8 coeffs  */
       MMB_LENS_3DE_ANAMORPHIC_STD_DEG4 = 3,     /* 13 coeffs */
       MMB_LENS_3DE_ANAMORPHIC_STD_DEG4_RESCALED = 4 /* 14 coeffs */ };
#define MMB_LENS_MAX_COEFFS 14
#define MMB_TFM_ATTRS 9  /* tx ty tz rx ry rz sx sy sz */
#define MMB_CAM_ATTRS 8  /* sensor_w_mm sensor_h_mm focal_mm lens_offset_x_mm
                            lens_offset_y_mm near far camera_scale
                            (dag.rs:120-172) */

/* ---- scene description ------------------------------------------------- */
typedef struct mmb_scene_desc {
    int32_t abi_version; /* MMB_ABI_VERSION */

    /* frames to solve (MTimeArray frameList / mmsgFrameList) */
    int32_t num_frames;
    const int32_t *frames; /* [F] frame numbers (informational) */

    /* attribute value store */
    int32_t num_static;
    const double *static_values; /* [num_static] */
    int32_t num_anim;
    const double *anim_values;   /* [num_anim * F] */

    /* transform hierarchy, parents before children (dag.rs:234-327) */
    int32_t num_transforms;
    const int32_t *tfm_parent;        /* [T] -1 = none, else < own index */
    const int32_t *tfm_rotate_order;  /* [T] MMB_ROO_* */
    const mmb_attr_ref *tfm_attrs;    /* [T*9] */

    /* cameras (flat.rs:263-283) */
    int32_t num_cameras;
    const int32_t *cam_tfm;           /* [C] transform index */
    const mmb_attr_ref *cam_attrs;    /* [C*8] */
    const int32_t *cam_film_fit;      /* [C] MMB_FILM_FIT_* */
    const int32_t *cam_render_width;  /* [C] */
    const int32_t *cam_render_height; /* [C] */
    const int32_t *cam_lens;          /* [C] lens index or -1 */

    /* bundles */
    int32_t num_bundles;
    const int32_t *bnd_tfm; /* [B] transform index */

    /* markers: dense marker x frame tables (marker-major), as cached by
     * countUpNumberOfErrors (adjust_relationships.cpp:75-221) */
    int32_t num_markers;
    const int32_t *mkr_cam;    /* [M] */
    const int32_t *mkr_bnd;    /* [M] */
    const double *mkr_x;       /* [M*F] -0.5..0.5, overscan applied */
    const double *mkr_y;       /* [M*F] */
    const double *mkr_weight;  /* [M*F] */
    const uint8_t *mkr_enable; /* [M*F] */

    /* lens models (mmlens LensModel + CameraParameters) */
    int32_t num_lenses;
    const int32_t *lens_model;        /* [L] MMB_LENS_* */
    const mmb_attr_ref *lens_attrs;   /* [L*MMB_LENS_MAX_COEFFS] */
    const double *lens_camera_params; /* [L*6] focal_cm fbw_cm fbh_cm
                                         pixel_aspect lco_x_cm lco_y_cm
                                         (lens_model.h:42 defaults
                                          3.0 3.6 2.4 1.0 0.0 0.0) */

    /* solver attributes, in the order of the reference's '-attr' flags
     * (countUpNumberOfUnknownParameters, adjust_relationships.cpp:223-337):
     * an anim ref expands to F parameters, a static ref to one. */
    int32_t num_attrs;
    const mmb_attr_ref *attr_ref; /* [A] */
    const double *attr_min;       /* [A] -FLT_MAX = none (maya_attr.cpp:88) */
    const double *attr_max;       /* [A] +FLT_MAX = none */
    const double *attr_offset;    /* [A] default 0 */
    const double *attr_scale;     /* [A] default 1 */

    /* `-attrStiffness / -attrSmoothness node.attr weightAttr varianceAttr valueAttr`
     * (v1 command; arg_flags_attr_details.cpp:60-200, StiffAttrs / SmoothAttrs
     * adjust_data.h): one extra residual per entry whose weight is > 0
     * (countUpNumberOfErrors, adjust_relationships.cpp:183-219), appended after the
     * marker errors: (1 / gaussian(attr value, value, variance) - 1) * weight
     * (adjust_measureErrors.cpp:311-387).  *_attr indexes the USED solver
     * attribute list, as the reference's attrIndex does; weight / variance /
     * value are attribute-store refs read -- like the attribute itself -- at
     * Maya's current time, here frame index `current_frame`.  Only
     * sceneGraphMode = 1 evaluates them; in mode 2 the rows exist and stay zero
     * (adjust_measureErrors.cpp:516 "TODO: Support stiffness and smoothness"). */
    int32_t num_stiffness;
    const int32_t *stiff_attr;            /* [num_stiffness] */
    const mmb_attr_ref *stiff_weight;     /* [num_stiffness] */
    const mmb_attr_ref *stiff_variance;   /* [num_stiffness] */
    const mmb_attr_ref *stiff_value;      /* [num_stiffness] */
    int32_t num_smoothness;
    const int32_t *smooth_attr;
    const mmb_attr_ref *smooth_weight;
    const mmb_attr_ref *smooth_variance;
    const mmb_attr_ref *smooth_value;
    int32_t current_frame;                /* index into frames (MAnimControl::currentTime) */
} mmb_scene_desc;

/* ---- solver options: SolverOptions, adjust_data.h:133-185 -------------- */
typedef struct mmb_options {
    int32_t iter_max;          /* -iterations, default 100 (maxfev) */
    double tau;                /* -tauFactor 1.0, clamped [0,1]; factor=tau*100 */
    double eps1;               /* ftol 1e-6 */
    double eps2;               /* xtol 1e-6 */
    double eps3;               /* gtol 1e-6 */
    double delta;              /* FD step 1e-4 */
    int32_t auto_diff_type;    /* 0 forward, 1 central */
    int32_t auto_param_scale;  /* 1 -> lmder mode 1, else mode 2 */
    int32_t robust_loss_type;  /* accepted, unused (adjust_defines.h:134) */
    double robust_loss_scale;
    int32_t scene_graph_mode;  /* 2 = mmscenegraph (the reference command's default is 1;
                                  mmb_options_default gives 2, the mode without Maya).
                                  1 = the Maya-DAG-convention arithmetic of the residual
                                  (maya_camera.cpp:75-414,863-894: row-vector matrices,
                                  film offsets in the projection, the C++ film-fit
                                  variant; behind-camera x1e6 and stiffness / smoothness
                                  rows, adjust_measureErrors.cpp:262-387) over the same
                                  flat scene: transforms are still T*R*S, not Maya's DG.
                                  Global solves on one GPU only. */
    int32_t solver_type;       /* 2 = cminpack_lmder (default), 1 = cminpack_lmdif */
    int32_t time_eval_mode;    /* accepted, unused without Maya */
    int32_t accept_only_better;/* default 1 */
    double image_width;        /* default 2048.0 */
    int32_t frame_solve_mode;  /* 0 all frames at once, 1 per frame */
    int32_t remove_unused_markers;
    int32_t remove_unused_attributes;
    /* extensions */
    int32_t lens_index_literal;/* 0 = intended indexing, 1 = the reference's
                                  literal read: reproduce the
                                  [markerIndex+frameIndex] lens table read
                                  (adjust_measureErrors.cpp:463); 0: intended
                                  [marker*F+frame] */
    int32_t device;            /* CUDA device ordinal */
} mmb_options;

/* ---- results: SolverResult / ErrorMetrics, adjust_results.h:119-286 ---- */
typedef struct mmb_result {
    int32_t success;
    int32_t reason_num;              /* cminpack info 0..8 */
    const char *reason_str;          /* static string, adjust_cminpack_base.h */
    double error_final;              /* enorm(fvec) */
    double error_avg, error_min, error_max; /* px, compute_error_stats */
    int32_t iteration_num;           /* nfev */
    int32_t iteration_function_num;  /* iflag=1 calls */
    int32_t iteration_jacobian_num;  /* FD re-evaluations (n per Jacobian) */
    int32_t iteration_attempt_num;   /* LM outer iterations (ours) */
    int32_t user_interrupted;
    int32_t error_is_better;         /* acceptOnlyBetter decision */
    double timer_solve_s, timer_jacobian_s, timer_function_s,
           timer_linear_s;           /* device time, CUDA events */
} mmb_result;

typedef struct mmb_problem mmb_problem; /* opaque */

/* fill with the reference defaults (adjust_defines.h:92-152,
 * arg_flags_solve_info.cpp) */
void mmb_options_default(mmb_options *opt);

/* Build index maps (bit-exact with adjust_relationships.cpp) and upload the
 * scene.  Replaces the head of solveFrames(), adjust_base.cpp:746-1074. */
mmb_status mmb_problem_create(const mmb_scene_desc *scene,
                              const mmb_options *opt, mmb_problem **out);
void mmb_problem_destroy(mmb_problem *p);

/* sizes: numberOfParameters n, numberOfMarkerErrors/2 (= Nobs), numberOfErrors m =
 * 2*Nobs + stiffness + smoothness rows */
mmb_status mmb_problem_sizes(const mmb_problem *p, int32_t *n_params,
                             int32_t *n_obs, int32_t *n_errors);

/* errorToMarkerList [Nobs*2] (marker, frame); paramToAttrList [n*2]
 * (attr, frame | -1).  Either pointer may be NULL. */
mmb_status mmb_get_index_maps(const mmb_problem *p, int32_t *error_to_marker,
                              int32_t *param_to_attr);
/* The marker <-> attribute relationship of the mmSolverAffects command
 * (src/mmSolver/cmd/MMSolverAffectsCmd.cpp; findMarkerToAttributeRelationship,
 * adjust_relationships.cpp:369-466): marker i is affected by solver attribute j
 * when j drives the marker's camera (transform chain, camera shape, lens) or its
 * bundle's transform chain.  marker_affects_attr [M*A] row-major, and the used
 * flags of splitUsedMarkersAndAttributes (adjust_base.cpp:574-610): marker_used
 * [M], attr_used [A].  Any output may be NULL.  Host only: needs no GPU. */
mmb_status mmb_affects(const mmb_scene_desc *scene, uint8_t *marker_affects_attr,
                       uint8_t *marker_used, uint8_t *attr_used);
/* solve_v1 drops markers affected by no attribute / attributes affecting no
 * marker before solveFrames when options.remove_unused_markers /
 * remove_unused_attributes are set (adjust_base.cpp:1322-1395); the index maps,
 * parameters and results of the handle then number the USED lists.  This gives
 * the position of each used marker / attribute in the caller's lists
 * (used_marker_index [num_markers_used], used_attr_index [num_attrs_used];
 * call with NULL arrays to get the counts). */
mmb_status mmb_problem_used_objects(const mmb_problem *p, int32_t *num_markers_used,
                                    int32_t *used_marker_index,
                                    int32_t *num_attrs_used,
                                    int32_t *used_attr_index);
/* normalised marker weights [Nobs] (adjust_relationships.cpp:166-182) */
mmb_status mmb_get_marker_weights(const mmb_problem *p, double *weights);

/* get_initial_parameters(), adjust_base.cpp:260: x[n] in solver space */
mmb_status mmb_get_initial_parameters(const mmb_problem *p, double *x);

/* solveFunc iflag=1 (adjust_solveFunc.cpp:262): fvec[m]; optional
 * error_dist[Nobs] (errorDistanceList, px).  x == NULL measures the scene at its
 * own attribute values, no parameter applied: the initial measureErrors of
 * solveFrames (adjust_base.cpp:1080-1103), which is all a printStatistics call does */
mmb_status mmb_eval_residuals(mmb_problem *p, const double *x, double *fvec,
                              double *error_dist);

/* solveFunc iflag=2 (adjust_solveFunc.cpp:482): dense column-major
 * fjac[n * ldfjac], ldfjac >= m, exactly what cminpack receives. */
mmb_status mmb_eval_jacobian_dense(mmb_problem *p, const double *x,
                                   double *fjac, int32_t ldfjac);

/* Same Jacobian in the compact block layout the device keeps.  For
 * observation i the k = k_max local columns are
 *   values [c * m + 2*i + {0,1}],  column ids [i * k_max + c] (-1 = unused).
 * Call with NULL outputs to query k_max. */
mmb_status mmb_eval_jacobian_blocks(mmb_problem *p, const double *x,
                                    double *values, int32_t *col_ids,
                                    int32_t *k_max);

/* The LM solve (solve_3d_cminpack_lmder + tail of solveFrames).
 * x_out[n] solver-space parameters, attr_values_out[n] external values
 * (what set_maya_attribute_values writes), error_dist_out[Nobs] px. */
mmb_status mmb_solve(mmb_problem *p, mmb_result *result, double *x_out,
                     double *attr_values_out, double *error_dist_out);

/* frameSolveMode = 1 (solve_v1's per-frame loop, adjust_base.cpp:1430-1484):
 * every frame is its own LM problem; the aggregate follows
 * CommandResult::add / divide (adjust_results.h:89-117).  All frames run in ONE
 * launch (one CTA per frame) when every solver attribute is animated; a scene
 * with static solver attributes returns MMB_ERR_UNSUPPORTED (the reference
 * carries their values from one frame's solve into the next).
 * per_frame_results: [F] or NULL.  mmb_solve() routes here when
 * options.frame_solve_mode == 1. */
mmb_status mmb_solve_per_frame(mmb_problem *p, mmb_result *result,
                               mmb_result *per_frame_results, double *x_out,
                               double *attr_values_out, double *error_dist_out);

/* ---- multi-GPU: frame-range shards (one process per GPU) -----------------
 * The reference is single-threaded; what shards is its own structure: a
 * frame's observations touch only that frame's animated parameters
 * (findErrorToParameterRelationship, adjust_relationships.cpp:565-617).
 *   per-frame mode (solve_v1's loop, adjust_base.cpp:1430-1484): shards are
 *     independent, no collective; merge with mmb_aggregate_results.
 *   global mode: every rank runs the same LM control law on its frames; the
 *     reduced (static-parameter) system and a few scalars are all-reduced
 *     once per lmpar trial over NCCL (NVLink / NVSwitch).
 * Every rank passes the FULL scene; the library cuts it to the rank's frames.
 * Parameters / observations of a shard handle are numbered locally;
 * mmb_shard_param_map gives their ids in the unsharded problem. */
#define MMB_COMM_ID_BYTES 128
/* contiguous balanced ranges: [*frame_begin, *frame_end) of rank */
mmb_status mmb_shard_frames(int32_t num_frames, int32_t world, int32_t rank,
                            int32_t *frame_begin, int32_t *frame_end);
/* host only (no GPU needed): n of the shard and, when global_param != NULL,
 * the unsharded parameter id of each of its n_local parameters */
mmb_status mmb_shard_param_map(const mmb_scene_desc *scene, int32_t world,
                               int32_t rank, int32_t *n_local,
                               int32_t *global_param);
mmb_status mmb_problem_create_shard(const mmb_scene_desc *scene,
                                    const mmb_options *opt, int32_t rank,
                                    int32_t world, mmb_problem **out);
/* The same for a caller that holds only its own frames: `slice` is frames
 * [frame_begin, frame_begin + slice->num_frames) of a scene of frames_global
 * frames (static attributes, objects and attr flags identical on every rank).
 * Nothing is cut; use it when the whole scene does not fit one host (cfg5:
 * 4e8 observations).  Unsharded id of local parameter: animated attribute a,
 * local frame f -> its global base + frame_begin + f (as mmb_shard_param_map). */
mmb_status mmb_problem_create_shard_local(const mmb_scene_desc *slice,
                                          const mmb_options *opt,
                                          int32_t frame_begin,
                                          int32_t frames_global,
                                          mmb_problem **out);
/* rank 0 makes the id (an ncclUniqueId; local_backend != 0: an in-process
 * group of handles on one device, used by single-GPU tests) and hands the
 * 128 bytes to the other ranks by any channel (torch.distributed, MPI, file) */
mmb_status mmb_comm_unique_id(uint8_t id[MMB_COMM_ID_BYTES],
                              int32_t local_backend);
/* collective: every rank of the id calls it (NCCL communicator + CUDA IPC mapping
 * of the ranks' exchange arenas, 104 MB of device memory each; MMB_PEER=0 keeps
 * NCCL alone, MMB_PEER_ARENA_MB sizes the arena).  Afterwards mmb_solve() on the
 * handle is the sharded global solve; results that describe the whole problem
 * (reason, iteration counts, error_final, error_avg/min/max) are identical on
 * every rank, x_out / attr_values_out / error_dist_out cover the shard. */
mmb_status mmb_problem_attach_comm(mmb_problem *p,
                                   const uint8_t id[MMB_COMM_ID_BYTES],
                                   int32_t rank, int32_t world);
/* A communicator that outlives handles.  Setting up an NCCL communicator and
 * its first collective costs ~1 s; a session that solves repeatedly (the
 * reference calls mmSolver once per solver step) creates it once -- collective
 * over the ranks of the id, connections warmed up -- and lends it to each
 * handle with mmb_problem_use_comm (the handle does not own it: destroy the
 * communicator after the handles that use it). */
typedef struct mmb_comm mmb_comm; /* opaque */
mmb_status mmb_comm_create(const uint8_t id[MMB_COMM_ID_BYTES], int32_t rank,
                           int32_t world, int32_t device, mmb_comm **out);
void mmb_comm_destroy(mmb_comm *c);
mmb_status mmb_problem_use_comm(mmb_problem *p, mmb_comm *c);
/* CommandResult::add / divide (adjust_results.h:89-117, 208-243) */
mmb_status mmb_aggregate_results(const mmb_result *per_frame, int32_t count,
                                 mmb_result *out);

/* fvec of the accepted point of the last mmb_solve(): errorList [m]
 * (SolveValuesResult, adjust_results.h:298-311) */
mmb_status mmb_get_last_residuals(mmb_problem *p, double *fvec);

/* The `key=value` strings the mmSolver command returns
 * (CommandResult::appendToMStringArray, adjust_results.h:871-938; split char
 * '#', adjust_defines.h:64), in the reference's order, one per line ('\n'),
 * so python/mmSolver/_api/solveresult.py parses them unchanged.  Host only.
 * error_to_marker [n_obs*2] / frames [num_frames] / marker_names [num_markers]
 * (NULL -> "marker<i>") name the per-marker-per-frame deviations;
 * solve_parameter_list [n_params], solve_error_list [2*n_obs], error_dist
 * [n_obs] may be NULL when their flag is off.  *needed = bytes incl. NUL;
 * MMB_ERR_INVALID_ARGUMENT with *needed set when cap is too small. */
#define MMB_FMT_ERROR_PER_MARKER_PER_FRAME 1
#define MMB_FMT_ERROR_PER_FRAME 2
#define MMB_FMT_SOLVE_VALUES 4
#define MMB_FMT_INPUT_STATS 8 /* printStatistics 'inputs': numberOf* counts */
/* stiffness / smoothness rows of the problem (0..2047 each), OR-ed into `flags`:
 * numberOfErrors and solve_error_list [2*n_obs + rows] then include them */
#define MMB_FMT_STIFFNESS(n) (((int32_t)(n) & 0x7ff) << 8)
#define MMB_FMT_SMOOTHNESS(n) (((int32_t)(n) & 0x7ff) << 19)
mmb_status mmb_format_results(const mmb_result *r, int32_t n_params,
                              int32_t n_obs, const int32_t *error_to_marker,
                              const int32_t *frames, int32_t num_frames,
                              const char *const *marker_names,
                              int32_t num_markers,
                              const double *solve_parameter_list,
                              const double *solve_error_list,
                              const double *error_dist, int32_t flags,
                              char *buf, int64_t cap, int64_t *needed);

/* MComputation::isInterruptRequested analogue; callable from another thread.
 * Sharded solve: a request on ANY rank travels through the iteration's all-reduce
 * and stops every rank at the end of the same LM iteration. */
void mmb_request_cancel(mmb_problem *p);

/* device-side benchmarking hooks (inputs already resident in HBM):
 * run `reps` back-to-back residual+FD-Jacobian evaluations at the initial
 * parameters on the problem's stream; returns mean ms per evaluation measured
 * with CUDA events (whole evaluation = setParameters + entity variants +
 * the residual/Jacobian kernel; and that last kernel alone) and the number of
 * kernel launches issued.  The L2 flush, when asked for, is outside the
 * timed brackets. */
mmb_status mmb_bench_eval(mmb_problem *p, int32_t reps, int32_t flush_l2,
                          double *ms_per_eval, double *ms_jacobian_kernel,
                          int32_t *launches);
/* Bulk lens (un)distortion of a point buffer: LensModel::applyModelUndistort
 * (direction 0) / applyModelDistort (direction 1) of
 * lib/cppbind/mmlens/src/lens_model_3de_*.cpp applied to n_points interleaved
 * (x, y) pairs in -0.5..0.5 units -- the per-pixel loop of tools/lensdistortion
 * (ST-maps).  coeffs [MMB_LENS_MAX_COEFFS] in the model's parameter order (NULL
 * = zeros), camera_params [6] as in mmb_scene_desc.lens_camera_params.  Host
 * buffers in and out; the arithmetic is the solve path's own lens code. */
mmb_status mmb_lens_apply(int32_t device, int32_t model, const double *coeffs,
                          const double *camera_params, int32_t direction,
                          int64_t n_points, const double *xy_in,
                          double *xy_out);
/* A chain of lens nodes (each LensModel applies its inputLensModel first,
 * lens_model_3de_*.cpp:44-50): layer 0 is the innermost; models [n_layers],
 * coeffs [n_layers * MMB_LENS_MAX_COEFFS], camera_params [n_layers * 6];
 * at most 8 layers.  (The SOLVER never connects lens chains --
 * maya_lens_model_utils.cpp:711 -- so mmb_scene_desc has one lens per camera.) */
mmb_status mmb_lens_apply_layered(int32_t device, int32_t n_layers,
                                  const int32_t *models, const double *coeffs,
                                  const double *camera_params,
                                  int32_t direction, int64_t n_points,
                                  const double *xy_in, double *xy_out);
/* mmb_lens_apply_layered for callers that must not allocate device memory
 * (tight serving loops, harnesses that forbid cudaMalloc; NOT usable under
 * CUDA graph capture: it takes host pointers and issues blocking copies on the
 * legacy default stream, which a capture cannot record).  No counterpart in the reference.
 * Same arguments, same kernel, same results; the layer table
 * and the staging of the points live in one static __device__ array of
 * 4096 + 2 * MMB_LENS_NO_ALLOC_MAX_POINTS * 16 bytes (8 MiB) that every GPU
 * which loads the library holds until exit.  n_points above
 * MMB_LENS_NO_ALLOC_MAX_POINTS is refused (MMB_ERR_INVALID_ARGUMENT, message in
 * mmb_last_error).  Calls no cudaMalloc / cudaFree / allocating routine: blocking
 * copies and one launch on the legacy default stream.  All calls share the
 * array, which keeps the previous call's contents (nothing in it needs to start
 * at zero); calls that overlap on one GPU -- from several host threads or
 * processes sharing a context -- race on it: serialise them. */
#define MMB_LENS_NO_ALLOC_MAX_POINTS 262144
mmb_status mmb_lens_apply_layered_no_alloc(int32_t device, int32_t n_layers,
                                           const int32_t *models,
                                           const double *coeffs,
                                           const double *camera_params,
                                           int32_t direction, int64_t n_points,
                                           const double *xy_in, double *xy_out);
/* the same kernel timed on a device-resident grid of n_points (CUDA events) */
mmb_status mmb_bench_lens(int32_t device, int32_t model, const double *coeffs,
                          const double *camera_params, int32_t direction,
                          int64_t n_points, int32_t reps, double *ms_per_pass);
/* time (median ms over reps, CUDA events, L2 flushed before each) of a plain
 * streaming kernel that reads read_bytes and writes write_bytes: what the memory
 * system gives a transfer of the SIZE of one Jacobian launch (the copy figure in
 * MEASURED_PEAKS.json is taken over 2 GiB) */
mmb_status mmb_bench_stream_floor(int32_t device, int64_t read_bytes,
                                  int64_t write_bytes, int32_t reps,
                                  double *ms_median);
/* debug: %globaltimer timeline (ns) of the critical path of the last reduced-
 * system Cholesky: [block column][0 diagonal tile, 1 tile below it][6 marks]
 * (csrc/lm_kernels.cuh g_chol_trace); count <= 768 */
mmb_status mmb_debug_chol_trace(uint64_t *out, int32_t count);
/* measured FP64 ceilings of the device in TFLOP/s (dependent-free DFMA chains; DMMA
 * m8n8k4): the roofline denominators of the compute-bound kernels (lens distortion in
 * the loop, the reduced-system SYRK / Cholesky updates) */
mmb_status mmb_bench_fp64_peaks(int32_t device, double *dfma_tflops,
                                double *dmma_tflops);
/* kernels launched by this handle so far (bench.py's gpu_launches) */
int64_t mmb_launch_count(const mmb_problem *p);
/* How the last mmb_solve() on this handle was driven.  The reference's loop is
 * host code calling back into the scene once per function / Jacobian evaluation
 * (lmder -> solveFunc, adjust_cminpack_lmder.cpp:114, adjust_solveFunc.cpp:529).
 * Here an unsharded global solve is one CUDA graph launch (conditional WHILE /
 * IF nodes steered from the device) and one synchronisation to read the result:
 * host_syncs = 1, graph_launches = 1.  A sharded solve is the same single graph
 * launch when the communicator's peer-memory all-reduce is on (CUDA IPC arenas over
 * NVLink: plain kernels, capturable); with NCCL collectives between the kernels
 * (fallback, or a reduced system beyond the arena) it is host-driven
 * (graph_launches = 0, a few host_syncs per LM iteration). */
mmb_status mmb_solve_stats(const mmb_problem *p, int64_t *host_syncs,
                           int64_t *graph_launches);

/* Device buffers come from the device's stream-ordered memory pool and stay
 * cached there when a handle is destroyed, so the next mmb_problem_create of a
 * session (the reference calls mmSolver once per solver step) does not pay for
 * allocation again; the page-locked host blocks the library stages its uploads
 * in are cached the same way (up to 1 GB idle).  This hands both back. */
mmb_status mmb_release_cached_memory(int32_t device);

const char *mmb_status_string(mmb_status s);
/* detail of the last failing call on this thread (static storage) */
const char *mmb_last_error(void);
const char *mmb_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MMSOLVER_B200_H */
