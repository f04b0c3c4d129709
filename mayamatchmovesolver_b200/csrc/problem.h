// Host-side problem description of the B200 mmSolver solve path.
//
// Replaces the head of solveFrames() (adjust_base.cpp:746-1074): the index
// maps of adjust_relationships.cpp are rebuilt bit-exactly on plain arrays
// and then compiled into the "variant plan" the CUDA kernels consume.
//
// Variant plan (ours, not in the reference): an observation (marker, frame)
// depends on its camera's 3x4 world-projection rows, its bundle's world
// position and its lens coefficients.  Every solver parameter perturbs one
// (or more) of those three *entities*, never the observation directly, so the
// finite-difference Jacobian (adjust_solveFunc.cpp:305-479) becomes
//   1. one small launch that evaluates every entity at x and at x + delta*e_p
//      for each parameter p it depends on ("variants"), and
//   2. one streaming launch over observations that combines variants.
#pragma once
#include <cstdint>
#include <string>
#include <memory>
#include <utility>
#include <vector>

#include "../../include/mmsolver_b200.h"

namespace mmb {

// per-marker plan entry: column c of marker m
// std::vector whose resize() leaves new elements uninitialised: the per-observation arrays
// (5e5 .. 5e7 entries) are filled right after the resize by several threads, a zero fill by
// one thread first would only take the page faults serially.
// Their memory comes from a process-wide cache of page-locked host blocks (host_block_acquire):
// the upload of a pageable vector is staged by the driver on the calling thread (16 MB at cfg3:
// 1.3 ms, and more when eight processes do it at once); from a page-locked block it is one
// asynchronous DMA.  Without a device (host-only entry points) the blocks are plain malloc.
void *host_block_acquire(size_t bytes);
void host_block_release(void *p);
void host_block_trim();   // free the cached blocks (mmb_release_cached_memory)
template <class T>
struct NoInitAlloc : std::allocator<T> {
    template <class U> struct rebind { using other = NoInitAlloc<U>; };
    NoInitAlloc() = default;
    template <class U> NoInitAlloc(const NoInitAlloc<U> &) {}
    T *allocate(size_t n) { return static_cast<T *>(host_block_acquire(n * sizeof(T))); }
    void deallocate(T *p, size_t) { host_block_release(p); }
    template <class U, class... Args> void construct(U *ptr, Args &&...args) {
        if constexpr (sizeof...(Args) == 0) ::new ((void *)ptr) U;
        else ::new ((void *)ptr) U(std::forward<Args>(args)...);
    }
};
template <class T> using RawVec = std::vector<T, NoInitAlloc<T>>;

struct PlanEntry {
    int32_t attr;      // solver attribute index (or -1 unused)
    int16_t cam_var;   // 0 = base, else 1-based index into camera dep list
    int16_t bnd_var;   // same for bundle
    int16_t lens_var;  // same for lens
    int16_t pad;
};

struct HostProblem {
    mmb_options opt;
    int F = 0, T = 0, C = 0, B = 0, M = 0, L = 0, A = 0;
    int num_static = 0, num_anim = 0;
    // deep copies of the scene
    std::vector<int32_t> frames;
    std::vector<double> stat0, anim0;
    std::vector<int32_t> tfm_parent, tfm_roo, tfm_attrs;
    std::vector<int32_t> cam_tfm, cam_attrs, cam_fit, cam_rw, cam_rh, cam_lens;
    std::vector<int32_t> bnd_tfm, mkr_cam, mkr_bnd;
    std::vector<int32_t> lens_model, lens_attrs;
    std::vector<double> lens_cam;
    std::vector<int32_t> attr_ref;
    std::vector<double> attr_min, attr_max, attr_off, attr_scl;

    // ---- used / unused split (adjust_base.cpp:574-610) -------------------
    int M_in = 0, A_in = 0;                       // markers / attributes the caller passed
    std::vector<uint8_t> marker_used, attr_used;  // [M_in], [A_in]
    std::vector<int32_t> mkr_src, attr_src;       // used marker / attr -> index in the caller's lists
    // ---- sceneGraphMode = 1 and the stiffness / smoothness rows ------------
    int maya_dag = 0;          // Maya-DAG-convention arithmetic of the residual
    int cur_frame = 0;         // index of "Maya's current time" in the frame list
    std::vector<int32_t> stiff_in, smooth_in;           // [count * 4] attr, weight, variance, value (as passed)
    std::vector<int32_t> ex_attr, ex_w, ex_var, ex_val;  // the counted rows: stiffness first, then smoothness
    int n_stiff = 0, n_smooth = 0, n_extra = 0;
    int m_total = 0;           // numberOfErrors = m + n_extra
    // ---- index maps (adjust_relationships.cpp) --------------------------
    int n = 0, nobs = 0, m = 0;   // m = numberOfMarkerErrors = 2 * nobs
    RawVec<int32_t> err_marker, err_frame;   // errorToMarkerList
    RawVec<double> obs_mx, obs_my, obs_w;    // markerPosList / weights
    std::vector<int32_t> par_attr, par_frame;     // paramToAttrList
    std::vector<int32_t> attr_param_base;         // first param of attr a
    std::vector<uint8_t> attr_anim;               // attr expands per frame
    std::vector<int32_t> stat_attr, anim_attr;    // attribute-store slot -> solver attr or -1
    bool dense_grid = false;                      // nobs == M*F, i = m*F+f

    // ---- entity dependency lists ---------------------------------------
    // solver attrs each entity depends on (sorted attr indices)
    std::vector<std::vector<int32_t>> cam_deps, bnd_deps, lens_deps;
    // does the entity vary per frame (any animated ref in its chain)?
    std::vector<uint8_t> cam_varies, bnd_varies, lens_varies;
    std::vector<uint8_t> cam_mscale_varies;  // sensor w/h is a solver attr
    // variant table offsets (in doubles) and frame strides
    std::vector<int64_t> cam_off, bnd_off, lens_off;
    std::vector<int32_t> cam_fs, bnd_fs, lens_fs;      // F or 1
    std::vector<int32_t> cam_nvar, bnd_nvar, lens_nvar;  // 1 + deps (*2 central)
    int64_t cam_tab_size = 0, bnd_tab_size = 0, lens_tab_size = 0;
    // flattened (entity, variant) work lists for the variant kernel
    std::vector<int32_t> cam_job_ent, cam_job_var, cam_job_attr;
    std::vector<int32_t> bnd_job_ent, bnd_job_var, bnd_job_attr;
    std::vector<int32_t> lens_job_ent, lens_job_var, lens_job_attr;
    int64_t cam_jobs_total = 0, bnd_jobs_total = 0, lens_jobs_total = 0;
    std::vector<int64_t> cam_job_start, bnd_job_start, lens_job_start;  // prefix (threads)
    std::vector<int32_t> cam_thread_job, bnd_thread_job, lens_thread_job;  // thread -> job

    // ---- per-marker column plan ---------------------------------------
    int kmax = 0;
    std::vector<PlanEntry> plan;  // [M * kmax]
    std::vector<int32_t> mkr_k;   // [M]
    bool any_lens = false;

    // ---- normal-equation partition (LM) --------------------------------
    // eliminated side: animated attrs grouped by frame; reduced side: static
    int na = 0;   // animated solver attrs (block size per frame)
    int ns = 0;   // static solver params (reduced system dimension)
    std::vector<int32_t> attr_slot;  // attr -> index within its class
    // solver param (reference order) <-> internal order
    // internal: [F*na eliminated (frame-major)] [ns reduced]
    std::vector<int32_t> par_to_int, int_to_par;

    std::string error;
};

// Build everything above from a scene description. Returns MMB_OK or error.
mmb_status build_host_problem(const mmb_scene_desc *s, const mmb_options *opt,
                              HostProblem *hp);

// findMarkerToAttributeRelationship (adjust_relationships.cpp:369-466) on the copied scene
// tables of hp, before any removal: table [M*A] (may be NULL), used flags
void marker_attr_usage(const HostProblem &hp, std::vector<uint8_t> *table, std::vector<uint8_t> *marker_used,
                       std::vector<uint8_t> *attr_used);
// copy + validate the scene tables only (what marker_attr_usage needs)
mmb_status copy_scene_tables(const mmb_scene_desc *s, const mmb_options *opt, HostProblem *hp);

// adjust_base.cpp:194-258 on the host (used for initial parameters / output)
double bound_internal_to_external(double v, double mn, double mx, double off, double scl);
double bound_external_to_internal(double v, double mn, double mx, double off, double scl);

}  // namespace mmb
