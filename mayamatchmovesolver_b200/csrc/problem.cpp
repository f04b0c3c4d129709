// Host C++: scene validation, reference index maps, variant plan.
// See problem.h.  Reference functions restated here:
//   countUpNumberOfErrors            adjust_relationships.cpp:75-221
//   countUpNumberOfUnknownParameters adjust_relationships.cpp:223-337
//   findErrorToParameterRelationship adjust_relationships.cpp:565-617
//     (kept implicit: the relationship is the per-marker column plan)
//   parameterBoundFromInternalToExternal / ExternalToInternal
//                                    adjust_base.cpp:194-258
#include "problem.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <thread>
#include <cmath>
#include <limits>
#include <set>
#include <map>
#include <mutex>

#ifdef MMB_BUILD   // the library; the test-only host emulation includes this file without CUDA
#include <cuda_runtime.h>
#endif

namespace mmb {

// ---- page-locked host blocks for the per-observation arrays (see NoInitAlloc) -------------------
#ifdef MMB_BUILD
namespace {
struct HostBlock { size_t bytes; bool pinned; bool in_use; };
std::mutex g_hb_mu;
std::map<void *, HostBlock> g_hb;           // every block handed out or cached
constexpr size_t kHbMin = 256 * 1024;       // smaller requests: plain malloc, not cached
constexpr size_t kHbCacheMax = (size_t)1 << 30;   // cached (idle) bytes kept at most
}  // namespace

void *host_block_acquire(size_t bytes) {
    if (bytes == 0) bytes = 1;
    if (bytes < kHbMin) {
        void *p = std::malloc(bytes);
        if (!p) throw std::bad_alloc();
        return p;
    }
    {
        std::lock_guard<std::mutex> lk(g_hb_mu);
        void *best = nullptr;
        size_t best_bytes = 0;
        for (auto &kv : g_hb)
            if (!kv.second.in_use && kv.second.bytes >= bytes && kv.second.bytes <= 2 * bytes && (!best || kv.second.bytes < best_bytes)) {
                best = kv.first; best_bytes = kv.second.bytes;
            }
        if (best) { g_hb[best].in_use = true; return best; }
    }
    void *p = nullptr;
    bool pinned = true;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0 || cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        pinned = false;
        p = std::malloc(bytes);
        if (!p) throw std::bad_alloc();
    }
    std::lock_guard<std::mutex> lk(g_hb_mu);
    g_hb[p] = HostBlock{bytes, pinned, true};
    return p;
}

void host_block_release(void *p) {
    if (!p) return;
    bool pinned = false, drop = false;
    {
        std::lock_guard<std::mutex> lk(g_hb_mu);
        auto it = g_hb.find(p);
        if (it == g_hb.end()) { drop = true; }   // a small (plain malloc) request
        else {
            size_t idle = 0;
            for (auto &kv : g_hb) if (!kv.second.in_use) idle += kv.second.bytes;
            if (!it->second.pinned || idle + it->second.bytes > kHbCacheMax) { pinned = it->second.pinned; g_hb.erase(it); drop = true; }
            else it->second.in_use = false;
        }
    }
    if (drop) { if (pinned) cudaFreeHost(p); else std::free(p); }
}

void host_block_trim() {
    std::vector<void *> dead;
    {
        std::lock_guard<std::mutex> lk(g_hb_mu);
        for (auto it = g_hb.begin(); it != g_hb.end();)
            if (!it->second.in_use && it->second.pinned) { dead.push_back(it->first); it = g_hb.erase(it); }
            else ++it;
    }
    for (void *p : dead) cudaFreeHost(p);
}
#else
void *host_block_acquire(size_t bytes) { void *p = std::malloc(bytes ? bytes : 1); if (!p) throw std::bad_alloc(); return p; }
void host_block_release(void *p) { std::free(p); }
void host_block_trim() {}
#endif

double bound_internal_to_external(double value, double xmin, double xmax,
                                  double offset, double scale) {
    const double float_max = std::numeric_limits<float>::max();
    if ((xmin <= -float_max) && (xmax >= float_max)) {
        value = (value / scale) - offset;
        value = std::max(value, xmin);
        value = std::min(value, xmax);
        return value;
    } else if (xmax >= float_max) {
        // lower bound only (formula as written at adjust_base.cpp:207)
        value = xmin - (1.0 + std::sqrt(value * value + 1.0));
    } else if (xmin <= -float_max) {
        value = xmax + (1.0 - std::sqrt(value * value + 1.0));
    } else {
        value = xmin + ((xmax - xmin) / 2.0) * (std::sin(value) + 1.0);
    }
    value = (value / scale) - offset;
    value = std::max(value, xmin);
    value = std::min(value, xmax);
    return value;
}

double bound_external_to_internal(double value, double xmin, double xmax,
                                  double offset, double scale) {
    value = std::max(value, xmin);
    value = std::min(value, xmax);
    value = (value * scale) + offset;
    xmin = (xmin * scale) + offset;
    xmax = (xmax * scale) + offset;
    const double float_max = std::numeric_limits<float>::max();
    // adjust_base.cpp:239 tests `xmin <= float_max` (sic); kept.
    if ((xmin <= float_max) && (xmax >= float_max)) {
        return value;
    } else if (xmax >= float_max) {
        value = std::sqrt(std::pow(((value - xmin) + 1.0), 2.0) - 1.0);
    } else if (xmin <= -float_max) {
        value = std::sqrt(std::pow((xmax - value) + 1.0, 2.0) - 1.0);
    } else {
        value = std::asin((2.0 * (value - xmin) / (xmax - xmin)) - 1.0);
    }
    return value;
}

static bool ref_ok(int32_t r, int ns, int na, bool allow_none) {
    if (r == MMB_ATTR_NONE) return allow_none;
    if (r >= 0) return r < ns;
    return (-(int64_t)r - 1) < na;
}

// refs read by a transform chain (leaf -> root)
static void chain_refs(const HostProblem &hp, int t, std::vector<int32_t> &out) {
    while (t >= 0) {
        for (int k = 0; k < MMB_TFM_ATTRS; ++k) out.push_back(hp.tfm_attrs[t * MMB_TFM_ATTRS + k]);
        t = hp.tfm_parent[t];
    }
}

// Solver attributes that write one of `refs` (sorted attr indices).  ref_attrs maps an
// attribute-store reference to the solver attributes naming it (duplicates allowed), so the
// cost is O(refs) per entity instead of a scan over all A attributes.
struct RefAttrs {
    std::vector<std::vector<int32_t>> stat, anim;
    const std::vector<int32_t> *find(int32_t r) const {
        if (r == MMB_ATTR_NONE) return nullptr;
        if (r >= 0) return (size_t)r < stat.size() ? &stat[r] : nullptr;
        const size_t i = (size_t)(-(int64_t)r - 1);
        return i < anim.size() ? &anim[i] : nullptr;
    }
};
static void deps_from_refs(const RefAttrs &ra, const std::vector<int32_t> &refs,
                           std::vector<int32_t> &deps, uint8_t &varies) {
    deps.clear();
    varies = 0;
    for (int32_t r : refs) {
        if (r != MMB_ATTR_NONE && r < 0) varies = 1;
        if (const std::vector<int32_t> *v = ra.find(r)) deps.insert(deps.end(), v->begin(), v->end());
    }
    std::sort(deps.begin(), deps.end());
    deps.erase(std::unique(deps.begin(), deps.end()), deps.end());
}

// marker x attribute "affects" table of the scene held in hp (tables already copied, before any
// removal): table[i * A + j] (may be NULL), plus which markers / attributes are used at all.
void marker_attr_usage(const HostProblem &hp, std::vector<uint8_t> *table, std::vector<uint8_t> *marker_used,
                       std::vector<uint8_t> *attr_used) {
    const int M = hp.M, A = hp.A;
    RefAttrs ra;
    ra.stat.resize(hp.num_static); ra.anim.resize(hp.num_anim);
    for (int a = 0; a < A; ++a) {
        const int32_t r = hp.attr_ref[a];
        if (r >= 0) ra.stat[r].push_back(a); else ra.anim[(size_t)(-(int64_t)r - 1)].push_back(a);
    }
    // attributes per camera (incl. its lens) and per bundle, once per entity
    std::vector<std::vector<int32_t>> cam_a(hp.C), bnd_a(hp.B);
    uint8_t varies;
    for (int c = 0; c < hp.C; ++c) {
        std::vector<int32_t> refs;
        chain_refs(hp, hp.cam_tfm[c], refs);
        for (int k = 0; k < MMB_CAM_ATTRS; ++k) refs.push_back(hp.cam_attrs[c * MMB_CAM_ATTRS + k]);
        const int l = hp.cam_lens[c];
        if (l >= 0) for (int k = 0; k < MMB_LENS_MAX_COEFFS; ++k) refs.push_back(hp.lens_attrs[l * MMB_LENS_MAX_COEFFS + k]);
        deps_from_refs(ra, refs, cam_a[c], varies);
    }
    for (int b = 0; b < hp.B; ++b) {
        std::vector<int32_t> refs;
        chain_refs(hp, hp.bnd_tfm[b], refs);
        deps_from_refs(ra, refs, bnd_a[b], varies);
    }
    if (table) table->assign((size_t)M * A, 0);
    if (marker_used) marker_used->assign(M, 0);
    if (attr_used) attr_used->assign(A, 0);
    for (int i = 0; i < M; ++i) {
        const std::vector<int32_t> *lists[2] = {&cam_a[hp.mkr_cam[i]], &bnd_a[hp.mkr_bnd[i]]};
        for (const std::vector<int32_t> *v : lists)
            for (int32_t a : *v) {
                if (table) (*table)[(size_t)i * A + a] = 1;
                if (marker_used) (*marker_used)[i] = 1;
                if (attr_used) (*attr_used)[a] = 1;
            }
    }
}

mmb_status copy_scene_tables(const mmb_scene_desc *s, const mmb_options *opt, HostProblem *hpp) {
    HostProblem &hp = *hpp;
    if (!s || !opt) return MMB_ERR_INVALID_ARGUMENT;
    if (s->abi_version != MMB_ABI_VERSION) {
        hp.error = "abi_version mismatch";
        return MMB_ERR_INVALID_ARGUMENT;
    }
    hp.opt = *opt;
    hp.F = s->num_frames; hp.T = s->num_transforms; hp.C = s->num_cameras;
    hp.B = s->num_bundles; hp.M = s->num_markers; hp.L = s->num_lenses;
    hp.A = s->num_attrs;
    hp.num_static = s->num_static; hp.num_anim = s->num_anim;
    if (hp.F <= 0 || hp.T <= 0 || hp.C <= 0 || hp.B <= 0 || hp.M <= 0 || hp.A < 0 ||
        hp.L < 0 || hp.num_static < 0 || hp.num_anim < 0) {
        // flat.rs:186-190 asserts the same counts are > 0
        hp.error = "scene needs >=1 frame, transform, camera, bundle and marker";
        return MMB_ERR_NOT_ENOUGH_OBJECTS;
    }
    if (opt->scene_graph_mode != 2 && opt->scene_graph_mode != 1) {
        hp.error = "sceneGraphMode must be 1 (Maya DAG conventions) or 2 (mmscenegraph)";
        return MMB_ERR_UNSUPPORTED;
    }
    hp.maya_dag = opt->scene_graph_mode == 1;
    if (hp.maya_dag && opt->frame_solve_mode == 1) {
        // solverstep.py:647-657: per-frame solving needs mmscenegraph
        hp.error = "frameSolveMode=1 (per frame) needs sceneGraphMode=2";
        return MMB_ERR_UNSUPPORTED;
    }
    if (opt->solver_type != 2 && opt->solver_type != 1) {
        hp.error = "only solverType=2 (cminpack_lmder) and 1 (cminpack_lmdif) are implemented";
        return MMB_ERR_UNSUPPORTED;
    }
    const size_t F = hp.F;
    if (s->frames) hp.frames.assign(s->frames, s->frames + F);
    else { hp.frames.resize(F); for (size_t f = 0; f < F; ++f) hp.frames[f] = (int32_t)f; }
    if (hp.num_static) hp.stat0.assign(s->static_values, s->static_values + hp.num_static);
    if (hp.num_anim) hp.anim0.assign(s->anim_values, s->anim_values + (size_t)hp.num_anim * F);
    hp.tfm_parent.assign(s->tfm_parent, s->tfm_parent + hp.T);
    hp.tfm_roo.assign(s->tfm_rotate_order, s->tfm_rotate_order + hp.T);
    hp.tfm_attrs.assign(s->tfm_attrs, s->tfm_attrs + hp.T * MMB_TFM_ATTRS);
    hp.cam_tfm.assign(s->cam_tfm, s->cam_tfm + hp.C);
    hp.cam_attrs.assign(s->cam_attrs, s->cam_attrs + hp.C * MMB_CAM_ATTRS);
    hp.cam_fit.assign(s->cam_film_fit, s->cam_film_fit + hp.C);
    hp.cam_rw.assign(s->cam_render_width, s->cam_render_width + hp.C);
    hp.cam_rh.assign(s->cam_render_height, s->cam_render_height + hp.C);
    hp.cam_lens.assign(hp.C, -1);
    if (s->cam_lens) hp.cam_lens.assign(s->cam_lens, s->cam_lens + hp.C);
    hp.bnd_tfm.assign(s->bnd_tfm, s->bnd_tfm + hp.B);
    hp.mkr_cam.assign(s->mkr_cam, s->mkr_cam + hp.M);
    hp.mkr_bnd.assign(s->mkr_bnd, s->mkr_bnd + hp.M);
    if (hp.L > 0) {
        hp.lens_model.assign(s->lens_model, s->lens_model + hp.L);
        hp.lens_attrs.assign(s->lens_attrs, s->lens_attrs + hp.L * MMB_LENS_MAX_COEFFS);
        hp.lens_cam.assign(s->lens_camera_params, s->lens_camera_params + hp.L * 6);
    }
    if (hp.A) {
        hp.attr_ref.assign(s->attr_ref, s->attr_ref + hp.A);
        hp.attr_min.assign(s->attr_min, s->attr_min + hp.A);
        hp.attr_max.assign(s->attr_max, s->attr_max + hp.A);
        hp.attr_off.assign(s->attr_offset, s->attr_offset + hp.A);
        hp.attr_scl.assign(s->attr_scale, s->attr_scale + hp.A);
    }
    // ---- validation ------------------------------------------------------
    for (int t = 0; t < hp.T; ++t) {
        if (hp.tfm_parent[t] >= t || hp.tfm_parent[t] < -1) { hp.error = "tfm_parent must precede child"; return MMB_ERR_INVALID_ARGUMENT; }
        if (hp.tfm_roo[t] < 0 || hp.tfm_roo[t] > 5) { hp.error = "bad rotate order"; return MMB_ERR_INVALID_ARGUMENT; }
        for (int k = 0; k < MMB_TFM_ATTRS; ++k)
            if (!ref_ok(hp.tfm_attrs[t * MMB_TFM_ATTRS + k], hp.num_static, hp.num_anim, false)) { hp.error = "bad transform attr ref"; return MMB_ERR_INVALID_ARGUMENT; }
    }
    for (int c = 0; c < hp.C; ++c) {
        if (hp.cam_tfm[c] < 0 || hp.cam_tfm[c] >= hp.T) { hp.error = "bad cam_tfm"; return MMB_ERR_INVALID_ARGUMENT; }
        if (hp.cam_lens[c] >= hp.L) { hp.error = "bad cam_lens"; return MMB_ERR_INVALID_ARGUMENT; }
        if (hp.cam_fit[c] < 0 || hp.cam_fit[c] > 3) { hp.error = "bad film fit"; return MMB_ERR_INVALID_ARGUMENT; }
        for (int k = 0; k < MMB_CAM_ATTRS; ++k)
            if (!ref_ok(hp.cam_attrs[c * MMB_CAM_ATTRS + k], hp.num_static, hp.num_anim, false)) { hp.error = "bad camera attr ref"; return MMB_ERR_INVALID_ARGUMENT; }
    }
    for (int b = 0; b < hp.B; ++b)
        if (hp.bnd_tfm[b] < 0 || hp.bnd_tfm[b] >= hp.T) { hp.error = "bad bnd_tfm"; return MMB_ERR_INVALID_ARGUMENT; }
    for (int m = 0; m < hp.M; ++m)
        if (hp.mkr_cam[m] < 0 || hp.mkr_cam[m] >= hp.C || hp.mkr_bnd[m] < 0 || hp.mkr_bnd[m] >= hp.B) { hp.error = "bad marker camera/bundle index"; return MMB_ERR_INVALID_ARGUMENT; }
    for (int l = 0; l < hp.L; ++l) {
        if (hp.lens_model[l] < 0 || hp.lens_model[l] > 4) { hp.error = "bad lens model"; return MMB_ERR_INVALID_ARGUMENT; }
        for (int k = 0; k < MMB_LENS_MAX_COEFFS; ++k)
            if (!ref_ok(hp.lens_attrs[l * MMB_LENS_MAX_COEFFS + k], hp.num_static, hp.num_anim, true)) { hp.error = "bad lens attr ref"; return MMB_ERR_INVALID_ARGUMENT; }
    }
    for (int a = 0; a < hp.A; ++a)
        if (!ref_ok(hp.attr_ref[a], hp.num_static, hp.num_anim, false)) { hp.error = "bad solver attr ref"; return MMB_ERR_INVALID_ARGUMENT; }

    return MMB_OK;
}

mmb_status build_host_problem(const mmb_scene_desc *s, const mmb_options *opt,
                              HostProblem *hpp) {
    HostProblem &hp = *hpp;
    const bool timing = getenv("MMB_TIMING") != nullptr;
    auto t_prev = std::chrono::steady_clock::now();
    auto mark = [&](const char *what) {
        if (!timing) return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[mmb timing]   host: %s %.3f ms\n", what, 1e3 * std::chrono::duration<double>(now - t_prev).count());
        t_prev = now;
    };
    {
        const mmb_status cs = copy_scene_tables(s, opt, hpp);
        if (cs != MMB_OK) return cs;
    }
    const size_t F = hp.F;
    hp.cur_frame = s->current_frame;
    if (hp.cur_frame < 0 || hp.cur_frame >= hp.F) { hp.error = "current_frame outside the frame list"; return MMB_ERR_INVALID_ARGUMENT; }
    hp.stiff_in.clear(); hp.smooth_in.clear();
    for (int pass = 0; pass < 2; ++pass) {
        const int cnt = pass ? s->num_smoothness : s->num_stiffness;
        const int32_t *a = pass ? s->smooth_attr : s->stiff_attr, *w = pass ? s->smooth_weight : s->stiff_weight;
        const int32_t *vr = pass ? s->smooth_variance : s->stiff_variance, *vl = pass ? s->smooth_value : s->stiff_value;
        if (cnt < 0 || (cnt > 0 && (!a || !w || !vr || !vl))) { hp.error = "bad stiffness / smoothness table"; return MMB_ERR_INVALID_ARGUMENT; }
        std::vector<int32_t> &dst = pass ? hp.smooth_in : hp.stiff_in;
        for (int i = 0; i < cnt; ++i) {
            if (!ref_ok(w[i], hp.num_static, hp.num_anim, false) || !ref_ok(vr[i], hp.num_static, hp.num_anim, false) ||
                !ref_ok(vl[i], hp.num_static, hp.num_anim, false)) { hp.error = "bad stiffness / smoothness attr ref"; return MMB_ERR_INVALID_ARGUMENT; }
            dst.insert(dst.end(), {a[i], w[i], vr[i], vl[i]});
        }
    }
    if (hp.maya_dag)
        for (int c = 0; c < hp.C; ++c)
            if (hp.cam_attrs[c * MMB_CAM_ATTRS + 7] < 0) {
                // Camera::getProjMatrix reads cameraScale without a time (maya_camera.cpp:790)
                hp.error = "sceneGraphMode=1: an animated cameraScale is not supported";
                return MMB_ERR_UNSUPPORTED;
            }
    mark("validation");
    // ---- used / unused solve objects (solve_v1, adjust_base.cpp:1322-1395) ------------------
    // markerToAttrList: marker i is affected by attribute j when j drives the marker's camera
    // (transform chain, camera shape, lens) or its bundle's transform chain -- what
    // findMarkerToAttributeRelationship (adjust_relationships.cpp:369-466) asks Maya's DG.
    // splitUsedMarkersAndAttributes (adjust_base.cpp:574-610): a marker affected by no attribute
    // and an attribute affecting no marker are dropped from the solve when the
    // removeUnused* options say so; the index maps then number the USED lists.
    hp.M_in = hp.M; hp.A_in = hp.A;
    hp.marker_used.assign(hp.M, 1); hp.attr_used.assign(hp.A, 1);
    hp.mkr_src.resize(hp.M);
    for (int i = 0; i < hp.M; ++i) hp.mkr_src[i] = i;
    hp.attr_src.resize(hp.A);
    for (int a = 0; a < hp.A; ++a) hp.attr_src[a] = a;
    if (opt->remove_unused_markers || opt->remove_unused_attributes) {
        std::vector<uint8_t> m_used, a_used;
        marker_attr_usage(hp, nullptr, &m_used, &a_used);
        hp.marker_used = m_used; hp.attr_used = a_used;
        if (opt->remove_unused_markers) {
            int k = 0;
            for (int i = 0; i < hp.M; ++i)
                if (m_used[i]) { hp.mkr_cam[k] = hp.mkr_cam[i]; hp.mkr_bnd[k] = hp.mkr_bnd[i]; hp.mkr_src[k] = i; ++k; }
            hp.M = k;
            hp.mkr_cam.resize(k); hp.mkr_bnd.resize(k); hp.mkr_src.resize(k);
        } else {
            hp.marker_used.assign(hp.M, 1);
        }
        if (opt->remove_unused_attributes) {
            int k = 0;
            for (int a = 0; a < hp.A; ++a)
                if (a_used[a]) {
                    hp.attr_ref[k] = hp.attr_ref[a]; hp.attr_min[k] = hp.attr_min[a]; hp.attr_max[k] = hp.attr_max[a];
                    hp.attr_off[k] = hp.attr_off[a]; hp.attr_scl[k] = hp.attr_scl[a]; hp.attr_src[k] = a; ++k;
                }
            hp.A = k;
            hp.attr_ref.resize(k); hp.attr_min.resize(k); hp.attr_max.resize(k); hp.attr_off.resize(k); hp.attr_scl.resize(k);
            hp.attr_src.resize(k);
        } else {
            hp.attr_used.assign(hp.A, 1);
        }
    }
    mark("used / unused objects");
    // ---- countUpNumberOfErrors (marker-major, frame-minor) ---------------
    std::vector<double> wmax(F, 0.0);
    std::vector<uint8_t> wseen(F, 0);
    {
        // Two passes over marker ranges on a few host threads (the scan is 5e5 .. 5e7 entries):
        // count the enabled observations per range, then fill at the prefix offsets.  The
        // order (marker-major, frame-minor) is the reference's.
        const size_t M = (size_t)hp.M;
        // host threads of the scan: the cores of the box divided among the ranks that share it (one
        // process per GPU: LOCAL_WORLD_SIZE as torchrun / mpirun export it), MMB_HOST_THREADS overrides
        unsigned hw = std::max(1u, std::thread::hardware_concurrency());
        if (const char *e = getenv("MMB_HOST_THREADS")) hw = (unsigned)std::max(1, atoi(e));
        else if (const char *lw = getenv("LOCAL_WORLD_SIZE")) hw = std::max(1u, hw / (unsigned)std::max(1, atoi(lw)));
        static int tcap = -1;
        if (tcap < 0) { const char *e = getenv("MMB_SCAN_THREADS_MAX"); tcap = e ? std::max(1, atoi(e)) : 8; }
        const int T = (int)std::max<size_t>(1, std::min<size_t>({(size_t)hw, (size_t)tcap, (M * F) / 65536 + 1}));
        std::vector<size_t> count(T + 1, 0);
        std::vector<std::vector<double>> twmax(T, std::vector<double>(F, 0.0));
        std::vector<std::vector<uint8_t>> tseen(T, std::vector<uint8_t>(F, 0));
        auto range = [&](int t, size_t *m0, size_t *m1) { *m0 = M * t / T; *m1 = M * (t + 1) / T; };
        auto enabled = [&](size_t o) { return s->mkr_enable[o] != 0 && (s->mkr_weight[o] > 0.0); };
        const std::vector<int32_t> &src = hp.mkr_src;  // row of marker i in the caller's tables
        // one set of threads for both passes: count, meet, (thread 0: prefix sums and the
        // uninitialised resize), meet, fill
        std::atomic<int> arrived{0}, released{0}, filled{0}, merged{0};
        auto body = [&](int t) {
            size_t m0, m1, k = 0;
            range(t, &m0, &m1);
            for (size_t i = m0; i < m1; ++i)
                for (size_t o = (size_t)src[i] * F, e = o + F; o < e; ++o) k += enabled(o) ? 1 : 0;
            count[t + 1] = k;
            arrived.fetch_add(1, std::memory_order_acq_rel);
            if (t == 0) {
                while (arrived.load(std::memory_order_acquire) < T) std::this_thread::yield();
                for (int u = 0; u < T; ++u) count[u + 1] += count[u];
                const size_t total = count[T];
                hp.err_marker.resize(total); hp.err_frame.resize(total);
                hp.obs_mx.resize(total); hp.obs_my.resize(total); hp.obs_w.resize(total);
                released.store(1, std::memory_order_release);
            } else {
                while (!released.load(std::memory_order_acquire)) std::this_thread::yield();
            }
            k = count[t];
            std::vector<double> &wm = twmax[t];
            std::vector<uint8_t> &ws = tseen[t];
            for (size_t i = m0; i < m1; ++i) {
                const size_t o0 = (size_t)src[i] * F;
                for (size_t j = 0; j < F; ++j) {
                    const size_t o = o0 + j;
                    if (!enabled(o)) continue;
                    const double weight = s->mkr_weight[o];
                    hp.err_marker[k] = (int32_t)i;
                    hp.err_frame[k] = (int32_t)j;
                    hp.obs_w[k] = weight;
                    hp.obs_mx[k] = s->mkr_x[o];
                    hp.obs_my[k] = s->mkr_y[o];
                    ++k;
                    if (!ws[j] || weight > wm[j]) wm[j] = weight;
                    ws[j] = 1;
                }
            }
            // meet again; thread 0 merges the per-frame weight maxima; then every thread normalises
            // the weights of its own observations (adjust_relationships.cpp: weight / max of the frame)
            filled.fetch_add(1, std::memory_order_acq_rel);
            if (t == 0) {
                while (filled.load(std::memory_order_acquire) < T) std::this_thread::yield();
                for (int u = 0; u < T; ++u)
                    for (size_t j = 0; j < F; ++j)
                        if (tseen[u][j] && (!wseen[j] || twmax[u][j] > wmax[j])) { wmax[j] = twmax[u][j]; wseen[j] = 1; }
                merged.store(1, std::memory_order_release);
            } else {
                while (!merged.load(std::memory_order_acquire)) std::this_thread::yield();
            }
            for (size_t q = count[t]; q < count[t + 1]; ++q) hp.obs_w[q] = hp.obs_w[q] / wmax[hp.err_frame[q]];
        };
        {
            std::vector<std::thread> th;
            for (int t = 1; t < T; ++t) th.emplace_back(body, t);
            body(0);
            for (auto &x : th) x.join();
        }
    }
    hp.nobs = (int)hp.err_marker.size();
    hp.m = 2 * hp.nobs;
    // stiffness / smoothness rows (adjust_relationships.cpp:183-219): one per entry whose weight,
    // read at the current time, is > 0; measureErrors then uses the FIRST that many entries of
    // each list (adjust_measureErrors.cpp:318-323), which is kept
    {
        auto value_of = [&](int32_t r) { return r >= 0 ? hp.stat0[r] : hp.anim0[(size_t)(-(int64_t)r - 1) * F + hp.cur_frame]; };
        hp.n_stiff = hp.n_smooth = 0;
        for (size_t i = 0; i + 3 < hp.stiff_in.size(); i += 4) if (value_of(hp.stiff_in[i + 1]) > 0.0) ++hp.n_stiff;
        for (size_t i = 0; i + 3 < hp.smooth_in.size(); i += 4) if (value_of(hp.smooth_in[i + 1]) > 0.0) ++hp.n_smooth;
        hp.ex_attr.clear(); hp.ex_w.clear(); hp.ex_var.clear(); hp.ex_val.clear();
        for (int pass = 0; pass < 2; ++pass) {
            const std::vector<int32_t> &src = pass ? hp.smooth_in : hp.stiff_in;
            const int cnt = pass ? hp.n_smooth : hp.n_stiff;
            for (int i = 0; i < cnt; ++i) {
                const int32_t a = src[4 * i];
                if (a < 0 || a >= hp.A) { hp.error = "stiffness / smoothness attribute index outside the used attribute list"; return MMB_ERR_INVALID_ARGUMENT; }
                hp.ex_attr.push_back(a); hp.ex_w.push_back(src[4 * i + 1]);
                hp.ex_var.push_back(src[4 * i + 2]); hp.ex_val.push_back(src[4 * i + 3]);
            }
        }
        hp.n_extra = hp.n_stiff + hp.n_smooth;
        hp.m_total = hp.m + hp.n_extra;
    }
    hp.dense_grid = ((size_t)hp.nobs == (size_t)hp.M * F);

    mark("observation scan");
    // ---- countUpNumberOfUnknownParameters (attr-major, frame-minor) -------
    hp.par_attr.clear(); hp.par_frame.clear();
    hp.attr_param_base.assign(hp.A, 0);
    hp.attr_anim.assign(hp.A, 0);
    hp.attr_slot.assign(hp.A, 0);
    hp.na = 0; hp.ns = 0;
    for (int a = 0; a < hp.A; ++a) {
        hp.attr_param_base[a] = (int32_t)hp.par_attr.size();
        if (hp.attr_ref[a] < 0) {
            hp.attr_anim[a] = 1;
            hp.attr_slot[a] = hp.na++;
            for (size_t j = 0; j < F; ++j) { hp.par_attr.push_back(a); hp.par_frame.push_back((int32_t)j); }
        } else {
            hp.attr_slot[a] = hp.ns++;
            hp.par_attr.push_back(a); hp.par_frame.push_back(-1);
        }
    }
    hp.n = (int)hp.par_attr.size();
    // later duplicates of an attribute win, as the last setParameters write would
    hp.stat_attr.assign(std::max(hp.num_static, 1), -1);
    hp.anim_attr.assign(std::max(hp.num_anim, 1), -1);
    for (int a = 0; a < hp.A; ++a) {
        if (hp.attr_ref[a] >= 0) hp.stat_attr[hp.attr_ref[a]] = a;
        else hp.anim_attr[-hp.attr_ref[a] - 1] = a;
    }
    hp.par_to_int.assign(hp.n, 0); hp.int_to_par.assign(hp.n, 0);
    for (int p = 0; p < hp.n; ++p) {
        const int a = hp.par_attr[p];
        const int q = hp.attr_anim[a] ? hp.par_frame[p] * hp.na + hp.attr_slot[a]
                                      : (int)F * hp.na + hp.attr_slot[a];
        hp.par_to_int[p] = q; hp.int_to_par[q] = p;
    }

    mark("parameters");
    // ---- entity dependencies ---------------------------------------------
    hp.cam_deps.assign(hp.C, {}); hp.bnd_deps.assign(hp.B, {}); hp.lens_deps.assign(hp.L, {});
    hp.cam_varies.assign(hp.C, 0); hp.bnd_varies.assign(hp.B, 0); hp.lens_varies.assign(hp.L, 0);
    hp.cam_mscale_varies.assign(hp.C, 0);
    RefAttrs ra;
    ra.stat.resize(hp.num_static); ra.anim.resize(hp.num_anim);
    for (int a = 0; a < hp.A; ++a) {
        const int32_t r = hp.attr_ref[a];
        if (r >= 0) ra.stat[r].push_back(a); else ra.anim[(size_t)(-(int64_t)r - 1)].push_back(a);
    }
    for (int c = 0; c < hp.C; ++c) {
        std::vector<int32_t> refs;
        chain_refs(hp, hp.cam_tfm[c], refs);
        for (int k = 0; k < MMB_CAM_ATTRS; ++k) refs.push_back(hp.cam_attrs[c * MMB_CAM_ATTRS + k]);
        deps_from_refs(ra, refs, hp.cam_deps[c], hp.cam_varies[c]);
        for (int a : hp.cam_deps[c])
            if (hp.attr_ref[a] == hp.cam_attrs[c * MMB_CAM_ATTRS + 0] || hp.attr_ref[a] == hp.cam_attrs[c * MMB_CAM_ATTRS + 1])
                hp.cam_mscale_varies[c] = 1;
    }
    for (int b = 0; b < hp.B; ++b) {
        std::vector<int32_t> refs;
        chain_refs(hp, hp.bnd_tfm[b], refs);
        deps_from_refs(ra, refs, hp.bnd_deps[b], hp.bnd_varies[b]);
    }
    hp.any_lens = false;
    for (int l = 0; l < hp.L; ++l) {
        std::vector<int32_t> refs;
        for (int k = 0; k < MMB_LENS_MAX_COEFFS; ++k) refs.push_back(hp.lens_attrs[l * MMB_LENS_MAX_COEFFS + k]);
        deps_from_refs(ra, refs, hp.lens_deps[l], hp.lens_varies[l]);
    }
    for (int c = 0; c < hp.C; ++c)
        if (hp.cam_lens[c] >= 0 && hp.lens_model[hp.cam_lens[c]] != MMB_LENS_PASSTHROUGH) hp.any_lens = true;

    mark("dependencies");
    // ---- variant tables ---------------------------------------------------
    const int vmul = (opt->auto_diff_type == 1) ? 2 : 1;
    auto layout = [&](const std::vector<std::vector<int32_t>> &deps, const std::vector<uint8_t> &varies,
                      int width, std::vector<int64_t> &off, std::vector<int32_t> &fs,
                      std::vector<int32_t> &nvar, int64_t &total, std::vector<int32_t> &job_ent,
                      std::vector<int32_t> &job_var, std::vector<int32_t> &job_attr,
                      std::vector<int64_t> &job_start, int64_t &jobs_total, int frame_block) {
        const int E = (int)deps.size();
        off.assign(E, 0); fs.assign(E, 1); nvar.assign(E, 1);
        job_ent.clear(); job_var.clear(); job_attr.clear(); job_start.clear();
        total = 0; jobs_total = 0;
        for (int e = 0; e < E; ++e) {
            // a variant of a static entity perturbed by an animated param
            // would differ per frame; `varies` already covers that because an
            // animated solver attr is an animated ref in the chain.
            fs[e] = varies[e] ? (int32_t)F : 1;
            nvar[e] = 1 + vmul * (int)deps[e].size();
            off[e] = total;
            // animated records may be stored in whole frame blocks (cam_idx, eval_bodies.cuh)
            const int64_t fpad = fs[e] == 1 ? 1 : ((int64_t)fs[e] + frame_block - 1) / frame_block * frame_block;
            total += (int64_t)nvar[e] * width * fpad;
            for (int v = 0; v < nvar[e]; ++v) {
                job_ent.push_back(e); job_var.push_back(v);
                job_attr.push_back(v == 0 ? -1 : deps[e][(v - 1) % deps[e].size()]);
                job_start.push_back(jobs_total);
                jobs_total += fs[e];
            }
        }
        job_start.push_back(jobs_total);
    };
    layout(hp.cam_deps, hp.cam_varies, 16, hp.cam_off, hp.cam_fs, hp.cam_nvar, hp.cam_tab_size,
           hp.cam_job_ent, hp.cam_job_var, hp.cam_job_attr, hp.cam_job_start, hp.cam_jobs_total, 32);
    layout(hp.bnd_deps, hp.bnd_varies, 4, hp.bnd_off, hp.bnd_fs, hp.bnd_nvar, hp.bnd_tab_size,
           hp.bnd_job_ent, hp.bnd_job_var, hp.bnd_job_attr, hp.bnd_job_start, hp.bnd_jobs_total, 1);
    layout(hp.lens_deps, hp.lens_varies, 24, hp.lens_off, hp.lens_fs, hp.lens_nvar, hp.lens_tab_size,
           hp.lens_job_ent, hp.lens_job_var, hp.lens_job_attr, hp.lens_job_start, hp.lens_jobs_total, 1);

    auto expand = [](const std::vector<int64_t> &start, std::vector<int32_t> &out) {
        out.assign((size_t)start.back(), 0);
        for (size_t j = 0; j + 1 < start.size(); ++j)
            for (int64_t t = start[j]; t < start[j + 1]; ++t) out[(size_t)t] = (int32_t)j;
    };
    expand(hp.cam_job_start, hp.cam_thread_job);
    expand(hp.bnd_job_start, hp.bnd_thread_job);
    expand(hp.lens_job_start, hp.lens_thread_job);

    mark("variant tables");
    // ---- per-marker column plan -------------------------------------------
    hp.mkr_k.assign(hp.M, 0);
    std::vector<std::vector<PlanEntry>> plans(hp.M);
    hp.kmax = 0;
    for (int mk = 0; mk < hp.M; ++mk) {
        const int c = hp.mkr_cam[mk], b = hp.mkr_bnd[mk], l = hp.cam_lens[c];
        std::set<int32_t> cols(hp.cam_deps[c].begin(), hp.cam_deps[c].end());
        cols.insert(hp.bnd_deps[b].begin(), hp.bnd_deps[b].end());
        if (l >= 0) cols.insert(hp.lens_deps[l].begin(), hp.lens_deps[l].end());
        auto find1 = [](const std::vector<int32_t> &v, int32_t a) -> int16_t {
            auto it = std::lower_bound(v.begin(), v.end(), a);
            return (it != v.end() && *it == a) ? (int16_t)(it - v.begin() + 1) : (int16_t)0;
        };
        for (int32_t a : cols) {
            PlanEntry e;
            e.attr = a;
            e.cam_var = find1(hp.cam_deps[c], a);
            e.bnd_var = find1(hp.bnd_deps[b], a);
            e.lens_var = l >= 0 ? find1(hp.lens_deps[l], a) : (int16_t)0;
            e.pad = 0;
            plans[mk].push_back(e);
        }
        hp.mkr_k[mk] = (int32_t)plans[mk].size();
        hp.kmax = std::max(hp.kmax, hp.mkr_k[mk]);
    }
    hp.plan.assign((size_t)hp.M * std::max(hp.kmax, 1), PlanEntry{-1, 0, 0, 0, 0});
    for (int mk = 0; mk < hp.M; ++mk)
        for (size_t c = 0; c < plans[mk].size(); ++c) hp.plan[(size_t)mk * hp.kmax + c] = plans[mk][c];
    mark("column plans");
    return MMB_OK;
}

}  // namespace mmb
