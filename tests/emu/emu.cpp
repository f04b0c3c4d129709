// TEST-ONLY host emulation of the CUDA per-thread bodies (eval_bodies.cuh).
// Lets the index/plan logic be debugged in the CPU container (no GPU here).
// Never linked into libmmsolver_b200.so; the product has no CPU path.
#include <vector>
#include <cstring>
#include "../../mayamatchmovesolver_b200/csrc/problem.cpp"
#include "../../mayamatchmovesolver_b200/csrc/eval_bodies.cuh"

using namespace mmb;

struct Emu {
    HostProblem hp;
    std::vector<uint8_t> plain;
    DevScene ds;
    std::vector<double> stat, anim, cam_tab, bnd_tab, lens_tab, da, db, cam_fwd;
    std::vector<int32_t> exa, exw, exv, exl;
    std::vector<int32_t> cnd, bnd, lnd;
};

extern "C" {
int emu_create(const mmb_scene_desc *s, const mmb_options *o, Emu **out) {
    Emu *e = new Emu();
    int st = build_host_problem(s, o, &e->hp);
    if (st != 0) { fprintf(stderr, "emu: %s\n", e->hp.error.c_str()); delete e; return st; }
    HostProblem &hp = e->hp; DevScene &d = e->ds;
    memset(&d, 0, sizeof(d));
    d.F = hp.F; d.T = hp.T; d.C = hp.C; d.B = hp.B; d.M = hp.M; d.L = hp.L; d.A = hp.A;
    d.n = hp.n; d.nobs = hp.nobs; d.kmax = std::max(hp.kmax, 1);
    d.central = o->auto_diff_type == 1; d.dense_grid = hp.dense_grid; d.any_lens = hp.any_lens;
    d.lens_literal = o->lens_index_literal; d.delta = o->delta; d.image_width = o->image_width;
    e->stat = hp.stat0; e->anim = hp.anim0; e->stat.resize(e->stat.size() + 1); e->anim.resize(e->anim.size() + 1);
    d.cam_thread_job = hp.cam_thread_job.data(); d.bnd_thread_job = hp.bnd_thread_job.data(); d.lens_thread_job = hp.lens_thread_job.data();
    e->plain.assign(std::max(hp.A, 1), 0); d.attr_plain = e->plain.data();  // emu: always the general transform
    d.stat_attr = hp.stat_attr.data(); d.anim_attr = hp.anim_attr.data();
    d.stat = e->stat.data(); d.anim = e->anim.data();
    d.tfm_parent = hp.tfm_parent.data(); d.tfm_roo = hp.tfm_roo.data(); d.tfm_attrs = hp.tfm_attrs.data();
    d.cam_tfm = hp.cam_tfm.data(); d.cam_attrs = hp.cam_attrs.data(); d.cam_fit = hp.cam_fit.data();
    d.cam_rw = hp.cam_rw.data(); d.cam_rh = hp.cam_rh.data(); d.cam_lens = hp.cam_lens.data();
    d.bnd_tfm = hp.bnd_tfm.data(); d.mkr_cam = hp.mkr_cam.data(); d.mkr_bnd = hp.mkr_bnd.data();
    d.lens_model = hp.lens_model.data(); d.lens_attrs = hp.lens_attrs.data(); d.lens_cam = hp.lens_cam.data();
    d.attr_ref = hp.attr_ref.data(); d.attr_param_base = hp.attr_param_base.data(); d.attr_anim = hp.attr_anim.data();
    d.attr_min = hp.attr_min.data(); d.attr_max = hp.attr_max.data(); d.attr_off = hp.attr_off.data(); d.attr_scl = hp.attr_scl.data();
    d.par_attr = hp.par_attr.data(); d.par_frame = hp.par_frame.data();
    d.err_marker = hp.err_marker.data(); d.err_frame = hp.err_frame.data();
    d.obs_mx = hp.obs_mx.data(); d.obs_my = hp.obs_my.data(); d.obs_w = hp.obs_w.data();
    e->cam_tab.assign(hp.cam_tab_size + 1, 0); e->bnd_tab.assign(hp.bnd_tab_size + 1, 0); e->lens_tab.assign(hp.lens_tab_size + 1, 0);
    d.cam_tab = e->cam_tab.data(); d.bnd_tab = e->bnd_tab.data(); d.lens_tab = e->lens_tab.data();
    d.cam_off = hp.cam_off.data(); d.bnd_off = hp.bnd_off.data(); d.lens_off = hp.lens_off.data();
    d.cam_fs = hp.cam_fs.data(); d.bnd_fs = hp.bnd_fs.data(); d.lens_fs = hp.lens_fs.data();
    for (auto &v : hp.cam_deps) e->cnd.push_back((int)v.size());
    for (auto &v : hp.bnd_deps) e->bnd.push_back((int)v.size());
    for (auto &v : hp.lens_deps) e->lnd.push_back((int)v.size());
    e->lnd.push_back(0);
    d.cam_ndep = e->cnd.data(); d.bnd_ndep = e->bnd.data(); d.lens_ndep = e->lnd.data();
    d.cam_job_ent = hp.cam_job_ent.data(); d.cam_job_var = hp.cam_job_var.data(); d.cam_job_attr = hp.cam_job_attr.data();
    d.cam_job_start = hp.cam_job_start.data(); d.cam_njobs = (int)hp.cam_job_ent.size(); d.cam_jobs_total = hp.cam_jobs_total;
    d.bnd_job_ent = hp.bnd_job_ent.data(); d.bnd_job_var = hp.bnd_job_var.data(); d.bnd_job_attr = hp.bnd_job_attr.data();
    d.bnd_job_start = hp.bnd_job_start.data(); d.bnd_njobs = (int)hp.bnd_job_ent.size(); d.bnd_jobs_total = hp.bnd_jobs_total;
    d.lens_job_ent = hp.lens_job_ent.data(); d.lens_job_var = hp.lens_job_var.data(); d.lens_job_attr = hp.lens_job_attr.data();
    d.lens_job_start = hp.lens_job_start.data(); d.lens_njobs = (int)hp.lens_job_ent.size(); d.lens_jobs_total = hp.lens_jobs_total;
    d.plan = (const PlanEntryD *)hp.plan.data(); d.mkr_k = hp.mkr_k.data();
    e->da.assign(hp.n + 1, 0); e->db.assign(hp.n + 1, 0);
    d.par_delta_a = e->da.data(); d.par_delta_b = e->db.data();
    d.maya_dag = hp.maya_dag; d.cur_frame = hp.cur_frame; d.n_extra = hp.n_extra;
    e->cam_fwd.assign(hp.cam_tab_size / 4 + 8, 0); d.cam_fwd = e->cam_fwd.data();
    e->exa = hp.ex_attr; e->exw = hp.ex_w; e->exv = hp.ex_var; e->exl = hp.ex_val;
    for (auto *v : {&e->exa, &e->exw, &e->exv, &e->exl}) v->push_back(0);
    d.ex_attr = e->exa.data(); d.ex_w = e->exw.data(); d.ex_var = e->exv.data(); d.ex_val = e->exl.data();
    d.lmdif = o->solver_type == 1;
    d.fd_eps = std::sqrt(std::max(std::fabs(o->delta), 2.220446049250313e-16));
    *out = e;
    return 0;
}
void emu_destroy(Emu *e) { delete e; }
void emu_sizes(Emu *e, int *n, int *nobs, int *kmax) { *n = e->hp.n; *nobs = e->hp.nobs; *kmax = e->ds.kmax; }
int emu_extra(Emu *e) { return e->hp.n_extra; }
static void prep(Emu *e, const double *x) {
    for (long long t = 0; t < e->ds.cam_jobs_total; ++t) camera_variant_thread(e->ds, x, t);
    for (long long t = 0; t < e->ds.bnd_jobs_total; ++t) bundle_variant_thread(e->ds, x, t);
    for (long long t = 0; t < e->ds.lens_jobs_total; ++t) lens_variant_thread(e->ds, x, t);
}
void emu_residuals(Emu *e, const double *x, double *fvec, double *dist) {
    prep(e, x);
    for (int i = 0; i < e->hp.nobs; ++i) residual_thread(e->ds, i, fvec, dist);
    for (int k = 0; k < e->hp.n_extra; ++k) extra_row_thread(e->ds, x, k, fvec + e->hp.m, nullptr, nullptr);
}
// the one Jacobian entry of every stiffness / smoothness row (after emu_jacobian at the same x)
void emu_extra_jacobian(Emu *e, const double *x, double *jext, int32_t *pid) {
    for (int k = 0; k < e->hp.n_extra; ++k) extra_row_thread(e->ds, x, k, nullptr, jext, pid);
}
void emu_jacobian(Emu *e, const double *x, double *fvec, double *dist, double *jac, int32_t *cols) {
    prep(e, x);
    for (int i = 0; i < e->hp.nobs; ++i) jacobian_thread(e->ds, i, fvec, dist, jac, cols);
}
}
