// api.cu -- the C ABI of libcvrecon_b200.so (include/cvrecon_b200.h): context, workspace, host<->device staging
// and the per-stage / batched entry points.  No CPU fallback exists: without a usable sm_100-class device
// cvr_create fails.
#include "cvr_internal.cuh"
#include <string.h>
#include <stdio.h>
#include <stdlib.h>
#include <new>
#include <nvtx3/nvToolsExt.h>   // header-only; a no-op unless a profiler injects its library

static const char* kVersion = "cvrecon_b200 0.1 (sm_100a)";
static thread_local std::string g_create_err;

extern "C" const char* cvr_version(void) { return kVersion; }
void cvr_set_thread_error(const char* msg) { g_create_err = msg ? msg : ""; }

extern "C" void cvr_default_params(cvr_params* p, int width, int height)
{
    memset(p, 0, sizeof(*p));
    p->width = width; p->height = height;
    // config_default.json:13-21 + Reconstruct3D.cpp:63-64
    p->sgbm.block_size = 27; p->sgbm.num_disparities = 128; p->sgbm.uniqueness_ratio = 10; p->sgbm.speckle_range = 2;
    p->sgbm.speckle_window_size = 100; p->sgbm.pre_filter_cap = 10; p->sgbm.min_disparity = 0;
    p->sgbm.P1 = 8 * 3 * 27 * 27; p->sgbm.P2 = 32 * 3 * 27 * 27; p->sgbm.disp12_max_diff = 0; p->sgbm.mode = CVR_MODE_SGBM;
    // Reconstruct3D.cpp:31-35 with resources/calib/kitti/calib_kitti_rectified.json
    const float fl = 959.791f, fr = 903.7596f, fyl = 956.9251f;
    for (int i = 0; i < 16; i++) p->Q[i] = (i % 5 == 0) ? 1.0f : 0.0f;
    const float f = (fl + fr) / 2.0f;
    p->Q[10] = (float)(0.005 * (double)f);
    p->tri_f = (fl + fyl) / 2.0f; p->tri_b = 0.532671f; p->tri_cx = 696.0217f; p->tri_cy = 224.1806f;
    p->prune_sigma_factor = 1.3;
    p->voxel_leaf[0] = p->voxel_leaf[1] = p->voxel_leaf[2] = 0.2f;
    p->sor_mean_k = 30; p->sor_stddev_mul = 0.5;
    p->max_batch = 1; p->cloud_mode = CVR_CLOUD_TRIANGULATE; p->max_points = 0; p->filter_only = 0;
    // config_default.json:8-12 ("block_matcher": "stereo_bm" is the file's default; this library's default stays SGBM, the path
    // named by the north star) + cv::StereoBM::create(16, 21) defaults
    p->block_matcher = CVR_MATCHER_SGBM; p->bm_window_size = 27; p->bm_num_disparities = 128; p->bm_min_disparity = 0;
    p->bm_pre_filter_cap = 31; p->bm_texture_threshold = 10; p->bm_uniqueness_ratio = 15; p->bm_speckle_window_size = 0; p->bm_speckle_range = 0;
    p->requires_rectification = 0;
}

void cvr_stage_mark(cvr_ctx* c, const char* name)
{
    if (!c->profiling) return;
    std::vector<StageTimer>& tv = c->timers[c->tset];
    int& used = c->timers_used[c->tset];
    if (used >= (int)tv.size()) {
        StageTimer t; t.name = name; cudaEventCreate(&t.ev);
        tv.push_back(t);
    }
    tv[used].name = name;
    cudaEventRecord(tv[used].ev, c->stream);
    nvtxMarkA(name);                                                   // the end of the stage's enqueue, for profiler timelines
    used++;
}
// ---- tail stream of the pipelined batches (see cvr_ctx::tail_stream) ----
int cvr_tail_split(cvr_ctx* c)
{
    if (!c->split_tail || c->path_stream) return CVR_OK;
    CVR_CUDA(c, cudaEventRecord(c->evPath, c->stream));
    CVR_CUDA(c, cudaStreamWaitEvent(c->tail_stream, c->evPath, 0));
    c->path_stream = c->stream; c->stream = c->tail_stream;           // every launcher enqueues on c->stream
    return CVR_OK;
}
int cvr_tail_aux_consumed(cvr_ctx* c)
{
    if (!c->split_tail || !c->path_stream) return CVR_OK;
    CVR_CUDA(c, cudaEventRecord(c->evAuxFree, c->stream));
    c->aux_pending = true;
    return CVR_OK;
}
int cvr_path_wait_aux(cvr_ctx* c)
{
    if (!c->aux_pending) return CVR_OK;
    CVR_CUDA(c, cudaStreamWaitEvent(c->stream, c->evAuxFree, 0));
    c->aux_pending = false;
    return CVR_OK;
}
// Entry points other than cvr_submit_batch work on c->stream alone: whatever a pipelined batch still has on the tail stream
// comes first (they share the workspace).
int cvr_join_tail(cvr_ctx* c)
{
    if (!c->tail_dirty) return CVR_OK;
    CVR_CUDA(c, cudaStreamWaitEvent(c->stream, c->evTailLast, 0));
    c->tail_dirty = false; c->aux_pending = false;
    return CVR_OK;
}
void cvr_timers_begin(cvr_ctx* c, int set) { c->tset = set; c->timers_used[set] = 0; cvr_stage_mark(c, "begin"); }
void cvr_timers_end(cvr_ctx* c, int set)                           // after the events of that set have completed
{
    c->t_names.clear(); c->t_ms.clear();
    if (!c->profiling) return;
    for (int i = 1; i < c->timers_used[set]; i++) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, c->timers[set][i - 1].ev, c->timers[set][i].ev);
        c->t_names.push_back(c->timers[set][i].name); c->t_ms.push_back(ms);
    }
    c->timers_used[set] = 0;
}

template <typename T>
static cudaError_t dalloc(T** p, size_t count) { *p = nullptr; if (count == 0) count = 1; return cudaMalloc((void**)p, count * sizeof(T)); }

static int validate(const cvr_params* p, std::string& err)
{
    if (!p || p->width <= 0 || p->height <= 0) { err = "width/height must be positive"; return CVR_ERR_INVALID_ARG; }
    if (p->width > 65535) { err = "width > 65535 unsupported"; return CVR_ERR_UNSUPPORTED; }
    if (p->max_batch < 1) { err = "max_batch must be >= 1"; return CVR_ERR_INVALID_ARG; }
    if (p->filter_only) return CVR_OK;
    if (p->block_matcher == CVR_MATCHER_BM) {
        // the checks of cv::StereoBM::compute
        if (p->bm_pre_filter_cap < 1 || p->bm_pre_filter_cap > 63) { err = "bm_pre_filter_cap must be within 1..63"; return CVR_ERR_INVALID_ARG; }
        if (p->bm_window_size < 5 || p->bm_window_size > 255 || p->bm_window_size % 2 == 0 || p->bm_window_size >= (p->width < p->height ? p->width : p->height)) {
            err = "bm_window_size must be odd, within 5..255 and smaller than the image"; return CVR_ERR_INVALID_ARG; }
        if (p->bm_num_disparities <= 0 || p->bm_num_disparities % 16 != 0) { err = "bm_num_disparities must be a positive multiple of 16"; return CVR_ERR_INVALID_ARG; }
        if (p->bm_num_disparities > 256) { err = "bm_num_disparities > 256 unsupported"; return CVR_ERR_UNSUPPORTED; }
        if (p->bm_texture_threshold < 0 || p->bm_uniqueness_ratio < 0) { err = "bm texture threshold / uniqueness ratio must be non-negative"; return CVR_ERR_INVALID_ARG; }
        if (p->bm_min_disparity > 0) { err = "bm_min_disparity > 0 unsupported (the reference never sets it; OpenCV's own loop overruns the row there)"; return CVR_ERR_UNSUPPORTED; }
        return CVR_OK;
    } else if (p->block_matcher != CVR_MATCHER_SGBM) { err = "block_matcher must be CVR_MATCHER_SGBM or CVR_MATCHER_BM"; return CVR_ERR_INVALID_ARG; }
    const cvr_sgbm_params& s = p->sgbm;
    if (p->width > 16384) { err = "width > 16384 unsupported by the SGBM matcher (the LR check keeps a row's right-view map in shared memory)"; return CVR_ERR_UNSUPPORTED; }
    if (s.num_disparities <= 0 || s.num_disparities % 16 != 0) { err = "num_disparities must be a positive multiple of 16 (OpenCV rejects others)"; return CVR_ERR_INVALID_ARG; }
    if (s.num_disparities > 256) { err = "num_disparities > 256 unsupported"; return CVR_ERR_UNSUPPORTED; }
    if (s.block_size > 0 && s.block_size % 2 == 0) { err = "block_size must be odd"; return CVR_ERR_INVALID_ARG; }
    if (s.pre_filter_cap > 63) { err = "pre_filter_cap > 63 unsupported"; return CVR_ERR_UNSUPPORTED; }
    if (s.mode != CVR_MODE_SGBM && s.mode != CVR_MODE_HH) { err = "mode must be CVR_MODE_SGBM or CVR_MODE_HH"; return CVR_ERR_UNSUPPORTED; }
    const int bs = s.block_size > 0 ? s.block_size : 5;
    const int ftzero = (s.pre_filter_cap > 15 ? s.pre_filter_cap : 15) | 1;
    // the horizontal block sum is evaluated in closed form, which equals OpenCV's saturating recurrence only
    // while a row sum cannot reach 32767
    if ((long long)bs * (2 * ftzero + 63) > 32767) { err = "block_size too large: horizontal block sum could saturate int16"; return CVR_ERR_UNSUPPORTED; }
    if (s.min_disparity < -32000 / 16 || s.min_disparity + s.num_disparities > 32000 / 16) { err = "disparity range does not fit the x16 int16 output"; return CVR_ERR_INVALID_ARG; }
    return CVR_OK;
}

static void fill_dev_params(cvr_ctx* c)
{
    const cvr_params& p = c->prm;
    SgbmDev& g = c->sg;
    memset(&g, 0, sizeof(g));
    g.W = p.width; g.H = p.height; g.Wp = (p.width + 15) & ~15;
    const cvr_sgbm_params& s = p.sgbm;
    g.minD = s.min_disparity; g.D = s.num_disparities; g.Dp = ((g.D + 63) / 64) * 64;
    g.bs = s.block_size > 0 ? s.block_size : 5; g.SW2 = g.SH2 = g.bs / 2;
    const int P1 = s.P1 > 0 ? s.P1 : 2;
    const int P2a = s.P2 > 0 ? s.P2 : 5;
    const int P2 = P2a > P1 + 1 ? P2a : P1 + 1;
    g.P1s = (int)(short)P1; g.P2 = P2; g.P2s = (int)(short)P2;
    g.uniq = s.uniqueness_ratio >= 0 ? s.uniqueness_ratio : 10;
    g.d12 = s.disp12_max_diff > 0 ? s.disp12_max_diff : 1;
    g.ftzero = (s.pre_filter_cap > 15 ? s.pre_filter_cap : 15) | 1;
    const int maxD = g.minD + g.D;
    g.minX1 = maxD > 0 ? maxD : 0;
    g.maxX1 = g.W + (g.minD < 0 ? g.minD : 0);
    g.W1 = g.maxX1 - g.minX1;
    g.INVALID = (g.minD - 1) * CVR_DISP_SCALE;
    g.mode = s.mode;
    g.ring = 1; while (g.ring < g.bs + 1) g.ring <<= 1;
    BmDev& m = c->bm;
    memset(&m, 0, sizeof(m));
    m.W = p.width; m.H = p.height; m.Wp = g.Wp;
    m.minD = p.bm_min_disparity; m.ndisp = p.bm_num_disparities; m.Dp = ((m.ndisp + 63) / 64) * 64;
    m.wsz = p.bm_window_size; m.wsz2 = m.wsz / 2; m.cap = p.bm_pre_filter_cap; m.textureThreshold = p.bm_texture_threshold; m.uniq = p.bm_uniqueness_ratio;
    { const int t = m.ndisp - 1 + m.minD; m.lofs = t > 0 ? t : 0; m.rofs = t < 0 ? -t : 0; }
    m.width1 = m.W - m.rofs - m.ndisp + 1;
    m.FILTERED = (m.minD - 1) * CVR_DISP_SCALE;
    m.ring = 1; while (m.ring < m.wsz + 1) m.ring <<= 1;
    CloudDev& cl = c->cl;
    memcpy(cl.Q, p.Q, sizeof(cl.Q));
    cl.f = p.tri_f; cl.b = p.tri_b; cl.cx = p.tri_cx; cl.cy = p.tri_cy;
    cl.sigma_factor = p.prune_sigma_factor;
    cl.leaf[0] = p.voxel_leaf[0]; cl.leaf[1] = p.voxel_leaf[1]; cl.leaf[2] = p.voxel_leaf[2];
    cl.sor_k = p.sor_mean_k; cl.sor_alpha = p.sor_stddev_mul;
}

extern "C" const char* cvr_last_error(const cvr_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }
extern "C" long long cvr_launch_count(const cvr_ctx* ctx) { return ctx ? ctx->launches : 0; }
extern "C" int cvr_suggest_batch(cvr_ctx* ctx, int around)
{
    if (!ctx) return CVR_ERR_INVALID_ARG;
    if (ctx->prm.filter_only) return around < 1 ? 1 : around;
    if (cudaSetDevice(ctx->device) != cudaSuccess) return CVR_ERR_CUDA;
    return cvr_sgbm_suggest_batch(ctx, around);
}
extern "C" int cvr_set_profiling(cvr_ctx* ctx, int enable) { if (!ctx) return CVR_ERR_INVALID_ARG; ctx->profiling = enable != 0; return CVR_OK; }
extern "C" int cvr_last_timings(const cvr_ctx* ctx, const char** names, float* ms, int max_stages)
{
    if (!ctx) return 0;
    int n = (int)ctx->t_ms.size();
    if (n > max_stages) n = max_stages;
    for (int i = 0; i < n; i++) { names[i] = ctx->t_names[i].c_str(); ms[i] = ctx->t_ms[i]; }
    return n;
}

static void free_all(cvr_ctx* c)
{
    void* ptrs[] = { c->d_in[0], c->d_in[1], c->d_gray, c->d_volA, c->d_volC, c->d_lr, c->d_minlr, c->d_raw, c->d_aux, c->d_disp, c->d_med,
                     c->d_label, c->d_size, c->d_pruned, c->d_dispf, c->d_mask, c->d_pstats, c->d_sel, c->d_boxv, c->d_boxs, c->d_cloud, c->d_ncloud, c->d_vhdr,
                     c->d_keys[0], c->d_keys[1], c->d_vals[0], c->d_vals[1], c->d_hist, c->d_segstart, c->d_vox, c->d_cellstart, c->d_scanaux, c->d_cellidx,
                     c->d_meand, c->d_sstats, c->d_sorout, c->d_nout, c->d_nvox, c->d_spts, c->d_flags, c->d_sgflags, c->d_bmpre,
                     c->d_rmap1[0], c->d_rmap1[1], c->d_rmap2[0], c->d_rmap2[1], c->d_rect[0], c->d_rect[1] };
    for (void* p : ptrs) if (p) cudaFree(p);
    if (c->h_counts) cudaFreeHost(c->h_counts);
    if (c->h_meta0) cudaFreeHost(c->h_meta0);
    for (uint8_t* q : { c->png_blob, c->png_z, c->png_raw, c->png_tab, c->png_out }) if (q) cudaFree(q);
    for (auto& st_ : c->png_stream) if (st_) cudaStreamDestroy(st_);
    for (auto& ev_ : c->png_ev) if (ev_) cudaEventDestroy(ev_);
    if (c->d_icp) cudaFree(c->d_icp);
    if (c->h_icp) cudaFreeHost(c->h_icp);
    if (c->d_meta0) cudaFree(c->d_meta0);
    cvr_comm_release(c);
    if (c->pipe_ready) {
        for (int s = 0; s < 2; s++) {
            PipeSlot& sl = c->slot[s];
            cudaFree(sl.d_in[0]); cudaFree(sl.d_in[1]); cudaFree(sl.d_out); cudaFree(sl.d_meta); cudaFree(sl.d_outdisp); if (sl.h_meta) cudaFreeHost(sl.h_meta);
            cudaEventDestroy(sl.evH2D); cudaEventDestroy(sl.evDone);
        }
        cudaStreamDestroy(c->copy_stream); cudaStreamDestroy(c->d2h_stream);
        if (c->tail_stream) cudaStreamDestroy(c->tail_stream);
        if (c->evPath) cudaEventDestroy(c->evPath);
        if (c->evAuxFree) cudaEventDestroy(c->evAuxFree);
        if (c->evTailLast) cudaEventDestroy(c->evTailLast);
    }
    for (auto& tv : c->timers) for (auto& t : tv) cudaEventDestroy(t.ev);
    if (c->stream) cudaStreamDestroy(c->stream);
}

extern "C" int cvr_create(const cvr_params* p, int device, cvr_ctx** out)
{
    if (!out) return CVR_ERR_INVALID_ARG;
    *out = nullptr;
    int rc = validate(p, g_create_err);
    if (rc) return rc;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { g_create_err = "no CUDA device: libcvrecon_b200 has no CPU fallback"; cudaGetLastError(); return CVR_ERR_NO_DEVICE; }
    if (device < 0 || device >= ndev) { g_create_err = "device index out of range"; return CVR_ERR_INVALID_ARG; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { g_create_err = "cudaGetDeviceProperties failed"; return CVR_ERR_CUDA; }
    if (prop.major != 10) { g_create_err = "device is not sm_100-class (this library is built for sm_100a only)"; return CVR_ERR_NO_DEVICE; }
    if (cudaSetDevice(device) != cudaSuccess) { g_create_err = "cudaSetDevice failed"; return CVR_ERR_CUDA; }
    cvr_ctx* c = new (std::nothrow) cvr_ctx();
    if (!c) return CVR_ERR_CUDA;
    c->prm = *p; c->device = device; c->maxB = p->max_batch; c->sm_count = prop.multiProcessorCount;
    c->launches = 0; c->profiling = false; c->timers_used[0] = c->timers_used[1] = c->timers_used[2] = 0; c->tset = 0; c->stream = nullptr; c->h_counts = nullptr; c->h_meta0 = nullptr; c->d_meta0 = nullptr; c->lp_valid = false; c->d_icp = nullptr; c->h_icp = nullptr; c->png_blob = c->png_z = c->png_raw = c->png_tab = c->png_out = nullptr; c->png_blob_cap = c->png_z_cap = c->png_raw_cap = c->png_tab_cap = c->png_out_cap = 0; for (auto& st_ : c->png_stream) st_ = nullptr; for (auto& ev_ : c->png_ev) ev_ = nullptr; cvr_comm_reset(c);
    c->pipe_ready = false; c->copy_stream = nullptr; c->d2h_stream = nullptr; c->n_submit = c->n_wait = 0;
    c->tail_stream = nullptr; c->path_stream = nullptr; c->evPath = c->evAuxFree = c->evTailLast = nullptr; c->split_tail = c->aux_pending = c->tail_dirty = false;
    { const char* e_ = getenv("CVR_TAIL_STREAM"); c->tail_mode = e_ ? atoi(e_) : 0; }
    memset(c->slot, 0, sizeof(c->slot));
    fill_dev_params(c);
    c->smem_optin = (int)prop.sharedMemPerBlockOptin; c->smem_sm = (int)prop.sharedMemPerMultiprocessor;
    { const char* e_ = getenv("CVR_PATH_KERNELS"); c->force_generic = e_ && strcmp(e_, "generic") == 0; c->path_v_variant = (e_ && strcmp(e_, "v2") == 0) ? 2 : ((e_ && strcmp(e_, "v3") == 0) ? 3 : 4); memset(c->pathv4_active, 0, sizeof(c->pathv4_active));
      const char* pf_ = getenv("CVR_PATH_H_FWD"); c->path_h_fwd = (pf_ && atoi(pf_) == 3) ? 3 : 2;   // -> : one row per warp already runs at the HBM rate
      const char* pr_ = getenv("CVR_PATH_H_REV"); c->path_h_rev = (pr_ && atoi(pr_) == 2) ? 2 : 3;   // <- + WTA: several rows per warp
      const char* ks_ = getenv("CVR_KEEP_S"); c->keep_S = ks_ && atoi(ks_) != 0;
      const char* nn_ = getenv("CVR_NO_NN"); c->force_no_nn = nn_ && atoi(nn_) != 0;
      const char* n_ = getenv("CVR_PATHV_NT"); c->pathv_max_nt = n_ ? atoi(n_) : 1024; memset(c->pathv_active, 0, sizeof(c->pathv_active)); }
    const SgbmDev& g = c->sg;
    const size_t B = (size_t)c->maxB, HW = (size_t)g.W * g.H;
    c->maxN = (int)(p->max_points > 0 ? (size_t)p->max_points : HW);
    if (!p->filter_only && (size_t)c->maxN < HW) c->maxN = (int)HW;
    const size_t N = (size_t)c->maxN;
    // search-grid cells per frame: a filter-only context (one large cloud) gets room for ~2 points per cell even when outliers
    // blow up the bounding box; the batch path's clouds are far smaller than W*H anyway
    { const long long cap = p->filter_only ? 4LL * c->maxN : (long long)c->maxN; c->sor_grid_cap = (int)(cap < 4096 ? 4096 : (cap > 0x3fffffffLL ? 0x3fffffffLL : cap)); }
    const size_t nRuns = (N + 511) / 512;
    cudaError_t e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
#define A(call) if (e == cudaSuccess) e = (call)
    if (!p->filter_only) {
        const bool isBM = p->block_matcher == CVR_MATCHER_BM;
        const size_t W1 = isBM ? (c->bm.width1 > 0 ? c->bm.width1 : 0) : (g.W1 > 0 ? g.W1 : 0);
        const size_t vol = (size_t)g.H * W1 * (isBM ? c->bm.Dp : g.Dp) * (isBM ? 1 : 1);
        A(dalloc(&c->d_in[0], B * HW * 4)); A(dalloc(&c->d_in[1], B * HW * 4));
        A(dalloc(&c->d_gray, B * 2 * g.H * g.Wp));
        A(dalloc(&c->d_bmpre, isBM ? B * 2 * g.H * g.Wp : 1));
        A(dalloc(&c->d_volA, B * vol)); A(dalloc(&c->d_volC, isBM ? 1 : B * vol));
        A(dalloc(&c->d_lr, B * 2 * (W1 + 2) * 3 * g.Dp)); A(dalloc(&c->d_minlr, B * 2 * (W1 + 2) * 4));
        A(dalloc(&c->d_raw, B * HW)); A(dalloc(&c->d_aux, B * g.H * (W1 + 1))); A(dalloc(&c->d_disp, B * HW)); A(dalloc(&c->d_med, B * HW));
        A(dalloc(&c->d_label, B * HW)); A(dalloc(&c->d_size, B * HW));
        A(dalloc(&c->d_pruned, B * HW)); A(dalloc(&c->d_dispf, B * HW)); A(dalloc(&c->d_mask, B * HW));
        A(dalloc(&c->d_pstats, B));
    }
    A(dalloc(&c->d_sel, 3 * B * cvr_sel_stride(c->maxN)));
    A(dalloc(&c->d_boxv, B)); A(dalloc(&c->d_boxs, B));
    A(dalloc(&c->d_cloud, B * N)); A(dalloc(&c->d_ncloud, B));
    A(dalloc(&c->d_vhdr, B));
    A(dalloc(&c->d_keys[0], B * N)); A(dalloc(&c->d_keys[1], B * N)); A(dalloc(&c->d_vals[0], B * N)); A(dalloc(&c->d_vals[1], B * N));
    A(dalloc(&c->d_hist, B * (256 * nRuns + 1)));
    A(dalloc(&c->d_segstart, B * (N + 1))); A(dalloc(&c->d_vox, B * N));
    A(dalloc(&c->d_cellstart, B * 2 * ((size_t)c->sor_grid_cap + 2))); A(dalloc(&c->d_scanaux, B * ((size_t)c->sor_grid_cap / 1024 + 4))); A(dalloc(&c->d_cellidx, B * N));
    A(dalloc(&c->d_meand, B * N)); A(dalloc(&c->d_sstats, B)); A(dalloc(&c->d_sorout, B * N));
    A(dalloc(&c->d_nout, B)); A(dalloc(&c->d_nvox, B)); A(dalloc(&c->d_spts, B * N));
    A(dalloc(&c->d_flags, B * 64));
    if (e == cudaSuccess) e = cudaMemset(c->d_flags, 0, B * 64 * sizeof(int));     // (holds counters that every use leaves at zero)
    A(dalloc(&c->d_sgflags, B));
    A(cudaHostAlloc((void**)&c->h_counts, sizeof(int) * B * 4, cudaHostAllocDefault));
    A(dalloc(&c->d_meta0, CVR_META_INTS(B))); A(cudaHostAlloc((void**)&c->h_meta0, CVR_META_INTS(B) * sizeof(int), cudaHostAllocDefault));
#undef A
    if (e != cudaSuccess) {
        g_create_err = std::string("workspace allocation failed: ") + cudaGetErrorString(e);
        free_all(c); delete c; cudaGetLastError();
        return CVR_ERR_CUDA;
    }
    *out = c;
    return CVR_OK;
}

extern "C" int cvr_destroy(cvr_ctx* c)
{
    if (!c) return CVR_OK;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->pipe_ready) { cudaStreamSynchronize(c->copy_stream); cudaStreamSynchronize(c->d2h_stream); if (c->tail_stream) cudaStreamSynchronize(c->tail_stream); }
    free_all(c);
    delete c;
    return CVR_OK;
}

#define REQUIRE(ctx, cond, msg) do { if (!(cond)) { (ctx)->err = (msg); return CVR_ERR_INVALID_ARG; } } while (0)
#define STEREO_ONLY(ctx) do { if ((ctx)->prm.filter_only) { (ctx)->err = "context was created with filter_only"; return CVR_ERR_INVALID_ARG; } } while (0)

static int sync_check(cvr_ctx* c)
{
    cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) { c->err = std::string("stream sync: ") + cudaGetErrorString(e); return CVR_ERR_CUDA; }
    return CVR_OK;
}

extern "C" int cvr_disparity(cvr_ctx* c, const uint8_t* left, size_t sl, const uint8_t* right, size_t sr, int ch, int16_t* disp, size_t sd)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    STEREO_ONLY(c);
    REQUIRE(c, left && right && disp, "null image pointer");
    // cv::cvtColor(BGR2GRAY) throws for anything but 3/4 channels; 1 = caller already has gray
    REQUIRE(c, ch == 1 || ch == 3 || ch == 4, "channels must be 1, 3 or 4");
    const int W = c->sg.W, H = c->sg.H;
    REQUIRE(c, sl >= (size_t)W * ch && sr >= (size_t)W * ch && sd >= (size_t)W * 2, "step smaller than a row");
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    cvr_timers_begin(c, 0);
    CVR_CUDA(c, cudaMemcpy2DAsync(c->d_in[0], (size_t)W * ch, left, sl, (size_t)W * ch, H, cudaMemcpyHostToDevice, c->stream));
    CVR_CUDA(c, cudaMemcpy2DAsync(c->d_in[1], (size_t)W * ch, right, sr, (size_t)W * ch, H, cudaMemcpyHostToDevice, c->stream));
    int rc = cvr_launch_gray(c, 1, c->d_in[0], (size_t)W * ch, 0, c->d_in[1], (size_t)W * ch, 0, ch);
    if (rc) return rc;
    cvr_stage_mark(c, "h2d_gray");
    rc = (c->prm.block_matcher == CVR_MATCHER_BM) ? cvr_launch_bm(c, 1) : cvr_launch_sgbm(c, 1);
    if (rc) return rc;
    CVR_CUDA(c, cudaMemcpy2DAsync(disp, sd, c->d_med, (size_t)W * 2, (size_t)W * 2, H, cudaMemcpyDeviceToHost, c->stream));
    cvr_stage_mark(c, "d2h");
    rc = sync_check(c);
    cvr_timers_end(c, 0);
    return rc;
}

static int rect_alloc(cvr_ctx* c)
{
    const size_t HW = (size_t)c->prm.width * c->prm.height;
    for (int cam = 0; cam < 2; cam++) {
        if (!c->d_rmap1[cam]) CVR_CUDA(c, dalloc(&c->d_rmap1[cam], HW));
        if (!c->d_rmap2[cam]) CVR_CUDA(c, dalloc(&c->d_rmap2[cam], HW));
        if (!c->d_rect[cam]) CVR_CUDA(c, dalloc(&c->d_rect[cam], (size_t)c->maxB * HW * 4));
    }
    return CVR_OK;
}

extern "C" int cvr_set_rectification(cvr_ctx* c, const double* K1, const double* D1, int nD1, const double* R1, const double* P1, int p1c,
                                     const double* K2, const double* D2, int nD2, const double* R2, const double* P2, int p2c)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    STEREO_ONLY(c);
    REQUIRE(c, K1 && R1 && P1 && K2 && R2 && P2, "null calibration matrix");
    REQUIRE(c, (p1c == 3 || p1c == 4) && (p2c == 3 || p2c == 4), "P must have 3 or 4 columns");
    REQUIRE(c, nD1 >= 0 && nD2 >= 0 && (nD1 == 0 || D1) && (nD2 == 0 || D2), "bad distortion vector");
    if (nD1 > 12 || nD2 > 12) { c->err = "distortion vectors with tilt terms (more than 12 coefficients) are not supported"; return CVR_ERR_UNSUPPORTED; }
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    int rc = rect_alloc(c);
    if (rc) return rc;
    c->rect_ready = false;
    rc = cvr_launch_rectify_maps(c, 0, K1, D1, nD1, R1, P1, p1c);
    if (rc) return rc;
    rc = cvr_launch_rectify_maps(c, 1, K2, D2, nD2, R2, P2, p2c);
    if (rc) return rc;
    rc = sync_check(c);
    if (rc) return rc;
    c->rect_ready = true;
    return CVR_OK;
}

extern "C" int cvr_get_rectification_maps(cvr_ctx* c, int cam, int16_t* map1, uint16_t* map2)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    REQUIRE(c, c->rect_ready, "cvr_set_rectification was not called");
    REQUIRE(c, cam == 0 || cam == 1, "camera must be 0 or 1");
    const size_t HW = (size_t)c->prm.width * c->prm.height;
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    if (map1) CVR_CUDA(c, cudaMemcpy(map1, c->d_rmap1[cam], HW * 4, cudaMemcpyDeviceToHost));
    if (map2) CVR_CUDA(c, cudaMemcpy(map2, c->d_rmap2[cam], HW * 2, cudaMemcpyDeviceToHost));
    return CVR_OK;
}

extern "C" int cvr_rectify(cvr_ctx* c, int ch, const uint8_t* left, size_t sl, const uint8_t* right, size_t sr, uint8_t* outL, size_t sol,
                           uint8_t* outR, size_t sor_)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    STEREO_ONLY(c);
    REQUIRE(c, c->rect_ready, "cvr_set_rectification was not called");
    REQUIRE(c, left && right && outL && outR, "null image pointer");
    REQUIRE(c, ch == 1 || ch == 3 || ch == 4, "channels must be 1, 3 or 4");
    const int W = c->sg.W, H = c->sg.H;
    const size_t rowBytes = (size_t)W * ch;
    REQUIRE(c, sl >= rowBytes && sr >= rowBytes && sol >= rowBytes && sor_ >= rowBytes, "step smaller than a row");
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    CVR_CUDA(c, cudaMemcpy2DAsync(c->d_in[0], rowBytes, left, sl, rowBytes, H, cudaMemcpyHostToDevice, c->stream));
    CVR_CUDA(c, cudaMemcpy2DAsync(c->d_in[1], rowBytes, right, sr, rowBytes, H, cudaMemcpyHostToDevice, c->stream));
    int rc = cvr_launch_remap(c, 1, ch, c->d_in[0], rowBytes, rowBytes * H, c->d_in[1], rowBytes, rowBytes * H, c->d_rect[0], c->d_rect[1], rowBytes, rowBytes * H);
    if (rc) return rc;
    CVR_CUDA(c, cudaMemcpy2DAsync(outL, sol, c->d_rect[0], rowBytes, rowBytes, H, cudaMemcpyDeviceToHost, c->stream));
    CVR_CUDA(c, cudaMemcpy2DAsync(outR, sor_, c->d_rect[1], rowBytes, rowBytes, H, cudaMemcpyDeviceToHost, c->stream));
    return sync_check(c);
}

extern "C" int cvr_debug_fetch(cvr_ctx* c, int which, void* dst, size_t bytes, int* w1, int* dp)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    if (which == 4 || which == 5) {                                    // cloud filters: usable on filter_only contexts too
        // 4: mean k-NN distance of every input point of the last cvr_sor call (frame 0, input order, `bytes`/4 floats);
        // 5: its SorStats record {double sum, sq, thr; int valid, kept}
        const void* src = which == 4 ? (const void*)c->d_meand : (const void*)c->d_sstats;
        const size_t need = which == 4 ? bytes : sizeof(SorStats);
        if (!dst) return CVR_OK;
        REQUIRE(c, bytes >= need && (which == 5 || bytes <= (size_t)c->maxN * 4), "debug buffer size");
        CVR_CUDA(c, cudaSetDevice(c->device));
        { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
        CVR_CUDA(c, cudaMemcpy(dst, src, need, cudaMemcpyDeviceToHost));
        return CVR_OK;
    }
    STEREO_ONLY(c);
    if (c->prm.block_matcher != CVR_MATCHER_SGBM) { c->err = "debug buffers exist for the SGBM matcher only"; return CVR_ERR_INVALID_ARG; }
    const SgbmDev& g = c->sg;
    if (w1) *w1 = g.W1;
    if (dp) *dp = g.Dp;
    const size_t vol = (size_t)g.H * (g.W1 > 0 ? g.W1 : 0) * g.Dp * 2;
    const void* src; size_t need;
    switch (which) {
        case 0: src = c->d_volC; need = vol; break;
        case 1:
            if (!c->keep_S) { c->err = "the final S volume is only kept when the context was created with CVR_KEEP_S=1 in the environment"; return CVR_ERR_INVALID_ARG; }
            src = c->d_volA; need = vol; break;
        case 2:
            if (!c->keep_S) { c->err = "the raw disparity is only kept when the context was created with CVR_KEEP_S=1 in the environment"; return CVR_ERR_INVALID_ARG; }
            src = c->d_raw; need = (size_t)g.W * g.H * 2; break;
        case 3: src = c->d_disp; need = (size_t)g.W * g.H * 2; break;
        default: c->err = "unknown debug buffer"; return CVR_ERR_INVALID_ARG;
    }
    if (!dst) return CVR_OK;
    REQUIRE(c, bytes >= need, "debug buffer too small");
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    CVR_CUDA(c, cudaMemcpy(dst, src, need, cudaMemcpyDeviceToHost));
    return CVR_OK;
}

extern "C" int cvr_prune(cvr_ctx* c, int16_t* disp, size_t sd, float* dispf, uint8_t* mask)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    STEREO_ONLY(c);
    REQUIRE(c, disp, "null disparity");
    const int W = c->sg.W, H = c->sg.H;
    REQUIRE(c, sd >= (size_t)W * 2, "step smaller than a row");
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    cvr_timers_begin(c, 0);
    CVR_CUDA(c, cudaMemcpy2DAsync(c->d_med, (size_t)W * 2, disp, sd, (size_t)W * 2, H, cudaMemcpyHostToDevice, c->stream));
    int rc = cvr_launch_prune(c, 1, c->d_med);
    if (rc) return rc;
    CVR_CUDA(c, cudaMemcpy2DAsync(disp, sd, c->d_pruned, (size_t)W * 2, (size_t)W * 2, H, cudaMemcpyDeviceToHost, c->stream));
    if (dispf) CVR_CUDA(c, cudaMemcpyAsync(dispf, c->d_dispf, (size_t)W * H * 4, cudaMemcpyDeviceToHost, c->stream));
    if (mask) CVR_CUDA(c, cudaMemcpyAsync(mask, c->d_mask, (size_t)W * H, cudaMemcpyDeviceToHost, c->stream));
    rc = sync_check(c);
    cvr_timers_end(c, 0);
    return rc;
}

static int fetch_cloud(cvr_ctx* c, const cvr_point* dsrc, const int* dcount, cvr_point* out, int capacity, int* n_out)
{
    CVR_CUDA(c, cudaMemcpyAsync(c->h_counts, dcount, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    int rc = sync_check(c);
    if (rc) return rc;
    const int n = c->h_counts[0];
    if (n_out) *n_out = n;
    if (n > capacity) { c->err = "output capacity too small for the cloud"; return CVR_ERR_CAPACITY; }
    if (n > 0) CVR_CUDA(c, cudaMemcpyAsync(out, dsrc, (size_t)n * sizeof(cvr_point), cudaMemcpyDeviceToHost, c->stream));
    return sync_check(c);
}

extern "C" int cvr_reproject_cloud(cvr_ctx* c, const int16_t* disp, size_t sd, const uint8_t* bgr, size_t sb, cvr_point* out, int capacity, int* n_out)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    STEREO_ONLY(c);
    REQUIRE(c, disp && bgr && out, "null pointer");
    const int W = c->sg.W, H = c->sg.H;
    REQUIRE(c, sd >= (size_t)W * 2 && sb >= (size_t)W * 3, "step smaller than a row");
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    cvr_timers_begin(c, 0);
    CVR_CUDA(c, cudaMemcpy2DAsync(c->d_pruned, (size_t)W * 2, disp, sd, (size_t)W * 2, H, cudaMemcpyHostToDevice, c->stream));
    CVR_CUDA(c, cudaMemcpy2DAsync(c->d_in[0], (size_t)W * 3, bgr, sb, (size_t)W * 3, H, cudaMemcpyHostToDevice, c->stream));
    int rc = cvr_launch_cloud(c, 1, CVR_CLOUD_REPROJECT, c->d_pruned, nullptr, nullptr, c->d_in[0], (size_t)W * 3, 0, 3, false);
    if (rc) return rc;
    rc = fetch_cloud(c, c->d_cloud, c->d_ncloud, out, capacity, n_out);
    cvr_timers_end(c, 0);
    return rc;
}

extern "C" int cvr_triangulate(cvr_ctx* c, const float* dispf, size_t sd, const uint8_t* bgr, size_t sb, const uint8_t* mask, size_t sm,
                               cvr_point* out, int capacity, int* n_out)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    STEREO_ONLY(c);
    REQUIRE(c, dispf && bgr && out, "null pointer");
    const int W = c->sg.W, H = c->sg.H;
    REQUIRE(c, sd >= (size_t)W * 4 && sb >= (size_t)W * 3 && (!mask || sm >= (size_t)W), "step smaller than a row");
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    cvr_timers_begin(c, 0);
    CVR_CUDA(c, cudaMemcpy2DAsync(c->d_dispf, (size_t)W * 4, dispf, sd, (size_t)W * 4, H, cudaMemcpyHostToDevice, c->stream));
    CVR_CUDA(c, cudaMemcpy2DAsync(c->d_in[0], (size_t)W * 3, bgr, sb, (size_t)W * 3, H, cudaMemcpyHostToDevice, c->stream));
    if (mask) CVR_CUDA(c, cudaMemcpy2DAsync(c->d_mask, (size_t)W, mask, sm, (size_t)W, H, cudaMemcpyHostToDevice, c->stream));
    int rc = cvr_launch_cloud(c, 1, CVR_CLOUD_TRIANGULATE, nullptr, c->d_dispf, mask ? c->d_mask : nullptr, c->d_in[0], (size_t)W * 3, 0, 3, false);
    if (rc) return rc;
    rc = fetch_cloud(c, c->d_cloud, c->d_ncloud, out, capacity, n_out);
    cvr_timers_end(c, 0);
    return rc;
}

extern "C" int cvr_voxel_grid(cvr_ctx* c, const cvr_point* in, int n, cvr_point* out, int capacity, int* m_out, int* passthrough)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    REQUIRE(c, n >= 0 && (n == 0 || (in && out)), "null pointer");
    if (n > c->maxN) { c->err = "cloud larger than the context's point capacity (cvr_params.max_points)"; return CVR_ERR_CAPACITY; }
    if (m_out) *m_out = 0;
    if (passthrough) *passthrough = 0;
    if (n == 0) return CVR_OK;
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    cvr_timers_begin(c, 0);
    CVR_CUDA(c, cudaMemcpyAsync(c->d_cloud, in, (size_t)n * sizeof(cvr_point), cudaMemcpyHostToDevice, c->stream));
    c->h_counts[1] = n;
    CVR_CUDA(c, cudaMemcpyAsync(c->d_ncloud, &c->h_counts[1], sizeof(int), cudaMemcpyHostToDevice, c->stream));
    int rc = cvr_launch_voxel(c, 1, c->d_cloud, (size_t)c->maxN, c->d_ncloud, false);
    if (rc) return rc;
    rc = fetch_cloud(c, c->d_vox, c->d_nvox, out, capacity, m_out);
    if (rc) return rc;
    if (passthrough) {
        VoxelHdr h;
        CVR_CUDA(c, cudaMemcpy(&h, c->d_vhdr, sizeof(h), cudaMemcpyDeviceToHost));
        *passthrough = h.passthrough;
    }
    cvr_timers_end(c, 0);
    return CVR_OK;
}

extern "C" int cvr_sor(cvr_ctx* c, const cvr_point* in, int n, cvr_point* out, int capacity, int* kept_out)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    REQUIRE(c, n >= 0 && (n == 0 || (in && out)), "null pointer");
    if (n > c->maxN) { c->err = "cloud larger than the context's point capacity (cvr_params.max_points)"; return CVR_ERR_CAPACITY; }
    if (kept_out) *kept_out = 0;
    if (n == 0) return CVR_OK;
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    cvr_timers_begin(c, 0);
    CVR_CUDA(c, cudaMemcpyAsync(c->d_vox, in, (size_t)n * sizeof(cvr_point), cudaMemcpyHostToDevice, c->stream));
    c->h_counts[1] = n;
    CVR_CUDA(c, cudaMemcpyAsync(c->d_nvox, &c->h_counts[1], sizeof(int), cudaMemcpyHostToDevice, c->stream));
    int rc = cvr_launch_sor(c, 1, c->d_vox, (size_t)c->maxN, c->d_nvox, false);
    if (rc) return rc;
    rc = fetch_cloud(c, c->d_sorout, c->d_nout, out, capacity, kept_out);
    cvr_timers_end(c, 0);
    return rc;
}

// VoxelGrid and/or StatisticalOutlierRemoval on one cloud in host OR device memory (the per-stage entry points above take host
// pointers only): the MappingSystem call sites run on clouds that are already on the device when the path produced them.
extern "C" int cvr_filter_cloud(cvr_ctx* c, int mem, const cvr_point* in, int n, int stages, cvr_point* out, int capacity, int* n_out)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    REQUIRE(c, n >= 0 && (n == 0 || (in && out)) && (stages & 12) != 0 && (mem == CVR_MEM_HOST || mem == CVR_MEM_DEVICE), "bad argument");
    if (n > c->maxN) { c->err = "cloud larger than the context's point capacity (cvr_params.max_points)"; return CVR_ERR_CAPACITY; }
    if (n_out) *n_out = 0;
    if (n == 0) return CVR_OK;
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    cvr_timers_begin(c, 0);
    const cudaMemcpyKind kin = mem == CVR_MEM_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
    const cudaMemcpyKind kout = mem == CVR_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
    const bool vox = (stages & 4) != 0, sor = (stages & 8) != 0;
    cvr_point* first = vox ? c->d_cloud : c->d_vox;                     // the slab the first enabled stage reads
    int* firstN = vox ? c->d_ncloud : c->d_nvox;
    CVR_CUDA(c, cudaMemcpyAsync(first, in, (size_t)n * sizeof(cvr_point), kin, c->stream));
    c->h_counts[1] = n;
    CVR_CUDA(c, cudaMemcpyAsync(firstN, &c->h_counts[1], sizeof(int), cudaMemcpyHostToDevice, c->stream));
    cvr_stage_mark(c, "in");
    int rc;
    const cvr_point* fin = first; const int* fcnt = firstN;
    if (vox) { rc = cvr_launch_voxel(c, 1, c->d_cloud, (size_t)c->maxN, c->d_ncloud, false); if (rc) return rc; fin = c->d_vox; fcnt = c->d_nvox; }
    if (sor) { rc = cvr_launch_sor(c, 1, fin, (size_t)c->maxN, fcnt, vox); if (rc) return rc; fin = c->d_sorout; fcnt = c->d_nout; }
    CVR_CUDA(c, cudaMemcpyAsync(c->h_counts, fcnt, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    rc = sync_check(c);
    if (rc) return rc;
    const int m = c->h_counts[0];
    if (n_out) *n_out = m;
    if (m > capacity) { c->err = "output capacity too small for the cloud"; return CVR_ERR_CAPACITY; }
    if (m > 0) CVR_CUDA(c, cudaMemcpyAsync(out, fin, (size_t)m * sizeof(cvr_point), kout, c->stream));
    cvr_stage_mark(c, "out");
    rc = sync_check(c);
    cvr_timers_end(c, 0);
    return rc;
}

extern "C" int cvr_radius_outlier_removal(cvr_ctx* c, const cvr_point* in, int n, float radius, int min_neighbors, cvr_point* out, int capacity, int* kept_out)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    REQUIRE(c, n >= 0 && (n == 0 || (in && out)) && radius > 0.f && min_neighbors >= 0, "bad argument");
    if (n > c->maxN) { c->err = "cloud larger than the context's point capacity (cvr_params.max_points)"; return CVR_ERR_CAPACITY; }
    if (kept_out) *kept_out = 0;
    if (n == 0) return CVR_OK;
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    cvr_timers_begin(c, 0);
    CVR_CUDA(c, cudaMemcpyAsync(c->d_vox, in, (size_t)n * sizeof(cvr_point), cudaMemcpyHostToDevice, c->stream));
    c->h_counts[1] = n;
    CVR_CUDA(c, cudaMemcpyAsync(c->d_nvox, &c->h_counts[1], sizeof(int), cudaMemcpyHostToDevice, c->stream));
    int rc = cvr_launch_ror(c, 1, c->d_vox, (size_t)c->maxN, c->d_nvox, radius, min_neighbors);
    if (rc) return rc;
    rc = fetch_cloud(c, c->d_sorout, c->d_nout, out, capacity, kept_out);
    cvr_timers_end(c, 0);
    return rc;
}

extern "C" int cvr_transform_cloud(cvr_ctx* c, const cvr_point* in, int n, const float* T, cvr_point* out)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    REQUIRE(c, n >= 0 && T && (n == 0 || (in && out)), "bad argument");
    if (n > c->maxN) { c->err = "cloud larger than the context's point capacity (cvr_params.max_points)"; return CVR_ERR_CAPACITY; }
    if (n == 0) return CVR_OK;
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    CVR_CUDA(c, cudaMemcpyAsync(c->d_vox, in, (size_t)n * sizeof(cvr_point), cudaMemcpyHostToDevice, c->stream));
    int rc = cvr_launch_transform(c, c->d_vox, c->d_sorout, n, T);
    if (rc) return rc;
    CVR_CUDA(c, cudaMemcpyAsync(out, c->d_sorout, (size_t)n * sizeof(cvr_point), cudaMemcpyDeviceToHost, c->stream));
    return sync_check(c);
}

// ---- output side of the batched calls ----------------------------------------------------------------------------
// The per-frame clouds of the last filter stage sit in [B][maxN] slabs.  k_pack_clouds writes them back to back into one
// payload (frame i = points [off[i], off[i+1])) and builds the record the cloud gather exchanges between ranks:
//   meta[0 .. maxB]                 off[0..B]   exclusive prefix of the per-frame counts (off[B] = total)
//   meta[maxB+1], meta[maxB+2]      nframes, total points            \  "header": 2 + maxB ints, the same size on every rank
//   meta[maxB+3 .. 2*maxB+2]        per-frame counts (0 beyond B)    /
//   meta[2*maxB+3]                  overflow flag: a cloud did not fit the caller's device buffer
// One payload serves the single device->host copy of the host path, the caller's packed output and the NCCL send.
__global__ void __launch_bounds__(256) k_pack_clouds(const cvr_point* __restrict__ src, size_t srcStride, const int* __restrict__ cnt, int B, int maxB,
                                                     cvr_point* __restrict__ dst, int* __restrict__ meta)
{
    __shared__ int ws[8];
    const int b = blockIdx.y, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int part = 0;
    for (int i = threadIdx.x; i < b; i += blockDim.x) part += cnt[i];
    part = __reduce_add_sync(0xffffffffu, part);
    if (lane == 0) ws[w] = part;
    __syncthreads();
    int off = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) off += ws[k];
    const int n = cnt[b];
    const uint4* s4 = (const uint4*)(src + (size_t)b * srcStride);
    uint4* d4 = (uint4*)(dst + off);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) d4[i] = s4[i];
    if (blockIdx.x == 0) {
        int* hdr = meta + maxB + 1;
        if (threadIdx.x == 0) {
            meta[b] = off; hdr[2 + b] = n;
            if (b == B - 1) { meta[B] = off + n; hdr[0] = B; hdr[1] = off + n; }
            if (b == 0) meta[2 * maxB + 3] = 0;
        }
        if (b == 0) for (int i = B + threadIdx.x; i < maxB; i += blockDim.x) hdr[2 + i] = 0;
    }
}
// device-resident callers, frame-strided layout: frame i at dst + i*cap; raises the overflow flag instead of truncating silently
__global__ void k_copy_clouds(const cvr_point* __restrict__ src, size_t srcStride, const int* __restrict__ cnt, cvr_point* __restrict__ dst,
                              size_t cap, int* __restrict__ cntOut, int* ovf)
{
    const int b = blockIdx.y, full = cnt[b], n = (int)min((size_t)full, cap);
    const uint4* s4 = (const uint4*)(src + (size_t)b * srcStride);
    uint4* d4 = (uint4*)(dst + (size_t)b * cap);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) d4[i] = s4[i];
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (cntOut) cntOut[b] = full;
        if (full > n) atomicExch(ovf, 1);
    }
}
// device-resident callers, packed layout: the payload and its offsets
__global__ void k_copy_packed(const cvr_point* __restrict__ src, const int* __restrict__ meta, int B, cvr_point* __restrict__ dst, size_t cap,
                              int* __restrict__ offOut, int* __restrict__ cntOut, int maxB, int* ovf)
{
    const int total = meta[B], n = (int)min((size_t)total, cap);
    const uint4* s4 = (const uint4*)src;
    uint4* d4 = (uint4*)dst;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) d4[i] = s4[i];
    if (blockIdx.x == 0) {
        for (int i = threadIdx.x; i <= B; i += blockDim.x) { if (offOut) offOut[i] = meta[i]; if (cntOut && i < B) cntOut[i] = meta[maxB + 3 + i]; }
        if (threadIdx.x == 0 && total > n) atomicExch(ovf, 1);
    }
}

// The per-frame chain in two halves, each enqueued on c->stream (whatever that currently is):
//   enqueue_matcher: (rectify ->) gray -> SGBM / BM, final disparity in c->d_med; *dl / *lstep / *lfs are updated to the
//                    left image the cloud must be coloured from (the rectified one when requires_rectification is set);
//   enqueue_filters: prune -> cloud -> voxel grid -> SOR from the disparity in `disp`; *fin / *fcnt receive the buffers
//                    that hold the last enabled stage's clouds (nullptr when only disparity was asked for).
static int enqueue_matcher(cvr_ctx* c, int B, const uint8_t** dl, size_t* lstep, size_t* lfs, const uint8_t* dr, size_t rstep, size_t rfs, int ch)
{
    int rc;
    if (c->prm.requires_rectification) {                               // ReconstructionSystem.cpp:42-44: everything downstream sees the rectified pair
        if (!c->rect_ready) { c->err = "requires_rectification is set but cvr_set_rectification was not called"; return CVR_ERR_INVALID_ARG; }
        const size_t orow = (size_t)c->sg.W * ch, ofs = orow * c->sg.H;
        rc = cvr_launch_remap(c, B, ch, *dl, *lstep, *lfs, dr, rstep, rfs, c->d_rect[0], c->d_rect[1], orow, ofs);
        if (rc) return rc;
        *dl = c->d_rect[0]; dr = c->d_rect[1]; *lstep = rstep = orow; *lfs = rfs = ofs;
        cvr_stage_mark(c, "rectify");
    }
    rc = cvr_launch_gray(c, B, *dl, *lstep, *lfs, dr, rstep, rfs, ch);
    if (rc) return rc;
    cvr_stage_mark(c, "gray");
    return (c->prm.block_matcher == CVR_MATCHER_BM) ? cvr_launch_bm(c, B) : cvr_launch_sgbm(c, B);
}

static int enqueue_filters(cvr_ctx* c, int B, int stages, const int16_t* disp, const uint8_t* dl, size_t lstep, size_t lfs, int ch,
                           const cvr_point** fin, const int** fcnt)
{
    const bool wantCloud = (stages & 2) != 0, wantVox = (stages & 4) != 0, wantSor = (stages & 8) != 0;
    *fin = nullptr; *fcnt = nullptr;
    if (!(wantCloud || wantVox || wantSor)) return CVR_OK;
    // per-frame records of the chain (prune statistics, the bounding boxes the cloud and voxel kernels accumulate)
    int rc = cvr_launch_chain_init(c, B, true, true, true);
    if (rc) return rc;
    if (c->prm.cloud_mode == CVR_CLOUD_REPROJECT) {
        rc = cvr_launch_prune(c, B, disp);                             // GeneratePointCloud reads the pruned CV_16S map (and its minimum)
        if (rc) return rc;
        rc = cvr_launch_cloud(c, B, CVR_CLOUD_REPROJECT, c->d_pruned, nullptr, nullptr, dl, lstep, lfs, ch, true);
    } else {
        rc = cvr_launch_prune_stats(c, B, disp);                       // the cloud kernel prunes on the fly from these
        if (rc) return rc;
        rc = cvr_launch_cloud(c, B, CVR_CLOUD_TRIANGULATE, disp, nullptr, nullptr, dl, lstep, lfs, ch, true);
    }
    if (rc) return rc;
    *fin = c->d_cloud; *fcnt = c->d_ncloud;
    bool haveBox = true;                                               // d_boxv by the cloud kernel, d_boxs by the centroid kernel
    if (wantVox) {
        rc = cvr_launch_voxel(c, B, *fin, (size_t)c->maxN, *fcnt, true);
        if (rc) return rc;
        *fin = c->d_vox; *fcnt = c->d_nvox;
    } else haveBox = false;
    if (wantSor) {
        rc = cvr_launch_sor(c, B, *fin, (size_t)c->maxN, *fcnt, haveBox);
        if (rc) return rc;
        *fin = c->d_sorout; *fcnt = c->d_nout;
    }
    return CVR_OK;
}

static int enqueue_chain(cvr_ctx* c, int B, int stages, const uint8_t* dl, size_t lstep, size_t lfs, const uint8_t* dr, size_t rstep, size_t rfs, int ch,
                         const cvr_point** fin, const int** fcnt)
{
    *fin = nullptr; *fcnt = nullptr;
    int rc = enqueue_matcher(c, B, &dl, &lstep, &lfs, dr, rstep, rfs, ch);
    if (rc) return rc;
    return enqueue_filters(c, B, stages, c->d_med, dl, lstep, lfs, ch, fin, fcnt);
}

static int check_batch_args(cvr_ctx* c, int nframes, const cvr_batch* io)
{
    REQUIRE(c, io && nframes >= 1, "bad batch");
    if (nframes > c->maxB) { c->err = "nframes exceeds max_batch"; return CVR_ERR_CAPACITY; }
    REQUIRE(c, io->left && io->right, "null image pointer");
    const int ch = io->channels;
    REQUIRE(c, ch == 1 || ch == 3 || ch == 4, "channels must be 1, 3 or 4");
    const size_t rowBytes = (size_t)c->sg.W * ch;
    REQUIRE(c, io->left_step >= rowBytes && io->right_step >= rowBytes, "step smaller than a row");
    REQUIRE(c, !io->disp || io->disp_step >= (size_t)c->sg.W * 2, "disp_step smaller than a row");
    return CVR_OK;
}

// The workspace slab that is free once the last enabled filter stage has written `fin`: the packed payload of the synchronous
// call goes there (the pipelined calls pack into their slot instead).
static cvr_point* free_slab(cvr_ctx* c, const cvr_point* fin) { return fin == c->d_cloud ? c->d_vox : c->d_cloud; }

static int enqueue_pack(cvr_ctx* c, int B, const cvr_point* fin, const int* fcnt, cvr_point* dst, int* meta)
{
    k_pack_clouds<<<dim3(32, B), 256, 0, c->stream>>>(fin, (size_t)c->maxN, fcnt, B, c->maxB, dst, meta);
    LAUNCH_CHECK(c);
    return CVR_OK;
}

// Device-resident outputs of one batch (strided or packed layout), from the packed payload `pk` / the slabs `fin`.
static int enqueue_device_outputs(cvr_ctx* c, int B, const cvr_batch* io, const cvr_point* fin, const int* fcnt, const cvr_point* pk, int* meta)
{
    int* ovf = meta + 2 * (size_t)c->maxB + 3;
    if (io->cloud_packed) {
        k_copy_packed<<<64, 256, 0, c->stream>>>(pk, meta, B, io->cloud_packed, io->cloud_packed_capacity, io->cloud_offsets, io->cloud_count, c->maxB, ovf);
        LAUNCH_CHECK(c);
    } else if (io->cloud) {
        k_copy_clouds<<<dim3(32, B), 256, 0, c->stream>>>(fin, (size_t)c->maxN, fcnt, io->cloud, io->cloud_capacity, io->cloud_count, ovf);
        LAUNCH_CHECK(c);
    } else if (io->cloud_count) {
        CVR_CUDA(c, cudaMemcpyAsync(io->cloud_count, fcnt, sizeof(int) * B, cudaMemcpyDeviceToDevice, c->stream));
    }
    return CVR_OK;
}

// Host-side outputs of one batch from its packed payload on the device (meta already on the host in hm).
static int fetch_host_outputs(cvr_ctx* c, int B, const cvr_batch* io, const cvr_point* pk, const int* hm, cudaStream_t st)
{
    const int* off = hm; const int* cnt = hm + c->maxB + 3;
    const size_t total = (size_t)off[B];
    if (io->cloud_count) memcpy(io->cloud_count, cnt, sizeof(int) * B);
    if (io->cloud_packed) {
        if (io->cloud_offsets) memcpy(io->cloud_offsets, off, sizeof(int) * (B + 1));
        if (total > io->cloud_packed_capacity) { c->err = "cloud_packed_capacity too small"; return CVR_ERR_CAPACITY; }
        if (total) CVR_CUDA(c, cudaMemcpyAsync(io->cloud_packed, pk, total * sizeof(cvr_point), cudaMemcpyDeviceToHost, st));   // ONE copy for the batch
    } else if (io->cloud) {
        for (int i = 0; i < B; i++)
            if ((size_t)cnt[i] > io->cloud_capacity) { c->err = "cloud_capacity too small"; return CVR_ERR_CAPACITY; }
        // frames that fill their slot exactly are rare: the strided layout costs one copy per frame
        for (int i = 0; i < B; i++)
            if (cnt[i] > 0)
                CVR_CUDA(c, cudaMemcpyAsync(io->cloud + (size_t)i * io->cloud_capacity, pk + off[i], (size_t)cnt[i] * sizeof(cvr_point), cudaMemcpyDeviceToHost, st));
    }
    return CVR_OK;
}

static void set_last_pack(cvr_ctx* c, const cvr_point* pk, const int* d_meta, const int* h_meta, int B, bool valid)
{
    c->lp_valid = valid; c->lp_pts = pk; c->lp_dhdr = d_meta ? d_meta + c->maxB + 1 : nullptr; c->lp_nframes = B;
    c->lp_total = (valid && h_meta) ? h_meta[B] : 0;
}

extern "C" int cvr_process_batch(cvr_ctx* c, int nframes, const cvr_batch* io)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    STEREO_ONLY(c);
    int rc = check_batch_args(c, nframes, io);
    if (rc) return rc;
    const int ch = io->channels;
    const int W = c->sg.W, H = c->sg.H, B = nframes;
    const size_t rowBytes = (size_t)W * ch;
    const int stages = io->stages ? io->stages : 15;
    const bool host = io->mem == CVR_MEM_HOST;
    CVR_CUDA(c, cudaSetDevice(c->device));
    { const int rcj_ = cvr_join_tail(c); if (rcj_) return rcj_; }
    c->lp_valid = false;
    cvr_timers_begin(c, 0);
    const uint8_t *dl, *dr; size_t lstep, lfs, rstep, rfs;
    if (host) {
        const bool dense = io->left_frame_stride == io->left_step * (size_t)H && io->right_frame_stride == io->right_step * (size_t)H;
        if (dense && io->left_step == rowBytes && io->right_step == rowBytes) {   // no row padding either: plain copies run at the full link rate
            CVR_CUDA(c, cudaMemcpyAsync(c->d_in[0], io->left, rowBytes * H * B, cudaMemcpyHostToDevice, c->stream));
            CVR_CUDA(c, cudaMemcpyAsync(c->d_in[1], io->right, rowBytes * H * B, cudaMemcpyHostToDevice, c->stream));
        } else if (dense) {                                            // frames back to back: one 2-D copy per camera for the whole batch
            CVR_CUDA(c, cudaMemcpy2DAsync(c->d_in[0], rowBytes, io->left, io->left_step, rowBytes, (size_t)H * B, cudaMemcpyHostToDevice, c->stream));
            CVR_CUDA(c, cudaMemcpy2DAsync(c->d_in[1], rowBytes, io->right, io->right_step, rowBytes, (size_t)H * B, cudaMemcpyHostToDevice, c->stream));
        } else {
            for (int i = 0; i < B; i++) {
                CVR_CUDA(c, cudaMemcpy2DAsync(c->d_in[0] + (size_t)i * rowBytes * H, rowBytes, io->left + (size_t)i * io->left_frame_stride, io->left_step, rowBytes, H, cudaMemcpyHostToDevice, c->stream));
                CVR_CUDA(c, cudaMemcpy2DAsync(c->d_in[1] + (size_t)i * rowBytes * H, rowBytes, io->right + (size_t)i * io->right_frame_stride, io->right_step, rowBytes, H, cudaMemcpyHostToDevice, c->stream));
            }
        }
        dl = c->d_in[0]; dr = c->d_in[1]; lstep = rstep = rowBytes; lfs = rfs = rowBytes * H;
        cvr_stage_mark(c, "h2d");
    } else {
        dl = io->left; dr = io->right; lstep = io->left_step; rstep = io->right_step; lfs = io->left_frame_stride; rfs = io->right_frame_stride;
    }
    const cvr_point* fin = nullptr; const int* fcnt = nullptr;
    rc = enqueue_chain(c, B, stages, dl, lstep, lfs, dr, rstep, rfs, ch, &fin, &fcnt);
    if (rc) return rc;
    // ---- outputs ----
    if (io->disp) {
        const cudaMemcpyKind kind = host ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
        const bool denseD = io->disp_frame_stride == io->disp_step * (size_t)H;
        if (denseD) CVR_CUDA(c, cudaMemcpy2DAsync(io->disp, io->disp_step, c->d_med, (size_t)W * 2, (size_t)W * 2, (size_t)H * B, kind, c->stream));
        else
            for (int i = 0; i < B; i++)
                CVR_CUDA(c, cudaMemcpy2DAsync((uint8_t*)io->disp + (size_t)i * io->disp_frame_stride, io->disp_step, c->d_med + (size_t)i * W * H, (size_t)W * 2,
                                              (size_t)W * 2, H, kind, c->stream));
    }
    if (fin) {
        cvr_point* pk = free_slab(c, fin);
        rc = enqueue_pack(c, B, fin, fcnt, pk, c->d_meta0);
        if (rc) return rc;
        if (!host) { rc = enqueue_device_outputs(c, B, io, fin, fcnt, pk, c->d_meta0); if (rc) return rc; }
        CVR_CUDA(c, cudaMemcpyAsync(c->h_meta0, c->d_meta0, CVR_META_INTS(c->maxB) * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
        rc = sync_check(c);
        if (rc) return rc;
        set_last_pack(c, pk, c->d_meta0, c->h_meta0, B, true);
        if (host) { rc = fetch_host_outputs(c, B, io, pk, c->h_meta0, c->stream); if (rc) return rc; }
        else if (c->h_meta0[2 * (size_t)c->maxB + 3]) { c->err = "a cloud did not fit the caller's device buffer (cloud_capacity / cloud_packed_capacity too small)"; return CVR_ERR_CAPACITY; }
    }
    cvr_stage_mark(c, "out");
    rc = sync_check(c);
    cvr_timers_end(c, 0);
    return rc;
}

static int pipe_init(cvr_ctx* c)
{
    if (c->pipe_ready) return CVR_OK;
    const size_t B = (size_t)c->maxB, HW = (size_t)c->sg.W * c->sg.H, N = (size_t)c->maxN;
    CVR_CUDA(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    CVR_CUDA(c, cudaStreamCreateWithFlags(&c->d2h_stream, cudaStreamNonBlocking));
    {
        int prLow = 0, prHigh = 0;                                     // numerically lowest = highest priority
        CVR_CUDA(c, cudaDeviceGetStreamPriorityRange(&prLow, &prHigh));
        CVR_CUDA(c, cudaStreamCreateWithPriority(&c->tail_stream, cudaStreamNonBlocking, prHigh));
        CVR_CUDA(c, cudaEventCreateWithFlags(&c->evPath, cudaEventDisableTiming));
        CVR_CUDA(c, cudaEventCreateWithFlags(&c->evAuxFree, cudaEventDisableTiming));
        CVR_CUDA(c, cudaEventCreateWithFlags(&c->evTailLast, cudaEventDisableTiming));
    }
    for (int s = 0; s < 2; s++) {
        PipeSlot& sl = c->slot[s];
        CVR_CUDA(c, dalloc(&sl.d_in[0], B * HW * 4)); CVR_CUDA(c, dalloc(&sl.d_in[1], B * HW * 4));
        CVR_CUDA(c, dalloc(&sl.d_out, B * N)); CVR_CUDA(c, dalloc(&sl.d_meta, CVR_META_INTS(B))); CVR_CUDA(c, dalloc(&sl.d_outdisp, B * HW));
        CVR_CUDA(c, cudaHostAlloc((void**)&sl.h_meta, CVR_META_INTS(B) * sizeof(int), cudaHostAllocDefault));
        CVR_CUDA(c, cudaEventCreateWithFlags(&sl.evH2D, cudaEventDisableTiming));
        CVR_CUDA(c, cudaEventCreateWithFlags(&sl.evDone, cudaEventDisableTiming));
        sl.busy = false; sl.used = false;
    }
    c->pipe_ready = true;
    return CVR_OK;
}

// ---- pipelined path: two batches in flight.  The upload of batch i+1 (copy stream) overlaps the kernels of batch i
// (compute stream) and the download of batch i-1 (d2h stream).  CVR_TAIL_STREAM=1 (off by default) issues everything after the
// path kernels of a batch on a high-priority stream of its own, beside the matcher of the next batch.  Measured on B200 (KITTI,
// batch 110): the tail then runs beside k_pix_hsum4, which slows down by what the tail saves (+0.6 % device-resident, +1 % end
// to end); started later, beside k_vsum / k_path_h2 / k_path_v4, it costs more than it saves (-1 .. -11 %): every matcher
// kernel already keeps the SMs or the HBM busy, the tail is not free to hide.
extern "C" int cvr_submit_batch(cvr_ctx* c, int nframes, const cvr_batch* io)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    STEREO_ONLY(c);
    int rc = check_batch_args(c, nframes, io);
    if (rc) return rc;
    CVR_CUDA(c, cudaSetDevice(c->device));
    rc = pipe_init(c);
    if (rc) return rc;
    PipeSlot& sl = c->slot[c->n_submit & 1];
    if (sl.busy) { c->err = "two batches are already in flight: call cvr_wait_batch first"; return CVR_ERR_CAPACITY; }
    const int ch = io->channels, W = c->sg.W, H = c->sg.H, B = nframes;
    const size_t rowBytes = (size_t)W * ch;
    const int stages = io->stages ? io->stages : 15;
    const bool host = io->mem == CVR_MEM_HOST;
    const uint8_t *dl, *dr; size_t lstep, lfs, rstep, rfs;
    const int tset = 1 + (c->n_submit & 1);
    if (host) {
        // upload on the copy stream, after the kernels of the batch that last used this slot's staging buffers
        if (sl.used) CVR_CUDA(c, cudaStreamWaitEvent(c->copy_stream, sl.evDone, 0));
        const bool dense = io->left_frame_stride == io->left_step * (size_t)H && io->right_frame_stride == io->right_step * (size_t)H;
        if (dense && io->left_step == rowBytes && io->right_step == rowBytes) {   // no row padding either: plain copies run at the full link rate
            CVR_CUDA(c, cudaMemcpyAsync(sl.d_in[0], io->left, rowBytes * H * B, cudaMemcpyHostToDevice, c->copy_stream));
            CVR_CUDA(c, cudaMemcpyAsync(sl.d_in[1], io->right, rowBytes * H * B, cudaMemcpyHostToDevice, c->copy_stream));
        } else if (dense) {                                            // frames back to back: one 2-D copy per camera for the whole batch
            CVR_CUDA(c, cudaMemcpy2DAsync(sl.d_in[0], rowBytes, io->left, io->left_step, rowBytes, (size_t)H * B, cudaMemcpyHostToDevice, c->copy_stream));
            CVR_CUDA(c, cudaMemcpy2DAsync(sl.d_in[1], rowBytes, io->right, io->right_step, rowBytes, (size_t)H * B, cudaMemcpyHostToDevice, c->copy_stream));
        } else {
            for (int i = 0; i < B; i++) {
                CVR_CUDA(c, cudaMemcpy2DAsync(sl.d_in[0] + (size_t)i * rowBytes * H, rowBytes, io->left + (size_t)i * io->left_frame_stride, io->left_step, rowBytes, H, cudaMemcpyHostToDevice, c->copy_stream));
                CVR_CUDA(c, cudaMemcpy2DAsync(sl.d_in[1] + (size_t)i * rowBytes * H, rowBytes, io->right + (size_t)i * io->right_frame_stride, io->right_step, rowBytes, H, cudaMemcpyHostToDevice, c->copy_stream));
            }
        }
        CVR_CUDA(c, cudaEventRecord(sl.evH2D, c->copy_stream));
        CVR_CUDA(c, cudaStreamWaitEvent(c->stream, sl.evH2D, 0));
        dl = sl.d_in[0]; dr = sl.d_in[1]; lstep = rstep = rowBytes; lfs = rfs = rowBytes * H;
    } else {                                                           // device-resident inputs: read in place (the caller keeps them until cvr_wait_batch)
        dl = io->left; dr = io->right; lstep = io->left_step; rstep = io->right_step; lfs = io->left_frame_stride; rfs = io->right_frame_stride;
    }
    cvr_timers_begin(c, tset);                                             // stage timers of this slot (read back by cvr_wait_batch)
    const cvr_point* fin = nullptr; const int* fcnt = nullptr;
    // From the matcher's hand-over point (cvr_tail_split) on, this batch is enqueued on the tail stream; c->stream is put back
    // when this function returns.  Not with rectification: the tail reads the rectified image the next batch's matcher rewrites.
    struct TailScope {
        cvr_ctx* c;
        ~TailScope() { if (c->path_stream) { c->stream = c->path_stream; c->path_stream = nullptr; } c->split_tail = false; }
    } tailScope{c};
    c->split_tail = c->tail_mode != 0 && !c->prm.requires_rectification && c->prm.block_matcher == CVR_MATCHER_SGBM;
    rc = enqueue_chain(c, B, stages, dl, lstep, lfs, dr, rstep, rfs, ch, &fin, &fcnt);
    if (rc) return rc;
    sl.hasCloud = fin != nullptr;
    if (fin) {
        // the packed payload is parked in the slot (the next batch may overwrite the workspace); it feeds the host copy,
        // device-resident packed outputs and cvr_gather_clouds
        rc = enqueue_pack(c, B, fin, fcnt, sl.d_out, sl.d_meta);
        if (rc) return rc;
        if (!host) { rc = enqueue_device_outputs(c, B, io, fin, fcnt, sl.d_out, sl.d_meta); if (rc) return rc; }
        CVR_CUDA(c, cudaMemcpyAsync(sl.h_meta, sl.d_meta, CVR_META_INTS(c->maxB) * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    }
    if (io->disp && host) CVR_CUDA(c, cudaMemcpyAsync(sl.d_outdisp, c->d_med, (size_t)B * W * H * 2, cudaMemcpyDeviceToDevice, c->stream));
    if (io->disp && !host) {
        if (io->disp_frame_stride == io->disp_step * (size_t)H)
            CVR_CUDA(c, cudaMemcpy2DAsync(io->disp, io->disp_step, c->d_med, (size_t)W * 2, (size_t)W * 2, (size_t)H * B, cudaMemcpyDeviceToDevice, c->stream));
        else
            for (int i = 0; i < B; i++)
                CVR_CUDA(c, cudaMemcpy2DAsync((uint8_t*)io->disp + (size_t)i * io->disp_frame_stride, io->disp_step, c->d_med + (size_t)i * W * H, (size_t)W * 2,
                                              (size_t)W * 2, H, cudaMemcpyDeviceToDevice, c->stream));
    }
    cvr_stage_mark(c, "out");
    CVR_CUDA(c, cudaEventRecord(sl.evDone, c->stream));
    if (c->path_stream) { CVR_CUDA(c, cudaEventRecord(c->evTailLast, c->stream)); c->tail_dirty = true; }
    c->tset = 0;
    sl.io = *io; sl.nframes = B; sl.busy = true; sl.used = true;
    c->n_submit++;
    return CVR_OK;
}

// Collects the oldest batch; whatever happens, its slot is released and the pipeline moves on.
static int wait_slot(cvr_ctx* c, PipeSlot& sl, int set)
{
    const cvr_batch& io = sl.io;
    const int B = sl.nframes, W = c->sg.W, H = c->sg.H;
    CVR_CUDA(c, cudaEventSynchronize(sl.evDone));
    cvr_timers_end(c, set);
    set_last_pack(c, sl.d_out, sl.d_meta, sl.h_meta, B, sl.hasCloud);
    if (io.mem != CVR_MEM_HOST) {                                      // device-resident batch: the results are already in the caller's buffers
        if (sl.hasCloud && sl.h_meta[2 * (size_t)c->maxB + 3]) {
            c->err = "a cloud did not fit the caller's device buffer (cloud_capacity / cloud_packed_capacity too small)";
            return CVR_ERR_CAPACITY;
        }
        return CVR_OK;
    }
    int rc = CVR_OK;
    if (sl.hasCloud) rc = fetch_host_outputs(c, B, &io, sl.d_out, sl.h_meta, c->d2h_stream);
    if (io.disp && rc == CVR_OK) {
        if (io.disp_frame_stride == io.disp_step * (size_t)H)
            CVR_CUDA(c, cudaMemcpy2DAsync(io.disp, io.disp_step, sl.d_outdisp, (size_t)W * 2, (size_t)W * 2, (size_t)H * B, cudaMemcpyDeviceToHost, c->d2h_stream));
        else
            for (int i = 0; i < B; i++)
                CVR_CUDA(c, cudaMemcpy2DAsync((uint8_t*)io.disp + (size_t)i * io.disp_frame_stride, io.disp_step, sl.d_outdisp + (size_t)i * W * H, (size_t)W * 2,
                                              (size_t)W * 2, H, cudaMemcpyDeviceToHost, c->d2h_stream));
    }
    cudaError_t e = cudaStreamSynchronize(c->d2h_stream);
    if (e != cudaSuccess && rc == CVR_OK) { c->err = std::string("d2h stream sync: ") + cudaGetErrorString(e); rc = CVR_ERR_CUDA; }
    return rc;
}

extern "C" int cvr_wait_batch(cvr_ctx* c)
{
    if (!c) return CVR_ERR_INVALID_ARG;
    if (!c->pipe_ready || c->n_wait >= c->n_submit) { c->err = "no batch in flight"; return CVR_ERR_INVALID_ARG; }
    CVR_CUDA(c, cudaSetDevice(c->device));
    PipeSlot& sl = c->slot[c->n_wait & 1];
    const int rc = wait_slot(c, sl, 1 + (int)(c->n_wait & 1));
    sl.busy = false;                                                   // also on errors: a failed batch must not wedge the pipeline
    c->n_wait++;
    return rc;
}

extern "C" int cvr_save_pcd(const char* path, const cvr_point* pts, int n)
{
    if (!path || n < 0 || (n > 0 && !pts)) { g_create_err = "cvr_save_pcd: bad arguments"; return CVR_ERR_INVALID_ARG; }
    FILE* f = fopen(path, "wb");
    if (!f) { g_create_err = std::string("cvr_save_pcd: cannot open ") + path; return CVR_ERR_INVALID_ARG; }
    // the header PCL's PCDWriter emits for pcl::PointXYZRGB (the 16 padding bytes of the in-memory point are not stored)
    fprintf(f, "# .PCD v0.7 - Point Cloud Data file format\nVERSION 0.7\nFIELDS x y z rgb\nSIZE 4 4 4 4\nTYPE F F F F\nCOUNT 1 1 1 1\n"
               "WIDTH %d\nHEIGHT 1\nVIEWPOINT 0 0 0 1 0 0 0\nPOINTS %d\nDATA binary\n", n, n);
    const size_t w = n ? fwrite(pts, sizeof(cvr_point), (size_t)n, f) : 0;
    const bool ok = w == (size_t)n && fclose(f) == 0;
    if (!ok) { g_create_err = std::string("cvr_save_pcd: short write to ") + path; return CVR_ERR_INVALID_ARG; }
    return CVR_OK;
}

extern "C" void cvr_points_to_pcl(const cvr_point* in, int n, void* pcl32)
{
    // pcl::PointXYZRGB: float x,y,z, float pad (=1.0f); then b,g,r,a packed in 4 bytes; 12 bytes of padding
    uint8_t* o = (uint8_t*)pcl32;
    for (int i = 0; i < n; i++, o += 32) {
        const float one = 1.0f;
        memcpy(o, &in[i].x, 12); memcpy(o + 12, &one, 4);
        o[16] = in[i].b; o[17] = in[i].g; o[18] = in[i].r; o[19] = in[i].a;
        memset(o + 20, 0, 12);
    }
}
