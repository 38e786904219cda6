// C ABI of libgforb.so (declared in include/gforb.h).
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <new>
#include <vector>

#include "engine.hpp"

using namespace gforb;

struct gforb_extractor {
    Tables tab;
    gforb_limits lim{};
    Engine eng;
    cudaStream_t stream = nullptr;
    int last_n = 0;
    bool extracted = false;
    // staging for host-provided key sets (gforb_stereo_match)
    gforb_keypoint* st_k = nullptr; uint8_t* st_d = nullptr; int32_t* st_n = nullptr; float* st_win = nullptr;
    int st_cap = 0;
    float *st_u = nullptr, *st_z = nullptr; int32_t *st_bd = nullptr, *st_bi = nullptr, *st_sad = nullptr, *st_nm = nullptr;
    int32_t *st_cut = nullptr, *st_ticket = nullptr;
    ~gforb_extractor() {
        free_stereo();
        eng.release();
        if (stream) cudaStreamDestroy(stream);
    }
    void free_stereo() {
        auto fr = [](auto*& p) { if (p) cudaFree(p); p = nullptr; };
        fr(st_k); fr(st_d); fr(st_n); fr(st_win); fr(st_u); fr(st_z); fr(st_bd); fr(st_bi); fr(st_sad); fr(st_nm); fr(st_cut); fr(st_ticket);
        st_cap = 0;
    }
};

struct gforb_batch {
    Tables tab;
    gforb_limits lim{};
    Engine eng;
    cudaStream_t own = nullptr, stream = nullptr;
    int max_frames = 0;
    bool timing = false;
    cudaEvent_t ev[7] = {};
    // chunked end-to-end pipeline (gforb_stereo_frames_batch): H2D, kernels and D2H of successive chunks overlap
    cudaStream_t s_up = nullptr, s_down = nullptr;
    cudaStream_t s_alt[3] = {nullptr, nullptr, nullptr};   // extra kernel streams: chunks rotate over stream + s_alt[]
    int kstreams = 4;            // kernel streams the chunk pipeline rotates over (1..4); measured best on B200
    std::vector<cudaEvent_t> ev_up, ev_done;
    cudaEvent_t ev_start = nullptr, ev_down = nullptr;
    int chunk_frames = 0;        // 0 = automatic
    bool copy_only = false;      // measurement switch: the pipeline's copies without its kernels (gforb_batch_debug_copy_only)
    int last_launches = 0;       // kernels enqueued by the last run / frames_batch call
    // Single-chunk calls (a frame or a few at a time: what a live Tracking does) are launch-bound -- 15 kernels, 7 copies
    // and a dozen event operations for ~0.25 ms of device work -- so their whole H2D -> kernels -> D2H sequence is
    // captured once into a CUDA graph per distinct argument set and replayed with one cudaGraphLaunch.
    struct GraphSlot { const void* args[8]; int n; gforb_stereo_params prm; bool has_prm; cudaGraphExec_t exec; int launches; };
    std::vector<GraphSlot> graphs;
    bool use_graphs = true;
    ~gforb_batch() {
        for (auto& gs : graphs) if (gs.exec) cudaGraphExecDestroy(gs.exec);
        for (auto& e : ev) if (e) cudaEventDestroy(e);
        for (auto& e : ev_up) if (e) cudaEventDestroy(e);
        for (auto& e : ev_done) if (e) cudaEventDestroy(e);
        if (ev_start) cudaEventDestroy(ev_start);
        if (ev_down) cudaEventDestroy(ev_down);
        eng.release();
        if (s_up) cudaStreamDestroy(s_up);
        if (s_down) cudaStreamDestroy(s_down);
        for (auto& sa : s_alt) if (sa) cudaStreamDestroy(sa);
        if (own) cudaStreamDestroy(own);
    }
};

struct gforb_matcher {
    float ratio = 0.6f;
    int sms = 148;
    cudaStream_t stream = nullptr, last = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timed = false;
    void* scratch = nullptr; size_t scratch_bytes = 0;   // per-segment partial results
    void* io = nullptr; size_t io_bytes = 0;             // staging of the host-pointer form
    void* pin = nullptr; size_t pin_bytes = 0;           // its pinned host mirror: one H2D and one D2H copy per call
    void* aux = nullptr; size_t aux_bytes = 0;           // by-products of composed entries (gforb_search_by_sim3)
    cudaEvent_t ev_spare = nullptr;
    float inv_sig2[32];                                  // mvInvLevelSigma2 for the Fuse gate (gforb_matcher_set_inv_level_sigma2)
    bool have_inv_sig2 = false;
    int reserve_pin(size_t want) {
        if (want <= pin_bytes) return GFORB_OK;
        if (pin) { cudaFreeHost(pin); pin = nullptr; pin_bytes = 0; }
        cudaError_t e = cudaHostAlloc(&pin, want, cudaHostAllocDefault);
        if (e != cudaSuccess) return cuda_fail(e, "cudaHostAlloc(matcher)");
        pin_bytes = want;
        return GFORB_OK;
    }
    int reserve(void*& p, size_t& have, size_t want) {
        if (want <= have) return GFORB_OK;
        if (p) { cudaFree(p); p = nullptr; have = 0; }
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc(matcher)");
        have = want;
        return GFORB_OK;
    }
    ~gforb_matcher() {
        if (scratch) cudaFree(scratch);
        if (io) cudaFree(io);
        if (aux) cudaFree(aux);
        if (ev_spare) cudaEventDestroy(ev_spare);
        if (pin) cudaFreeHost(pin);
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
        if (stream) cudaStreamDestroy(stream);
    }
};

namespace {
int ensure_stream(cudaStream_t& s) {
    if (s) return GFORB_OK;
    cudaError_t e = cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
    if (e != cudaSuccess) return cuda_fail(e, "cudaStreamCreate");
    return GFORB_OK;
}
int ensure_engine(gforb_extractor* ex, int w, int h) {
    int rc = ensure_stream(ex->stream);
    if (rc) return rc;
    if (ex->eng.ready && ex->eng.g.W == w && ex->eng.g.H == h) return GFORB_OK;
    ex->extracted = false;
    return ex->eng.init(ex->tab, ex->lim, w, h, 1, 1);
}
gforb_stereo_params norm_params(const gforb_stereo_params& p) {
    gforb_stereo_params q = p;
    if (q.th_high <= 0) q.th_high = GFORB_TH_HIGH;
    if (q.th_low <= 0) q.th_low = GFORB_TH_LOW;
    q.th_high = std::min(q.th_high, 256);
    return q;
}
void fill_view_levels(StereoView& v, const Engine& L, const Engine& R, size_t frame_mul, size_t right_img_off) {
    const Geom& g = L.g;
    for (int l = 0; l < g.nlevels; ++l) {
        const LevelGeom& a = g.lv[l];
        const LevelGeom& c = R.g.lv[l];
        v.pyrL[l] = L.b.pyr + a.off + (size_t)kEdge * a.pitch + kXPad;
        v.pyrR[l] = R.b.pyr + c.off + right_img_off * c.img_stride + (size_t)kEdge * c.pitch + kXPad;
        v.frame_stride[l] = frame_mul * a.img_stride;
        v.pitch[l] = a.pitch; v.w[l] = a.w;
        v.scale[l] = a.scale; v.inv_scale[l] = a.inv_scale;
    }
    v.nRows = g.lv[0].h;
}
}  // namespace

extern "C" {

const char* gforb_last_error_string(void) { return last_error(); }
int gforb_version(void) { return GFORB_VERSION; }

int gforb_hamming256(const void* a, const void* b) {
    uint64_t x[4], y[4];
    std::memcpy(x, a, 32); std::memcpy(y, b, 32);
    int d = 0;
    for (int i = 0; i < 4; ++i) d += __builtin_popcountll(x[i] ^ y[i]);
    return d;
}

int gforb_host_alloc(void** ptr, size_t bytes) {
    if (!ptr) return GFORB_ERR_ARG;
    cudaError_t e = cudaHostAlloc(ptr, bytes, cudaHostAllocDefault);
    return e == cudaSuccess ? GFORB_OK : cuda_fail(e, "cudaHostAlloc");
}
void gforb_host_free(void* ptr) { if (ptr) cudaFreeHost(ptr); }

// --------------------------------------------------------------------------------------------- extractor
int gforb_extractor_create(const gforb_params* p, const gforb_limits* lim, gforb_extractor** out) {
    if (!p || !out) { set_error("gforb_extractor_create: null argument"); return GFORB_ERR_ARG; }
    std::unique_ptr<gforb_extractor> ex(new (std::nothrow) gforb_extractor);
    if (!ex) return GFORB_ERR_ARG;
    int rc = make_tables(*p, ex->tab);
    if (rc) return rc;
    if (lim) ex->lim = *lim;
    *out = ex.release();
    return GFORB_OK;
}
void gforb_extractor_destroy(gforb_extractor* ex) { delete ex; }

int gforb_extractor_tables(const gforb_extractor* ex, float* scale, float* inv_scale, float* sigma2,
                           float* inv_sigma2, int32_t* fpl, int32_t* umax) {
    if (!ex) return GFORB_ERR_ARG;
    const size_t n = (size_t)ex->tab.prm.nlevels;
    if (scale) std::memcpy(scale, ex->tab.scale.data(), n * 4);
    if (inv_scale) std::memcpy(inv_scale, ex->tab.inv_scale.data(), n * 4);
    if (sigma2) std::memcpy(sigma2, ex->tab.sigma2.data(), n * 4);
    if (inv_sigma2) std::memcpy(inv_sigma2, ex->tab.inv_sigma2.data(), n * 4);
    if (fpl) std::memcpy(fpl, ex->tab.quota.data(), n * 4);
    if (umax) std::memcpy(umax, ex->tab.umax, 64);
    return GFORB_OK;
}
int gforb_extractor_params(const gforb_extractor* ex, gforb_params* out) {
    if (!ex || !out) return GFORB_ERR_ARG;
    *out = ex->tab.prm;
    return GFORB_OK;
}
int gforb_extractor_max_keypoints(const gforb_extractor* ex, int w, int h) {
    if (!ex) return GFORB_ERR_ARG;
    Geom g; std::vector<int32_t> tabs; size_t a, b;
    int rc = make_geometry(ex->tab, ex->lim, w, h, 1, g, tabs, a, b);
    return rc ? rc : g.K;
}

int gforb_compute_pyramid(gforb_extractor* ex, const uint8_t* image, int w, int h, size_t pitch) {
    if (!ex) return GFORB_ERR_ARG;
    if (!image || w <= 0 || h <= 0) return GFORB_OK;   // reference: no-op on empty input
    if (pitch < (size_t)w) { set_error("gforb: pitch < width"); return GFORB_ERR_ARG; }
    int rc = ensure_engine(ex, w, h);
    if (rc) return rc;
    rc = ex->eng.upload_images(image, pitch, 0, 1, 1, ex->stream);
    if (rc) return rc;
    rc = ex->eng.run_pyramid(1, ex->stream);
    if (rc) return rc;
    ex->extracted = false;
    cudaError_t e = cudaStreamSynchronize(ex->stream);
    return e == cudaSuccess ? GFORB_OK : cuda_fail(e, "cudaStreamSynchronize");
}

int gforb_extract(gforb_extractor* ex, const uint8_t* image, int w, int h, size_t pitch,
                  gforb_keypoint* kps, uint8_t* desc, int cap, int* n) {
    if (!ex || !n) { set_error("gforb_extract: null argument"); return GFORB_ERR_ARG; }
    *n = 0;
    if (!image || w <= 0 || h <= 0) return GFORB_OK;   // src/ORBextractor.cc:1115-1116
    if (pitch < (size_t)w || cap < 0 || (cap > 0 && (!kps || !desc))) {
        set_error("gforb_extract: bad pitch/capacity/buffers");
        return GFORB_ERR_ARG;
    }
    int rc = ensure_engine(ex, w, h);
    if (rc) return rc;
    Engine& E = ex->eng;
    rc = E.upload_images(image, pitch, 0, 1, 1, ex->stream);
    if (rc) return rc;
    rc = E.run_extract(1, ex->stream);
    if (rc) return rc;
    cudaError_t e = cudaMemcpyAsync(E.h_small + 4, E.b.nkp, 4, cudaMemcpyDeviceToHost, ex->stream);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpyAsync(nkp)");
    rc = E.check_errors(ex->stream);   // synchronises
    if (rc) return rc;
    const int cnt = E.h_small[4];
    *n = cnt;
    ex->last_n = cnt;
    ex->extracted = true;
    if (cnt > cap) { set_error("gforb_extract: keypoint capacity too small"); return GFORB_ERR_CAPACITY; }
    if (cnt > 0) {
        e = cudaMemcpyAsync(kps, E.b.kps, (size_t)cnt * sizeof(gforb_keypoint), cudaMemcpyDeviceToHost, ex->stream);
        if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpyAsync(kps)");
        e = cudaMemcpyAsync(desc, E.b.desc, (size_t)cnt * 32, cudaMemcpyDeviceToHost, ex->stream);
        if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpyAsync(desc)");
        e = cudaStreamSynchronize(ex->stream);
        if (e != cudaSuccess) return cuda_fail(e, "cudaStreamSynchronize");
    }
    return GFORB_OK;
}

int gforb_pyramid_level(const gforb_extractor* ex, int level, const uint8_t** dev_ptr, int* w, int* h, size_t* pitch) {
    if (!ex || !ex->eng.ready || !ex->eng.pyramid_valid) { set_error("gforb: no pyramid yet"); return GFORB_ERR_STATE; }
    if (level < 0 || level >= ex->eng.g.nlevels) return GFORB_ERR_ARG;
    const LevelGeom& L = ex->eng.g.lv[level];
    if (dev_ptr) *dev_ptr = ex->eng.b.pyr + L.off + (size_t)kEdge * L.pitch + kXPad;
    if (w) *w = L.w;
    if (h) *h = L.h;
    if (pitch) *pitch = (size_t)L.pitch;
    return GFORB_OK;
}

int gforb_download_pyramid(gforb_extractor* ex, int level, uint8_t* host, size_t host_pitch, int with_border) {
    if (!ex || !host) return GFORB_ERR_ARG;
    if (!ex->eng.ready || !ex->eng.pyramid_valid) { set_error("gforb: no pyramid yet"); return GFORB_ERR_STATE; }
    if (level < 0 || level >= ex->eng.g.nlevels) return GFORB_ERR_ARG;
    const LevelGeom& L = ex->eng.g.lv[level];
    const int bw = with_border ? L.w + 2 * kEdge : L.w, bh = with_border ? L.h + 2 * kEdge : L.h;
    if (host_pitch < (size_t)bw) return GFORB_ERR_ARG;
    const uint8_t* src = ex->eng.b.pyr + L.off +
                         (with_border ? (size_t)(kXPad - kEdge) : (size_t)kEdge * L.pitch + kXPad);
    cudaError_t e = cudaMemcpy2DAsync(host, host_pitch, src, L.pitch, bw, bh, cudaMemcpyDeviceToHost, ex->stream);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpy2DAsync(pyramid)");
    e = cudaStreamSynchronize(ex->stream);
    return e == cudaSuccess ? GFORB_OK : cuda_fail(e, "cudaStreamSynchronize");
}

int gforb_debug_blurred_level(gforb_extractor* ex, int level, uint8_t* host, size_t host_pitch) {
    if (!ex || !host) return GFORB_ERR_ARG;
    if (!ex->eng.ready || !ex->extracted) { set_error("gforb: no extraction yet"); return GFORB_ERR_STATE; }
    if (level < 0 || level >= ex->eng.g.nlevels) return GFORB_ERR_ARG;
    const LevelGeom& L = ex->eng.g.lv[level];
    if (host_pitch < (size_t)L.w) return GFORB_ERR_ARG;
    cudaError_t e = cudaMemcpy2DAsync(host, host_pitch, ex->eng.b.blur + L.boff, L.bpitch, L.w, L.h, cudaMemcpyDeviceToHost, ex->stream);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpy2DAsync(blurred level)");
    e = cudaStreamSynchronize(ex->stream);
    return e == cudaSuccess ? GFORB_OK : cuda_fail(e, "cudaStreamSynchronize");
}

int gforb_debug_candidates(gforb_extractor* ex, int level, int32_t* xyr, int cap, int* n) {
    if (!ex || !n) return GFORB_ERR_ARG;
    if (!ex->eng.ready || !ex->extracted) { set_error("gforb: no extraction yet"); return GFORB_ERR_STATE; }
    const Geom& g = ex->eng.g;
    if (level < 0 || level >= g.nlevels) return GFORB_ERR_ARG;
    const LevelGeom& L = g.lv[level];
    std::vector<uint32_t> cnt((size_t)std::max(L.ncells, 1)), pts((size_t)std::max(L.ncells, 1) * g.cell_cap);
    if (L.ncells > 0) {
        cudaError_t e = cudaMemcpy(cnt.data(), ex->eng.b.cell_cnt + L.cell_base, (size_t)L.ncells * 4, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpy(cell_cnt)");
        e = cudaMemcpy(pts.data(), ex->eng.b.cell_pts + (size_t)L.cell_base * g.cell_cap,
                       (size_t)L.ncells * g.cell_cap * 4, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpy(cell_pts)");
    }
    int k = 0;
    for (int c = 0; c < L.ncells; ++c)
        for (uint32_t i = 0; i < cnt[c]; ++i, ++k)
            if (k < cap && xyr) {
                const uint32_t p = pts[(size_t)c * g.cell_cap + i];
                xyr[3 * k] = cand_x(p); xyr[3 * k + 1] = cand_y(p); xyr[3 * k + 2] = cand_r(p);
            }
    *n = k;
    return k > cap && xyr ? GFORB_ERR_CAPACITY : GFORB_OK;
}

int gforb_debug_level_counts(gforb_extractor* ex, int32_t* counts) {
    if (!ex || !counts) return GFORB_ERR_ARG;
    if (!ex->eng.ready || !ex->extracted) { set_error("gforb: no extraction yet"); return GFORB_ERR_STATE; }
    cudaError_t e = cudaMemcpy(counts, ex->eng.b.lvl_cnt, (size_t)ex->eng.g.nlevels * 4, cudaMemcpyDeviceToHost);
    return e == cudaSuccess ? GFORB_OK : cuda_fail(e, "cudaMemcpy(lvl_cnt)");
}

// ------------------------------------------------------------------------------------------------ stereo
int gforb_stereo_match(gforb_extractor* exL, gforb_extractor* exR, const gforb_stereo_params* prm,
                       const gforb_keypoint* kL, int nL, const uint8_t* dL,
                       const gforb_keypoint* kR, int nR, const uint8_t* dR,
                       const float* min_d, const float* max_d, float* u_right, float* depth, int* n_matched) {
    if (!exL || !exR || !prm || nL < 0 || nR < 0 || (nL > 0 && (!kL || !dL || !u_right || !depth)) ||
        (nR > 0 && (!kR || !dR)) || ((min_d == nullptr) != (max_d == nullptr))) {
        set_error("gforb_stereo_match: bad arguments");
        return GFORB_ERR_ARG;
    }
    if (!exL->eng.ready || !exR->eng.ready || !exL->eng.pyramid_valid || !exR->eng.pyramid_valid) {
        set_error("gforb_stereo_match: both extractors need a pyramid (extract or compute_pyramid first)");
        return GFORB_ERR_STATE;
    }
    const Geom& gl = exL->eng.g; const Geom& gr = exR->eng.g;
    if (gl.W != gr.W || gl.H != gr.H || gl.nlevels != gr.nlevels) {
        set_error("gforb_stereo_match: left/right pyramids differ in geometry");
        return GFORB_ERR_ARG;
    }
    if (nL > 65535 || nR > 65535) { set_error("gforb_stereo_match: more than 65535 keys"); return GFORB_ERR_ARG; }
    for (int i = 0; i < nL; ++i) if (kL[i].octave < 0 || kL[i].octave >= gl.nlevels) { set_error("gforb_stereo_match: bad octave"); return GFORB_ERR_ARG; }
    for (int i = 0; i < nR; ++i) if (kR[i].octave < 0 || kR[i].octave >= gl.nlevels) { set_error("gforb_stereo_match: bad octave"); return GFORB_ERR_ARG; }
    if (n_matched) *n_matched = 0;
    if (nL == 0) return GFORB_OK;

    const int need = std::max(std::max(nL, nR), 8);
    cudaStream_t s = exL->stream;
    cudaError_t e;
    if (need > exL->st_cap) {
        exL->free_stereo();
        const size_t K = (size_t)need;
#define GA(p, bytes) do { e = cudaMalloc(&(p), (bytes)); if (e != cudaSuccess) { exL->free_stereo(); return cuda_fail(e, "cudaMalloc(stereo staging)"); } } while (0)
        GA(exL->st_k, 2 * K * sizeof(gforb_keypoint)); GA(exL->st_d, 2 * K * 32); GA(exL->st_n, 16);
        GA(exL->st_win, 2 * K * 4); GA(exL->st_u, K * 4); GA(exL->st_z, K * 4); GA(exL->st_bd, K * 4);
        GA(exL->st_bi, K * 4); GA(exL->st_sad, K * 4); GA(exL->st_nm, 16);
        GA(exL->st_cut, K * 4); GA(exL->st_ticket, 16);
        cudaMemsetAsync(exL->st_ticket, 0, 16, s);
#undef GA
        exL->st_cap = need;
    }
    const size_t K = (size_t)exL->st_cap;
    // the right pyramid is produced on exR's stream: make sure it is complete
    e = cudaStreamSynchronize(exR->stream);
    if (e != cudaSuccess) return cuda_fail(e, "cudaStreamSynchronize(right)");
    int32_t nn[2] = {nL, nR};
    cudaMemcpyAsync(exL->st_k, kL, (size_t)nL * sizeof(gforb_keypoint), cudaMemcpyHostToDevice, s);
    cudaMemcpyAsync(exL->st_d, dL, (size_t)nL * 32, cudaMemcpyHostToDevice, s);
    if (nR > 0) {
        cudaMemcpyAsync(exL->st_k + K, kR, (size_t)nR * sizeof(gforb_keypoint), cudaMemcpyHostToDevice, s);
        cudaMemcpyAsync(exL->st_d + K * 32, dR, (size_t)nR * 32, cudaMemcpyHostToDevice, s);
    }
    cudaMemcpyAsync(exL->st_n, nn, 8, cudaMemcpyHostToDevice, s);
    if (min_d) {
        cudaMemcpyAsync(exL->st_win, min_d, (size_t)nL * 4, cudaMemcpyHostToDevice, s);
        cudaMemcpyAsync(exL->st_win + K, max_d, (size_t)nL * 4, cudaMemcpyHostToDevice, s);
    }
    StereoView v{};
    v.kL = exL->st_k; v.kR = exL->st_k + K; v.dL = exL->st_d; v.dR = exL->st_d + K * 32;
    v.nL = exL->st_n; v.nR = exL->st_n + 1; v.k_stride = 0; v.n_stride = 0;
    fill_view_levels(v, exL->eng, exR->eng, 0, 0);
    v.min_d = min_d ? exL->st_win : nullptr; v.max_d = min_d ? exL->st_win + K : nullptr;
    v.uright = exL->st_u; v.depth = exL->st_z; v.best_dist = exL->st_bd; v.best_idx = exL->st_bi;
    v.sad = exL->st_sad; v.nmatch = exL->st_nm; v.out_stride = 0;
    v.cut_dist = exL->st_cut; v.ticket = exL->st_ticket; v.proc_count = exL->st_ticket + 1;
    GFORB_RANGE_PUSH("gforb ComputeStereoMatches", 5);
    launch_stereo(v, norm_params(*prm), 1, (int)K, s);
    GFORB_RANGE_POP();
    e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "stereo launch");
    int32_t nm = 0;
    cudaMemcpyAsync(u_right, exL->st_u, (size_t)nL * 4, cudaMemcpyDeviceToHost, s);
    cudaMemcpyAsync(depth, exL->st_z, (size_t)nL * 4, cudaMemcpyDeviceToHost, s);
    cudaMemcpyAsync(&nm, exL->st_nm, 4, cudaMemcpyDeviceToHost, s);
    e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return cuda_fail(e, "cudaStreamSynchronize(stereo)");
    if (n_matched) *n_matched = nm;
    return GFORB_OK;
}

int gforb_debug_stereo(gforb_extractor* exL, int nL, int32_t* bd, int32_t* bi, int32_t* sad) {
    if (!exL || nL < 0 || nL > exL->st_cap) return GFORB_ERR_ARG;
    cudaError_t e = cudaSuccess;
    if (bd) e = cudaMemcpy(bd, exL->st_bd, (size_t)nL * 4, cudaMemcpyDeviceToHost);
    if (bi && e == cudaSuccess) e = cudaMemcpy(bi, exL->st_bi, (size_t)nL * 4, cudaMemcpyDeviceToHost);
    if (sad && e == cudaSuccess) e = cudaMemcpy(sad, exL->st_sad, (size_t)nL * 4, cudaMemcpyDeviceToHost);
    return e == cudaSuccess ? GFORB_OK : cuda_fail(e, "cudaMemcpy(stereo debug)");
}

// ----------------------------------------------------------------------------------------------- matcher
#define M_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return cuda_fail(e_, #call); } while (0)

int gforb_matcher_create(float nnratio, gforb_matcher** out) {
    if (!out || !(nnratio > 0.f)) { set_error("gforb_matcher_create: bad arguments"); return GFORB_ERR_ARG; }
    int dev = 0, sms = 0;
    M_CUDA(cudaGetDevice(&dev));
    M_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    std::unique_ptr<gforb_matcher> m(new (std::nothrow) gforb_matcher);
    if (!m) return GFORB_ERR_ARG;
    m->ratio = nnratio; m->sms = sms;
    int rc = ensure_stream(m->stream);
    if (rc) return rc;
    M_CUDA(cudaEventCreate(&m->ev0));
    M_CUDA(cudaEventCreate(&m->ev1));
    M_CUDA(cudaEventCreate(&m->ev_spare));
    *out = m.release();
    return GFORB_OK;
}
void gforb_matcher_destroy(gforb_matcher* m) { delete m; }

int gforb_match_bruteforce_device(gforb_matcher* m, const uint8_t* dA, int nA, const uint8_t* dB, int nB, int threshold,
                                  int32_t* best_idx, int32_t* best_dist, int32_t* second_dist, int32_t* match,
                                  void* cuda_stream) {
    if (!m || nA < 0 || nB < 0 || (nA > 0 && !dA) || (nB > 0 && !dB)) { set_error("gforb_match_bruteforce: bad arguments"); return GFORB_ERR_ARG; }
    if (nB > match_max_train()) { set_error("gforb_match_bruteforce: n_b exceeds 4,194,304"); return GFORB_ERR_CAPACITY; }
    if (((uintptr_t)dA | (uintptr_t)dB) & 15) { set_error("gforb_match_bruteforce: descriptor arrays must be 16-byte aligned"); return GFORB_ERR_ARG; }
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : m->stream;
    m->last = s;
    if (match_tc_eligible(nA, nB)) {                 // large sets: the contraction runs on the tensor cores (k_match_tc.cu)
        const int tseg = match_tc_segments(nA, nB, m->sms);
        int rct = m->reserve(m->scratch, m->scratch_bytes, match_tc_scratch_bytes(nA, nB, tseg));
        if (rct) return rct;
        M_CUDA(cudaEventRecord(m->ev0, s));
        if (launch_match_tc(dA, nA, dB, nB, m->ratio, threshold, tseg, m->scratch, best_idx, best_dist, second_dist, match, s) == 0) {
            M_CUDA(cudaEventRecord(m->ev1, s));
            M_CUDA(cudaGetLastError());
            m->timed = true;
            return GFORB_OK;
        }
        (void)cudaGetLastError();                    // tensor map / opt-in refused: the POPC kernels below
    }
    const int nseg = match_segments(nA, nB, m->sms);
    int rc = m->reserve(m->scratch, m->scratch_bytes, match_scratch_bytes(nA, nseg));
    if (rc) return rc;
    M_CUDA(cudaEventRecord(m->ev0, s));
    launch_match(dA, nA, dB, nB, m->ratio, threshold, nseg, m->scratch, best_idx, best_dist, second_dist, match, s);
    M_CUDA(cudaEventRecord(m->ev1, s));
    M_CUDA(cudaGetLastError());
    m->timed = true;
    return GFORB_OK;
}

int gforb_matcher_sync(gforb_matcher* m, float* last_ms) {
    if (!m) return GFORB_ERR_ARG;
    M_CUDA(cudaStreamSynchronize(m->last ? m->last : m->stream));
    if (last_ms) {
        *last_ms = 0.f;
        if (m->timed) M_CUDA(cudaEventElapsedTime(last_ms, m->ev0, m->ev1));
    }
    return GFORB_OK;
}

int gforb_match_bruteforce(gforb_matcher* m, const uint8_t* a, int nA, const uint8_t* b, int nB, int threshold,
                           int32_t* best_idx, int32_t* best_dist, int32_t* second_dist, int32_t* match) {
    if (!m || nA < 0 || nB < 0 || (nA > 0 && !a) || (nB > 0 && !b)) { set_error("gforb_match_bruteforce: bad arguments"); return GFORB_ERR_ARG; }
    if (nA == 0) return GFORB_OK;
    const size_t ba = (size_t)nA * 32, bb = (size_t)std::max(nB, 1) * 32, bo = (size_t)nA * 4;
    int rc = m->reserve(m->io, m->io_bytes, ba + bb + 4 * bo + 64);
    if (rc) return rc;
    uint8_t* dA = static_cast<uint8_t*>(m->io);
    uint8_t* dB = dA + ((ba + 15) & ~(size_t)15);
    int32_t* o = reinterpret_cast<int32_t*>(dB + ((bb + 15) & ~(size_t)15));
    cudaStream_t s = m->stream;
    M_CUDA(cudaMemcpyAsync(dA, a, ba, cudaMemcpyHostToDevice, s));
    if (nB > 0) M_CUDA(cudaMemcpyAsync(dB, b, (size_t)nB * 32, cudaMemcpyHostToDevice, s));
    rc = gforb_match_bruteforce_device(m, dA, nA, dB, nB, threshold, o, o + nA, o + 2 * (size_t)nA, o + 3 * (size_t)nA, nullptr);
    if (rc) return rc;
    int32_t* outs[4] = {best_idx, best_dist, second_dist, match};
    for (int i = 0; i < 4; ++i)
        if (outs[i]) M_CUDA(cudaMemcpyAsync(outs[i], o + (size_t)i * nA, bo, cudaMemcpyDeviceToHost, s));
    M_CUDA(cudaStreamSynchronize(s));
    return GFORB_OK;
}

// ------------------------------------------------------------------------ frame grid + SearchForInitialization
namespace {
int check_bounds(const gforb_frame_bounds* b, const char* who) {
    if (!b || !(b->max_x > b->min_x) || !(b->max_y > b->min_y)) { set_error(std::string(who) + ": bad image bounds"); return GFORB_ERR_ARG; }
    return GFORB_OK;
}
inline size_t al16(size_t v) { return (v + 15) & ~(size_t)15; }
}  // namespace

int gforb_assign_features_to_grid(gforb_matcher* m, const gforb_frame_bounds* bounds, const gforb_keypoint* keys, int n,
                                  int32_t* cell_start, int32_t* cell_idx) {
    if (!m || n < 0 || (n > 0 && (!keys || !cell_idx)) || !cell_start) { set_error("gforb_assign_features_to_grid: bad arguments"); return GFORB_ERR_ARG; }
    int rc = check_bounds(bounds, "gforb_assign_features_to_grid");
    if (rc) return rc;
    if (n > gridmatch_max_keys()) { set_error("gforb_assign_features_to_grid: more than 8192 keys"); return GFORB_ERR_CAPACITY; }
    const int cells = gridmatch_cells();
    const size_t bk = al16((size_t)std::max(n, 1) * sizeof(gforb_keypoint)), bs = al16((size_t)(cells + 1) * 4), bi = al16((size_t)std::max(n, 1) * 4);
    rc = m->reserve(m->io, m->io_bytes, bk + bs + bi);
    if (rc) return rc;
    uint8_t* base = static_cast<uint8_t*>(m->io);
    gforb_keypoint* dk = reinterpret_cast<gforb_keypoint*>(base);
    int32_t* ds = reinterpret_cast<int32_t*>(base + bk);
    int32_t* di = reinterpret_cast<int32_t*>(base + bk + bs);
    cudaStream_t s = m->stream;
    const float bb[4] = {bounds->min_x, bounds->max_x, bounds->min_y, bounds->max_y};
    if (n > 0) M_CUDA(cudaMemcpyAsync(dk, keys, (size_t)n * sizeof(gforb_keypoint), cudaMemcpyHostToDevice, s));
    launch_grid_build(dk, nullptr, n, std::max(n, 1), 1, bb, ds, di, s);
    M_CUDA(cudaGetLastError());
    M_CUDA(cudaMemcpyAsync(cell_start, ds, (size_t)(cells + 1) * 4, cudaMemcpyDeviceToHost, s));
    M_CUDA(cudaStreamSynchronize(s));
    const int inside = cell_start[cells];
    if (inside > 0) M_CUDA(cudaMemcpy(cell_idx, di, (size_t)inside * 4, cudaMemcpyDeviceToHost));
    return GFORB_OK;
}

int gforb_features_in_area(gforb_matcher* m, const gforb_frame_bounds* bounds, const gforb_keypoint* keys, int n,
                           const gforb_area_query* queries, int nq, int32_t* out_start, int32_t* out_idx, int cap) {
    if (!m || n < 0 || nq < 0 || cap < 0 || (n > 0 && !keys) || (nq > 0 && !queries) || !out_start || (cap > 0 && !out_idx)) {
        set_error("gforb_features_in_area: bad arguments");
        return GFORB_ERR_ARG;
    }
    int rc = check_bounds(bounds, "gforb_features_in_area");
    if (rc) return rc;
    if (n > gridmatch_max_keys()) { set_error("gforb_features_in_area: more than 8192 keys"); return GFORB_ERR_CAPACITY; }
    out_start[0] = 0;
    if (nq == 0) return GFORB_OK;
    const int cells = gridmatch_cells();
    const size_t bk = al16((size_t)std::max(n, 1) * sizeof(gforb_keypoint)), bs = al16((size_t)(cells + 1) * 4), bi = al16((size_t)std::max(n, 1) * 4);
    const size_t bq = al16((size_t)nq * sizeof(gforb_area_query)), bc = al16((size_t)(nq + 1) * 4), bo = al16((size_t)std::max(cap, 1) * 4);
    rc = m->reserve(m->io, m->io_bytes, bk + bs + bi + bq + bc + bo);
    if (rc) return rc;
    uint8_t* base = static_cast<uint8_t*>(m->io);
    gforb_keypoint* dk = reinterpret_cast<gforb_keypoint*>(base);
    int32_t* ds = reinterpret_cast<int32_t*>(base + bk);
    int32_t* di = reinterpret_cast<int32_t*>(base + bk + bs);
    void* dq = base + bk + bs + bi;
    int32_t* dc = reinterpret_cast<int32_t*>(base + bk + bs + bi + bq);
    int32_t* dout = reinterpret_cast<int32_t*>(base + bk + bs + bi + bq + bc);
    cudaStream_t s = m->stream;
    const float bb[4] = {bounds->min_x, bounds->max_x, bounds->min_y, bounds->max_y};
    if (n > 0) M_CUDA(cudaMemcpyAsync(dk, keys, (size_t)n * sizeof(gforb_keypoint), cudaMemcpyHostToDevice, s));
    M_CUDA(cudaMemcpyAsync(dq, queries, (size_t)nq * sizeof(gforb_area_query), cudaMemcpyHostToDevice, s));
    launch_grid_build(dk, nullptr, n, std::max(n, 1), 1, bb, ds, di, s);
    launch_area_query(dk, bb, ds, di, dq, nq, dc, nullptr, nullptr, s);                 // pass 1: counts
    M_CUDA(cudaGetLastError());
    M_CUDA(cudaMemcpyAsync(out_start + 1, dc, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
    M_CUDA(cudaStreamSynchronize(s));
    long long total = 0;
    for (int q = 0; q < nq; ++q) { total += out_start[q + 1]; out_start[q + 1] = (int32_t)std::min<long long>(total, INT32_MAX); }
    if (total > cap) { set_error("gforb_features_in_area: output capacity too small (out_start holds the sizes)"); return GFORB_ERR_CAPACITY; }
    if (total == 0) return GFORB_OK;
    M_CUDA(cudaMemcpyAsync(dc, out_start, (size_t)(nq + 1) * 4, cudaMemcpyHostToDevice, s));
    launch_area_query(dk, bb, ds, di, dq, nq, nullptr, dc, dout, s);                    // pass 2: ordered fill
    M_CUDA(cudaGetLastError());
    M_CUDA(cudaMemcpyAsync(out_idx, dout, (size_t)total * 4, cudaMemcpyDeviceToHost, s));
    M_CUDA(cudaStreamSynchronize(s));
    return GFORB_OK;
}

int gforb_search_for_initialization_batch_device(gforb_matcher* m, const gforb_frame_bounds* bounds, int n_pairs,
                                    const gforb_keypoint* k1, const uint8_t* d1, const int32_t* n1, int stride1,
                                    const gforb_keypoint* k2, const uint8_t* d2, const int32_t* n2, int stride2,
                                    float* prev_xy, int window_size, int check_orientation,
                                    int32_t* matches12, int32_t* nmatches, void* cuda_stream) {
    if (!m || n_pairs < 0 || stride1 < 0 || stride2 < 0 || !nmatches || (stride1 > 0 && (!k1 || !d1 || !prev_xy || !matches12)) ||
        (stride2 > 0 && (!k2 || !d2))) {
        set_error("gforb_search_for_initialization: bad arguments");
        return GFORB_ERR_ARG;
    }
    int rc = check_bounds(bounds, "gforb_search_for_initialization");
    if (rc) return rc;
    if (stride1 > gridmatch_max_keys() || stride2 > gridmatch_max_keys()) { set_error("gforb_search_for_initialization: more than 8192 keys per frame"); return GFORB_ERR_CAPACITY; }
    if (((uintptr_t)d1 | (uintptr_t)d2) & 15) { set_error("gforb_search_for_initialization: descriptor arrays must be 16-byte aligned"); return GFORB_ERR_ARG; }
    if (n_pairs == 0) return GFORB_OK;
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : m->stream;
    m->last = s;
    const int cells = gridmatch_cells(), st2 = std::max(stride2, 1);
    const size_t bs = al16((size_t)n_pairs * (cells + 1) * 4), bi = al16((size_t)n_pairs * st2 * 4);
    rc = m->reserve(m->scratch, m->scratch_bytes, bs + bi + (size_t)n_pairs * std::max(stride1, 1) * 32);
    if (rc) return rc;
    int32_t* ds = static_cast<int32_t*>(m->scratch);
    int32_t* di = reinterpret_cast<int32_t*>(static_cast<uint8_t*>(m->scratch) + bs);
    const float bb[4] = {bounds->min_x, bounds->max_x, bounds->min_y, bounds->max_y};
    M_CUDA(cudaEventRecord(m->ev0, s));
    launch_grid_build(k2, n2, stride2, st2, n_pairs, bb, ds, di, s);
    InitSearchArgs a{};
    a.k1 = k1; a.d1 = d1; a.n1 = n1; a.k2 = k2; a.d2 = d2; a.n2 = n2;
    a.n1_fixed = stride1; a.n2_fixed = stride2; a.stride1 = stride1; a.stride2 = st2;
    a.cell_start = ds; a.cell_idx = di;
    a.prev_xy = prev_xy; a.window = (float)window_size; a.ratio = m->ratio;
    a.check_ori = check_orientation ? 1 : 0; a.th_low = GFORB_TH_LOW;
    a.matches12 = matches12; a.nmatches = nmatches;
    a.spec = reinterpret_cast<uint4*>(static_cast<uint8_t*>(m->scratch) + bs + bi);
    if (launch_init_search(a, bb, n_pairs, s)) { set_error("gforb_search_for_initialization: shared-memory opt-in failed"); return GFORB_ERR_CUDA; }
    M_CUDA(cudaEventRecord(m->ev1, s));
    M_CUDA(cudaGetLastError());
    m->timed = true;
    return GFORB_OK;
}

int gforb_search_for_initialization(gforb_matcher* m, const gforb_frame_bounds* bounds, const gforb_keypoint* keys1,
                                    const uint8_t* desc1, int n1, const gforb_keypoint* keys2, const uint8_t* desc2, int n2,
                                    float* prev_xy, int window_size, int check_orientation, int32_t* matches12, int* nmatches) {
    if (!m || n1 < 0 || n2 < 0 || !nmatches || (n1 > 0 && (!keys1 || !desc1 || !prev_xy || !matches12)) || (n2 > 0 && (!keys2 || !desc2))) {
        set_error("gforb_search_for_initialization: bad arguments");
        return GFORB_ERR_ARG;
    }
    *nmatches = 0;
    if (n1 == 0) return check_bounds(bounds, "gforb_search_for_initialization");
    // inputs packed into one pinned block (one H2D), outputs into another (one D2H); vbPrevMatched is in both
    struct Seg { const void* src; size_t bytes, off; };
    Seg in[5] = {{keys1, (size_t)n1 * sizeof(gforb_keypoint), 0}, {desc1, (size_t)n1 * 32, 0}, {keys2, (size_t)n2 * sizeof(gforb_keypoint), 0},
                 {desc2, (size_t)n2 * 32, 0}, {prev_xy, (size_t)n1 * 8, 0}};
    size_t total = 0;
    for (Seg& g : in) { g.off = total; total += al16(std::max<size_t>(g.bytes, 16)); }
    const size_t in_bytes = total, om = total, on = om + al16((size_t)n1 * 4);
    total = on + 16;
    int rc = m->reserve(m->io, m->io_bytes, total);
    if (rc) return rc;
    rc = m->reserve_pin(total);
    if (rc) return rc;
    uint8_t* dev = static_cast<uint8_t*>(m->io);
    uint8_t* pin = static_cast<uint8_t*>(m->pin);
    for (const Seg& g : in)
        if (g.bytes) std::memcpy(pin + g.off, g.src, g.bytes);
    cudaStream_t s = m->stream;
    M_CUDA(cudaMemcpyAsync(dev, pin, in_bytes, cudaMemcpyHostToDevice, s));
    float* dp = reinterpret_cast<float*>(dev + in[4].off);
    rc = gforb_search_for_initialization_batch_device(m, bounds, 1, reinterpret_cast<gforb_keypoint*>(dev + in[0].off), dev + in[1].off,
                                                      nullptr, n1, reinterpret_cast<gforb_keypoint*>(dev + in[2].off), dev + in[3].off,
                                                      nullptr, n2, dp, window_size, check_orientation,
                                                      reinterpret_cast<int32_t*>(dev + om), reinterpret_cast<int32_t*>(dev + on), nullptr);
    if (rc) return rc;
    M_CUDA(cudaMemcpyAsync(pin + in[4].off, dp, total - in[4].off, cudaMemcpyDeviceToHost, s));       // prev_xy | matches12 | nmatches
    M_CUDA(cudaStreamSynchronize(s));
    std::memcpy(prev_xy, pin + in[4].off, (size_t)n1 * 8);
    std::memcpy(matches12, pin + om, (size_t)n1 * 4);
    *nmatches = *reinterpret_cast<const int32_t*>(pin + on);
    return GFORB_OK;
}

int gforb_matcher_set_inv_level_sigma2(gforb_matcher* m, const float* inv_level_sigma2, int n_levels) {
    if (!m || !inv_level_sigma2 || n_levels < 1 || n_levels > 32) { set_error("gforb_matcher_set_inv_level_sigma2: bad arguments"); return GFORB_ERR_ARG; }
    for (int l = 0; l < 32; ++l) m->inv_sig2[l] = inv_level_sigma2[std::min(l, n_levels - 1)];
    m->have_inv_sig2 = true;
    return GFORB_OK;
}

int gforb_search_by_projection_batch_device(gforb_matcher* m, const gforb_frame_bounds* bounds, int n_frames,
                                    const gforb_keypoint* keys, const uint8_t* desc, const float* u_right,
                                    const uint8_t* occupied, const int32_t* n, int key_stride,
                                    const gforb_proj_query* queries, const uint8_t* qdesc, const int32_t* nq, int q_stride,
                                    int th_dist, int use_ratio, int check_orientation,
                                    int32_t* key_query, int32_t* key_dist, int32_t* query_key, int32_t* nmatches,
                                    void* cuda_stream) {
    const char* fn = "gforb_search_by_projection";
    if (!m || n_frames < 0 || key_stride < 0 || q_stride < 0 || !nmatches ||
        (key_stride > 0 && (!keys || !desc || !key_query || !key_dist)) || (q_stride > 0 && (!queries || !qdesc || !query_key))) {
        set_error(std::string(fn) + ": bad arguments");
        return GFORB_ERR_ARG;
    }
    if (th_dist < 0 || th_dist > 255) { set_error(std::string(fn) + ": th_dist outside 0..255"); return GFORB_ERR_ARG; }
    int rc = check_bounds(bounds, fn);
    if (rc) return rc;
    if (key_stride > gridmatch_max_keys() || q_stride > gridmatch_max_keys()) {
        set_error(std::string(fn) + ": more than 8192 keys or queries per frame");
        return GFORB_ERR_CAPACITY;
    }
    if (((uintptr_t)desc | (uintptr_t)qdesc) & 15) { set_error(std::string(fn) + ": descriptor arrays must be 16-byte aligned"); return GFORB_ERR_ARG; }
    if (n_frames == 0) return GFORB_OK;
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : m->stream;
    m->last = s;
    const int cells = gridmatch_cells(), ks = std::max(key_stride, 1), qs = std::max(q_stride, 1);
    const size_t bs = al16((size_t)n_frames * (cells + 1) * 4), bi = al16((size_t)n_frames * ks * 4), bp = (size_t)n_frames * qs * 32;
    rc = m->reserve(m->scratch, m->scratch_bytes, bs + bi + bp);
    if (rc) return rc;
    uint8_t* base = static_cast<uint8_t*>(m->scratch);
    int32_t* ds = reinterpret_cast<int32_t*>(base);
    int32_t* di = reinterpret_cast<int32_t*>(base + bs);
    const float bb[4] = {bounds->min_x, bounds->max_x, bounds->min_y, bounds->max_y};
    M_CUDA(cudaEventRecord(m->ev0, s));
    launch_grid_build(keys, n, key_stride, ks, n_frames, bb, ds, di, s);
    ProjSearchArgs a{};
    a.keys = keys; a.desc = desc; a.u_right = u_right; a.occupied = occupied; a.n = n;
    a.q = queries; a.qdesc = qdesc; a.nq = nq;
    a.n_fixed = key_stride; a.key_stride = ks; a.nq_fixed = q_stride; a.q_stride = qs;
    a.cell_start = ds; a.cell_idx = di;
    a.ratio = m->ratio; a.th_dist = th_dist; a.use_ratio = use_ratio ? 1 : 0; a.check_ori = check_orientation ? 1 : 0;
    for (int l = 0; l < 32; ++l) a.inv_sig2[l] = m->have_inv_sig2 ? m->inv_sig2[l] : 1.0f;
    a.spec = reinterpret_cast<uint4*>(base + bs + bi);
    a.key_query = key_query; a.key_dist = key_dist; a.query_key = query_key; a.nmatches = nmatches;
    if (launch_proj_search(a, bb, n_frames, s)) { set_error(std::string(fn) + ": shared-memory opt-in failed"); return GFORB_ERR_CUDA; }
    M_CUDA(cudaEventRecord(m->ev1, s));
    M_CUDA(cudaGetLastError());
    m->timed = true;
    return GFORB_OK;
}

int gforb_search_by_projection(gforb_matcher* m, const gforb_frame_bounds* bounds, const gforb_keypoint* keys,
                               const uint8_t* desc, const float* u_right, const uint8_t* occupied, int n,
                               const gforb_proj_query* queries, const uint8_t* qdesc, int nq, int th_dist, int use_ratio,
                               int check_orientation, int32_t* key_query, int32_t* key_dist, int32_t* query_key,
                               int* nmatches) {
    if (!m || n < 0 || nq < 0 || !nmatches || (n > 0 && (!keys || !desc || !key_query || !key_dist)) ||
        (nq > 0 && (!queries || !qdesc || !query_key))) {
        set_error("gforb_search_by_projection: bad arguments");
        return GFORB_ERR_ARG;
    }
    *nmatches = 0;
    // inputs and outputs are packed into one block each, mirrored in pinned host memory: one H2D, one D2H per call
    struct Seg { const void* src; size_t bytes, off; };
    Seg in[6] = {{keys, (size_t)n * sizeof(gforb_keypoint), 0}, {desc, (size_t)n * 32, 0}, {u_right, u_right ? (size_t)n * 4 : 0, 0},
                 {occupied, occupied ? (size_t)n : 0, 0}, {queries, (size_t)nq * sizeof(gforb_proj_query), 0}, {qdesc, (size_t)nq * 32, 0}};
    size_t total = 0;
    for (Seg& g : in) { g.off = total; total += al16(std::max<size_t>(g.bytes, 16)); }
    const size_t in_bytes = total, okq = total, okd = okq + al16((size_t)std::max(n, 1) * 4), oqk = okd + al16((size_t)std::max(n, 1) * 4),
                 onm = oqk + al16((size_t)std::max(nq, 1) * 4);
    total = onm + 16;
    int rc = m->reserve(m->io, m->io_bytes, total);
    if (rc) return rc;
    rc = m->reserve_pin(total);
    if (rc) return rc;
    uint8_t* dev = static_cast<uint8_t*>(m->io);
    uint8_t* pin = static_cast<uint8_t*>(m->pin);
    for (const Seg& g : in)
        if (g.bytes) std::memcpy(pin + g.off, g.src, g.bytes);
    cudaStream_t s = m->stream;
    M_CUDA(cudaMemcpyAsync(dev, pin, in_bytes, cudaMemcpyHostToDevice, s));
    rc = gforb_search_by_projection_batch_device(
        m, bounds, 1, reinterpret_cast<gforb_keypoint*>(dev + in[0].off), dev + in[1].off,
        u_right ? reinterpret_cast<float*>(dev + in[2].off) : nullptr, occupied ? dev + in[3].off : nullptr, nullptr, n,
        reinterpret_cast<gforb_proj_query*>(dev + in[4].off), dev + in[5].off, nullptr, nq, th_dist, use_ratio, check_orientation,
        reinterpret_cast<int32_t*>(dev + okq), reinterpret_cast<int32_t*>(dev + okd), reinterpret_cast<int32_t*>(dev + oqk),
        reinterpret_cast<int32_t*>(dev + onm), nullptr);
    if (rc) return rc;
    M_CUDA(cudaMemcpyAsync(pin + in_bytes, dev + in_bytes, total - in_bytes, cudaMemcpyDeviceToHost, s));
    M_CUDA(cudaStreamSynchronize(s));
    if (n > 0) { std::memcpy(key_query, pin + okq, (size_t)n * 4); std::memcpy(key_dist, pin + okd, (size_t)n * 4); }
    if (nq > 0) std::memcpy(query_key, pin + oqk, (size_t)nq * 4);
    *nmatches = *reinterpret_cast<const int32_t*>(pin + onm);
    return GFORB_OK;
}

// ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1214-1438): the two windowed searches are the projection entry with plain
// GFORB_PQ_ACTIVE queries (nothing blocks, no gate, no ratio test, first minimum in window order, bestDist <= TH_HIGH),
// the agreement test is one more kernel; nothing returns to the host in between.
int gforb_search_by_sim3_device(gforb_matcher* m, const gforb_frame_bounds* bounds,
                                const gforb_keypoint* keys1, const uint8_t* desc1, int n1,
                                const gforb_keypoint* keys2, const uint8_t* desc2, int n2,
                                const gforb_proj_query* queries12, const uint8_t* qdesc12,
                                const gforb_proj_query* queries21, const uint8_t* qdesc21, int th_dist,
                                int32_t* match1, int32_t* match2, int32_t* match12, int32_t* nfound, void* cuda_stream) {
    const char* fn = "gforb_search_by_sim3";
    if (!m || n1 < 0 || n2 < 0 || !nfound || (n1 > 0 && (!keys1 || !desc1 || !queries12 || !qdesc12 || !match1 || !match12)) ||
        (n2 > 0 && (!keys2 || !desc2 || !queries21 || !qdesc21 || !match2))) {
        set_error(std::string(fn) + ": bad arguments");
        return GFORB_ERR_ARG;
    }
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : m->stream;
    if (n1 == 0 || n2 == 0) {                      // one side has no keys: no window holds a candidate, nothing can agree
        if (n1 > 0) { M_CUDA(cudaMemsetAsync(match1, 0xFF, (size_t)n1 * 4, s)); M_CUDA(cudaMemsetAsync(match12, 0xFF, (size_t)n1 * 4, s)); }
        if (n2 > 0) M_CUDA(cudaMemsetAsync(match2, 0xFF, (size_t)n2 * 4, s));
        M_CUDA(cudaMemsetAsync(nfound, 0, 4, s));
        m->last = s;
        m->timed = false;
        return GFORB_OK;
    }
    // key_query / key_dist / nmatches of the two searches are by-products: they live in the aux block
    const size_t kq = al16((size_t)std::max(std::max(n1, n2), 1) * 4);
    int rc = m->reserve(m->aux, m->aux_bytes, 2 * kq + 16);
    if (rc) return rc;
    uint8_t* aux = static_cast<uint8_t*>(m->aux);
    int32_t* key_query = reinterpret_cast<int32_t*>(aux);
    int32_t* key_dist = reinterpret_cast<int32_t*>(aux + kq);
    int32_t* nm = reinterpret_cast<int32_t*>(aux + 2 * kq);
    // KF1's map points into KF2 (:1259-1337), then KF2's into KF1 (:1339-1417)
    rc = gforb_search_by_projection_batch_device(m, bounds, 1, keys2, desc2, nullptr, nullptr, nullptr, n2, queries12, qdesc12,
                                                 nullptr, n1, th_dist, 0, 0, key_query, key_dist, match1, nm, s);
    if (rc) return rc;
    cudaEvent_t first = m->ev0;                    // time the whole entry: keep the first call's start event
    m->ev0 = m->ev_spare;
    rc = gforb_search_by_projection_batch_device(m, bounds, 1, keys1, desc1, nullptr, nullptr, nullptr, n1, queries21, qdesc21,
                                                 nullptr, n2, th_dist, 0, 0, key_query, key_dist, match2, nm + 1, s);
    m->ev_spare = m->ev0;
    m->ev0 = first;
    if (rc) return rc;
    launch_sim3_agree(match1, n1, match2, n2, match12, nfound, s);
    M_CUDA(cudaEventRecord(m->ev1, s));
    M_CUDA(cudaGetLastError());
    m->last = s;
    m->timed = true;
    return GFORB_OK;
}

int gforb_search_by_sim3(gforb_matcher* m, const gforb_frame_bounds* bounds,
                         const gforb_keypoint* keys1, const uint8_t* desc1, int n1,
                         const gforb_keypoint* keys2, const uint8_t* desc2, int n2,
                         const gforb_proj_query* queries12, const uint8_t* qdesc12,
                         const gforb_proj_query* queries21, const uint8_t* qdesc21, int th_dist,
                         int32_t* match1, int32_t* match2, int32_t* match12, int* nfound) {
    if (!m || n1 < 0 || n2 < 0 || !nfound || (n1 > 0 && (!keys1 || !desc1 || !queries12 || !qdesc12 || !match12)) ||
        (n2 > 0 && (!keys2 || !desc2 || !queries21 || !qdesc21))) {
        set_error("gforb_search_by_sim3: bad arguments");
        return GFORB_ERR_ARG;
    }
    *nfound = 0;
    struct Seg { const void* src; size_t bytes, off; };
    Seg in[8] = {{keys1, (size_t)n1 * sizeof(gforb_keypoint), 0}, {desc1, (size_t)n1 * 32, 0},
                 {keys2, (size_t)n2 * sizeof(gforb_keypoint), 0}, {desc2, (size_t)n2 * 32, 0},
                 {queries12, (size_t)n1 * sizeof(gforb_proj_query), 0}, {qdesc12, (size_t)n1 * 32, 0},
                 {queries21, (size_t)n2 * sizeof(gforb_proj_query), 0}, {qdesc21, (size_t)n2 * 32, 0}};
    size_t total = 0;
    for (Seg& g : in) { g.off = total; total += al16(std::max<size_t>(g.bytes, 16)); }
    const size_t in_bytes = total, o1 = total, o2 = o1 + al16((size_t)std::max(n1, 1) * 4), o12 = o2 + al16((size_t)std::max(n2, 1) * 4),
                 onf = o12 + al16((size_t)std::max(n1, 1) * 4);
    total = onf + 16;
    int rc = m->reserve(m->io, m->io_bytes, total);
    if (rc) return rc;
    rc = m->reserve_pin(total);
    if (rc) return rc;
    uint8_t* dev = static_cast<uint8_t*>(m->io);
    uint8_t* pin = static_cast<uint8_t*>(m->pin);
    for (const Seg& g : in)
        if (g.bytes) std::memcpy(pin + g.off, g.src, g.bytes);
    cudaStream_t s = m->stream;
    M_CUDA(cudaMemcpyAsync(dev, pin, in_bytes, cudaMemcpyHostToDevice, s));
    auto K = [&](int i) { return reinterpret_cast<gforb_keypoint*>(dev + in[i].off); };
    auto Q = [&](int i) { return reinterpret_cast<gforb_proj_query*>(dev + in[i].off); };
    rc = gforb_search_by_sim3_device(m, bounds, K(0), dev + in[1].off, n1, K(2), dev + in[3].off, n2, Q(4), dev + in[5].off, Q(6),
                                     dev + in[7].off, th_dist, reinterpret_cast<int32_t*>(dev + o1), reinterpret_cast<int32_t*>(dev + o2),
                                     reinterpret_cast<int32_t*>(dev + o12), reinterpret_cast<int32_t*>(dev + onf), nullptr);
    if (rc) return rc;
    M_CUDA(cudaMemcpyAsync(pin + in_bytes, dev + in_bytes, total - in_bytes, cudaMemcpyDeviceToHost, s));
    M_CUDA(cudaStreamSynchronize(s));
    if (n1 > 0) {
        if (match1) std::memcpy(match1, pin + o1, (size_t)n1 * 4);
        std::memcpy(match12, pin + o12, (size_t)n1 * 4);
    }
    if (n2 > 0 && match2) std::memcpy(match2, pin + o2, (size_t)n2 * 4);
    *nfound = *reinterpret_cast<const int32_t*>(pin + onf);
    return GFORB_OK;
}

namespace {
// SearchByBoW (ep == nullptr) and SearchForTriangulation (ep != nullptr) share the node walk; all pointers device
int bow_device_impl(const char* fn, gforb_matcher* m, const gforb_keypoint* k1, const uint8_t* d1, const uint8_t* valid1,
                    const uint8_t* stereo1, int n1, const int32_t* node_id1, const int32_t* node_start1, const int32_t* node_idx1,
                    int nn1, const gforb_keypoint* k2, const uint8_t* d2, const uint8_t* valid2, const uint8_t* stereo2, int n2,
                    const int32_t* node_id2, const int32_t* node_start2, const int32_t* node_idx2, int nn2,
                    const gforb_epipolar* ep, int th_dist, int check_orientation, int32_t* matches12, int32_t* nmatches,
                    void* cuda_stream) {
    if (!m || n1 < 0 || n2 < 0 || nn1 < 0 || nn2 < 0 || !nmatches || (n1 > 0 && (!k1 || !d1 || !matches12)) || (n2 > 0 && (!k2 || !d2)) ||
        (nn1 > 0 && (!node_id1 || !node_start1 || !node_idx1)) || (nn2 > 0 && (!node_id2 || !node_start2 || !node_idx2))) {
        set_error(std::string(fn) + ": bad arguments");
        return GFORB_ERR_ARG;
    }
    if (th_dist < 0 || th_dist > 255) { set_error(std::string(fn) + ": th_dist outside 0..255"); return GFORB_ERR_ARG; }
    if (ep && (ep->n_levels < 1 || ep->n_levels > 32)) { set_error(std::string(fn) + ": n_levels outside 1..32"); return GFORB_ERR_ARG; }
    if (((uintptr_t)d1 | (uintptr_t)d2) & 15) { set_error(std::string(fn) + ": descriptor arrays must be 16-byte aligned"); return GFORB_ERR_ARG; }
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : m->stream;
    m->last = s;
    int rc = m->reserve(m->scratch, m->scratch_bytes, al16((size_t)std::max(n2, 1)));
    if (rc) return rc;
    BowSearchArgs a{};
    a.k1 = k1; a.d1 = d1; a.valid1 = valid1; a.n1 = n1; a.node_id1 = node_id1; a.node_start1 = node_start1; a.node_idx1 = node_idx1; a.nn1 = nn1;
    a.k2 = k2; a.d2 = d2; a.valid2 = valid2; a.n2 = n2; a.node_id2 = node_id2; a.node_start2 = node_start2; a.node_idx2 = node_idx2; a.nn2 = nn2;
    a.th_dist = th_dist; a.ratio = m->ratio; a.check_ori = check_orientation ? 1 : 0;
    if (ep) {
        a.tri = 1; a.stereo1 = stereo1; a.stereo2 = stereo2;
        std::memcpy(a.F12, ep->F12, sizeof(a.F12));
        a.ex = ep->ex; a.ey = ep->ey;
        for (int l = 0; l < 32; ++l) {                       // levels past n_levels repeat the last one (never indexed by valid keys)
            const int c = std::min(l, ep->n_levels - 1);
            a.sf[l] = ep->scale_factors[c]; a.sig2[l] = ep->level_sigma2[c];
        }
    }
    a.matched2 = static_cast<uint8_t*>(m->scratch); a.matches12 = matches12; a.nmatches = nmatches;
    M_CUDA(cudaEventRecord(m->ev0, s));
    launch_bow_search(a, s);
    M_CUDA(cudaEventRecord(m->ev1, s));
    M_CUDA(cudaGetLastError());
    m->timed = true;
    return GFORB_OK;
}

// host-pointer form: inputs packed into one pinned block (one H2D), results into another (one D2H)
int bow_host_impl(const char* fn, gforb_matcher* m, const gforb_keypoint* k1, const uint8_t* d1, const uint8_t* valid1,
                  const uint8_t* stereo1, int n1, const int32_t* node_id1, const int32_t* node_start1, const int32_t* node_idx1, int nn1,
                  const gforb_keypoint* k2, const uint8_t* d2, const uint8_t* valid2, const uint8_t* stereo2, int n2,
                  const int32_t* node_id2, const int32_t* node_start2, const int32_t* node_idx2, int nn2, const gforb_epipolar* ep,
                  int th_dist, int check_orientation, int32_t* matches12, int* nmatches) {
    if (!m || n1 < 0 || n2 < 0 || nn1 < 0 || nn2 < 0 || !nmatches || (n1 > 0 && (!k1 || !d1 || !matches12)) || (n2 > 0 && (!k2 || !d2)) ||
        (nn1 > 0 && (!node_id1 || !node_start1 || !node_idx1)) || (nn2 > 0 && (!node_id2 || !node_start2 || !node_idx2))) {
        set_error(std::string(fn) + ": bad arguments");
        return GFORB_ERR_ARG;
    }
    *nmatches = 0;
    const int e1 = nn1 > 0 ? node_start1[nn1] : 0, e2 = nn2 > 0 ? node_start2[nn2] : 0;      // feature-vector entries
    for (int i = 0; i < e1; ++i) if (node_idx1[i] < 0 || node_idx1[i] >= n1) { set_error(std::string(fn) + ": feature index out of range"); return GFORB_ERR_ARG; }
    for (int i = 0; i < e2; ++i) if (node_idx2[i] < 0 || node_idx2[i] >= n2) { set_error(std::string(fn) + ": feature index out of range"); return GFORB_ERR_ARG; }
    struct Seg { const void* src; size_t bytes, off; };
    Seg in[14] = {{k1, (size_t)n1 * sizeof(gforb_keypoint), 0}, {d1, (size_t)n1 * 32, 0}, {valid1, valid1 ? (size_t)n1 : 0, 0},
                  {node_id1, (size_t)nn1 * 4, 0}, {node_start1, nn1 > 0 ? (size_t)(nn1 + 1) * 4 : 0, 0}, {node_idx1, (size_t)e1 * 4, 0},
                  {k2, (size_t)n2 * sizeof(gforb_keypoint), 0}, {d2, (size_t)n2 * 32, 0}, {valid2, valid2 ? (size_t)n2 : 0, 0},
                  {node_id2, (size_t)nn2 * 4, 0}, {node_start2, nn2 > 0 ? (size_t)(nn2 + 1) * 4 : 0, 0}, {node_idx2, (size_t)e2 * 4, 0},
                  {stereo1, stereo1 ? (size_t)n1 : 0, 0}, {stereo2, stereo2 ? (size_t)n2 : 0, 0}};
    size_t total = 0;
    for (Seg& g : in) { g.off = total; total += al16(std::max<size_t>(g.bytes, 16)); }
    const size_t in_bytes = total, om = total, on = om + al16((size_t)std::max(n1, 1) * 4);
    total = on + 16;
    int rc = m->reserve(m->io, m->io_bytes, total);
    if (rc) return rc;
    rc = m->reserve_pin(total);
    if (rc) return rc;
    uint8_t* dev = static_cast<uint8_t*>(m->io);
    uint8_t* pin = static_cast<uint8_t*>(m->pin);
    for (const Seg& g : in)
        if (g.bytes) std::memcpy(pin + g.off, g.src, g.bytes);
    cudaStream_t s = m->stream;
    M_CUDA(cudaMemcpyAsync(dev, pin, in_bytes, cudaMemcpyHostToDevice, s));
    auto i32 = [&](int i) { return reinterpret_cast<int32_t*>(dev + in[i].off); };
    rc = bow_device_impl(fn, m, reinterpret_cast<gforb_keypoint*>(dev + in[0].off), dev + in[1].off, valid1 ? dev + in[2].off : nullptr,
                         stereo1 ? dev + in[12].off : nullptr, n1, i32(3), i32(4), i32(5), nn1,
                         reinterpret_cast<gforb_keypoint*>(dev + in[6].off), dev + in[7].off, valid2 ? dev + in[8].off : nullptr,
                         stereo2 ? dev + in[13].off : nullptr, n2, i32(9), i32(10), i32(11), nn2, ep, th_dist, check_orientation,
                         reinterpret_cast<int32_t*>(dev + om), reinterpret_cast<int32_t*>(dev + on), nullptr);
    if (rc) return rc;
    M_CUDA(cudaMemcpyAsync(pin + in_bytes, dev + in_bytes, total - in_bytes, cudaMemcpyDeviceToHost, s));
    M_CUDA(cudaStreamSynchronize(s));
    if (n1 > 0) std::memcpy(matches12, pin + om, (size_t)n1 * 4);
    *nmatches = *reinterpret_cast<const int32_t*>(pin + on);
    return GFORB_OK;
}
}  // namespace

int gforb_search_by_bow_device(gforb_matcher* m, const gforb_keypoint* k1, const uint8_t* d1, const uint8_t* valid1, int n1,
                               const int32_t* node_id1, const int32_t* node_start1, const int32_t* node_idx1, int nn1,
                               const gforb_keypoint* k2, const uint8_t* d2, const uint8_t* valid2, int n2,
                               const int32_t* node_id2, const int32_t* node_start2, const int32_t* node_idx2, int nn2,
                               int th_dist, int check_orientation, int32_t* matches12, int32_t* nmatches, void* cuda_stream) {
    return bow_device_impl("gforb_search_by_bow", m, k1, d1, valid1, nullptr, n1, node_id1, node_start1, node_idx1, nn1, k2, d2, valid2,
                           nullptr, n2, node_id2, node_start2, node_idx2, nn2, nullptr, th_dist, check_orientation, matches12, nmatches,
                           cuda_stream);
}

int gforb_search_by_bow(gforb_matcher* m, const gforb_keypoint* k1, const uint8_t* d1, const uint8_t* valid1, int n1,
                        const int32_t* node_id1, const int32_t* node_start1, const int32_t* node_idx1, int nn1,
                        const gforb_keypoint* k2, const uint8_t* d2, const uint8_t* valid2, int n2,
                        const int32_t* node_id2, const int32_t* node_start2, const int32_t* node_idx2, int nn2,
                        int th_dist, int check_orientation, int32_t* matches12, int* nmatches) {
    return bow_host_impl("gforb_search_by_bow", m, k1, d1, valid1, nullptr, n1, node_id1, node_start1, node_idx1, nn1, k2, d2, valid2,
                         nullptr, n2, node_id2, node_start2, node_idx2, nn2, nullptr, th_dist, check_orientation, matches12, nmatches);
}

int gforb_search_for_triangulation_device(gforb_matcher* m, const gforb_keypoint* k1, const uint8_t* d1, const uint8_t* valid1,
                               const uint8_t* stereo1, int n1, const int32_t* node_id1, const int32_t* node_start1,
                               const int32_t* node_idx1, int nn1, const gforb_keypoint* k2, const uint8_t* d2, const uint8_t* valid2,
                               const uint8_t* stereo2, int n2, const int32_t* node_id2, const int32_t* node_start2,
                               const int32_t* node_idx2, int nn2, const gforb_epipolar* ep, int th_low, int check_orientation,
                               int32_t* matches12, int32_t* nmatches, void* cuda_stream) {
    if (!ep) { set_error("gforb_search_for_triangulation: null epipolar geometry"); return GFORB_ERR_ARG; }
    return bow_device_impl("gforb_search_for_triangulation", m, k1, d1, valid1, stereo1, n1, node_id1, node_start1, node_idx1, nn1, k2, d2,
                           valid2, stereo2, n2, node_id2, node_start2, node_idx2, nn2, ep, th_low, check_orientation, matches12, nmatches,
                           cuda_stream);
}

int gforb_search_for_triangulation(gforb_matcher* m, const gforb_keypoint* k1, const uint8_t* d1, const uint8_t* valid1,
                               const uint8_t* stereo1, int n1, const int32_t* node_id1, const int32_t* node_start1,
                               const int32_t* node_idx1, int nn1, const gforb_keypoint* k2, const uint8_t* d2, const uint8_t* valid2,
                               const uint8_t* stereo2, int n2, const int32_t* node_id2, const int32_t* node_start2,
                               const int32_t* node_idx2, int nn2, const gforb_epipolar* ep, int th_low, int check_orientation,
                               int32_t* matches12, int* nmatches) {
    if (!ep) { set_error("gforb_search_for_triangulation: null epipolar geometry"); return GFORB_ERR_ARG; }
    return bow_host_impl("gforb_search_for_triangulation", m, k1, d1, valid1, stereo1, n1, node_id1, node_start1, node_idx1, nn1, k2, d2,
                         valid2, stereo2, n2, node_id2, node_start2, node_idx2, nn2, ep, th_low, check_orientation, matches12, nmatches);
}

// ------------------------------------------------------------------------------------------- pre-processing
struct gforb_rectifier {
    int w = 0, h = 0;
    float* mapx = nullptr; float* mapy = nullptr;
    cudaStream_t stream = nullptr, last = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timed = false;
    void* io = nullptr; size_t io_bytes = 0;
    ~gforb_rectifier() {
        if (mapx) cudaFree(mapx);
        if (mapy) cudaFree(mapy);
        if (io) cudaFree(io);
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
        if (stream) cudaStreamDestroy(stream);
    }
};

int gforb_rectifier_create(const float* map_x, const float* map_y, int w, int h, gforb_rectifier** out) {
    if (!out || (map_x == nullptr) != (map_y == nullptr) || (map_x && (w <= 0 || h <= 0))) { set_error("gforb_rectifier_create: bad arguments"); return GFORB_ERR_ARG; }
    std::unique_ptr<gforb_rectifier> r(new (std::nothrow) gforb_rectifier);
    if (!r) return GFORB_ERR_ARG;
    int rc = ensure_stream(r->stream);
    if (rc) return rc;
    M_CUDA(cudaEventCreate(&r->ev0));
    M_CUDA(cudaEventCreate(&r->ev1));
    if (map_x) {
        r->w = w; r->h = h;
        const size_t bytes = (size_t)w * h * sizeof(float);
        M_CUDA(cudaMalloc(&r->mapx, bytes));
        M_CUDA(cudaMalloc(&r->mapy, bytes));
        M_CUDA(cudaMemcpy(r->mapx, map_x, bytes, cudaMemcpyHostToDevice));
        M_CUDA(cudaMemcpy(r->mapy, map_y, bytes, cudaMemcpyHostToDevice));
    }
    *out = r.release();
    return GFORB_OK;
}
void gforb_rectifier_destroy(gforb_rectifier* r) { delete r; }

int gforb_preprocess_batch_device(gforb_rectifier* r, const void* src, int sw, int sh, size_t spitch, size_t sstride, int fmt,
                                  int n, uint8_t* dst, size_t dpitch, size_t dstride, void* cuda_stream) {
    const int bpp = preprocess_bytes_per_pixel(fmt);
    if (!r || !src || !dst || sw <= 0 || sh <= 0 || n < 0 || n > 65535 || bpp == 0 || spitch < (size_t)sw * bpp) { set_error("gforb_preprocess: bad arguments"); return GFORB_ERR_ARG; }
    const int dw = r->mapx ? r->w : sw, dh = r->mapx ? r->h : sh;
    if (dpitch < (size_t)dw) { set_error("gforb_preprocess: dst_pitch < output width"); return GFORB_ERR_ARG; }
    if ((unsigned long long)spitch * (unsigned long long)sh > 0xFFFFFFFFull) { set_error("gforb_preprocess: source image larger than 4 GB"); return GFORB_ERR_ARG; }
    if (fmt == GFORB_PIX_GRAY16 && (((uintptr_t)src | spitch | sstride) & 1)) { set_error("gforb_preprocess: 16-bit source must be 2-byte aligned"); return GFORB_ERR_ARG; }
    if (n == 0) return GFORB_OK;
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : r->stream;
    r->last = s;
    M_CUDA(cudaEventRecord(r->ev0, s));
    launch_preprocess(static_cast<const uint8_t*>(src), sw, sh, spitch, sstride, fmt, r->mapx, r->mapy, dst, dw, dh, dpitch, dstride, n, s);
    M_CUDA(cudaEventRecord(r->ev1, s));
    M_CUDA(cudaGetLastError());
    r->timed = true;
    return GFORB_OK;
}

int gforb_rectifier_sync(gforb_rectifier* r, float* last_ms) {
    if (!r) return GFORB_ERR_ARG;
    M_CUDA(cudaStreamSynchronize(r->last ? r->last : r->stream));
    if (last_ms) {
        *last_ms = 0.f;
        if (r->timed) M_CUDA(cudaEventElapsedTime(last_ms, r->ev0, r->ev1));
    }
    return GFORB_OK;
}

int gforb_preprocess(gforb_rectifier* r, const void* src, int sw, int sh, size_t spitch, int fmt, uint8_t* dst, size_t dpitch) {
    const int bpp = preprocess_bytes_per_pixel(fmt);
    if (!r || !src || !dst || sw <= 0 || sh <= 0 || bpp == 0 || spitch < (size_t)sw * bpp) { set_error("gforb_preprocess: bad arguments"); return GFORB_ERR_ARG; }
    const int dw = r->mapx ? r->w : sw, dh = r->mapx ? r->h : sh;
    if (dpitch < (size_t)dw) { set_error("gforb_preprocess: dst_pitch < output width"); return GFORB_ERR_ARG; }
    const size_t srow = ((size_t)sw * bpp + 15) & ~(size_t)15, drow = ((size_t)dw + 15) & ~(size_t)15;
    const size_t need = srow * sh + drow * dh;
    if (need > r->io_bytes) {
        if (r->io) { cudaFree(r->io); r->io = nullptr; r->io_bytes = 0; }
        M_CUDA(cudaMalloc(&r->io, need));
        r->io_bytes = need;
    }
    uint8_t* dsrc = static_cast<uint8_t*>(r->io);
    uint8_t* ddst = dsrc + srow * sh;
    cudaStream_t s = r->stream;
    M_CUDA(cudaMemcpy2DAsync(dsrc, srow, src, spitch, (size_t)sw * bpp, sh, cudaMemcpyHostToDevice, s));
    int rc = gforb_preprocess_batch_device(r, dsrc, sw, sh, srow, 0, fmt, 1, ddst, drow, 0, nullptr);
    if (rc) return rc;
    M_CUDA(cudaMemcpy2DAsync(dst, dpitch, ddst, drow, (size_t)dw, dh, cudaMemcpyDeviceToHost, s));
    M_CUDA(cudaStreamSynchronize(s));
    return GFORB_OK;
}

// --------------------------------------------------------------------------------- undistortion + image bounds
namespace {
int check_camera(const gforb_camera* c, const char* who) {
    if (!c || !(c->K[0] != 0.0) || !(c->K[4] != 0.0)) { set_error(std::string(who) + ": bad camera"); return GFORB_ERR_ARG; }
    return GFORB_OK;
}
}  // namespace

int gforb_undistort_keypoints_batch_device(gforb_matcher* m, const gforb_camera* cam, const gforb_keypoint* keys,
                                           const int32_t* counts, int stride, int n_images, gforb_keypoint* out, void* cuda_stream) {
    if (!m || stride < 0 || n_images < 0 || n_images > 65535 || (stride > 0 && n_images > 0 && (!keys || !out))) { set_error("gforb_undistort_keypoints: bad arguments"); return GFORB_ERR_ARG; }
    int rc = check_camera(cam, "gforb_undistort_keypoints");
    if (rc) return rc;
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : m->stream;
    m->last = s;
    M_CUDA(cudaEventRecord(m->ev0, s));
    launch_undistort_keys(cam->K, cam->dist, cam->R, cam->P, keys, counts, stride, stride, n_images, out, s);
    M_CUDA(cudaEventRecord(m->ev1, s));
    M_CUDA(cudaGetLastError());
    m->timed = true;
    return GFORB_OK;
}

int gforb_undistort_keypoints(gforb_matcher* m, const gforb_camera* cam, const gforb_keypoint* keys, int n, gforb_keypoint* out) {
    if (!m || n < 0 || (n > 0 && (!keys || !out))) { set_error("gforb_undistort_keypoints: bad arguments"); return GFORB_ERR_ARG; }
    int rc = check_camera(cam, "gforb_undistort_keypoints");
    if (rc || n == 0) return rc;
    const size_t bytes = (size_t)n * sizeof(gforb_keypoint);
    rc = m->reserve(m->io, m->io_bytes, bytes);
    if (rc) return rc;
    gforb_keypoint* d = static_cast<gforb_keypoint*>(m->io);
    cudaStream_t s = m->stream;
    M_CUDA(cudaMemcpyAsync(d, keys, bytes, cudaMemcpyHostToDevice, s));
    rc = gforb_undistort_keypoints_batch_device(m, cam, d, nullptr, n, 1, d, nullptr);
    if (rc) return rc;
    M_CUDA(cudaMemcpyAsync(out, d, bytes, cudaMemcpyDeviceToHost, s));
    M_CUDA(cudaStreamSynchronize(s));
    return GFORB_OK;
}

int gforb_compute_image_bounds(gforb_matcher* m, const gforb_camera* cam, int cols, int rows, gforb_frame_bounds* out) {
    if (!m || !out || cols <= 0 || rows <= 0) { set_error("gforb_compute_image_bounds: bad arguments"); return GFORB_ERR_ARG; }
    int rc = check_camera(cam, "gforb_compute_image_bounds");
    if (rc) return rc;
    if ((float)cam->dist[0] == 0.0f) {                      // src/Frame.cc:822-828
        out->min_x = 0.0f; out->max_x = (float)cols; out->min_y = 0.0f; out->max_y = (float)rows;
        return GFORB_OK;
    }
    const float c = (float)cols, r = (float)rows, hc = (float)((float)cols / 2.0), hr = (float)((float)rows / 2.0);
    float pts[16] = {0.f, 0.f, c, 0.f, 0.f, r, c, r, 0.f, hr, hc, 0.f, hc, r, c, hr};     // src/Frame.cc:770-787
    rc = m->reserve(m->io, m->io_bytes, sizeof(pts));
    if (rc) return rc;
    float* d = static_cast<float*>(m->io);
    cudaStream_t s = m->stream;
    M_CUDA(cudaMemcpyAsync(d, pts, sizeof(pts), cudaMemcpyHostToDevice, s));
    launch_undistort_xy(cam->K, cam->dist, cam->R, cam->P, d, 8, d, s);
    M_CUDA(cudaGetLastError());
    M_CUDA(cudaMemcpyAsync(pts, d, sizeof(pts), cudaMemcpyDeviceToHost, s));
    M_CUDA(cudaStreamSynchronize(s));
    // src/Frame.cc:803-818: float members initialised from INT_MAX / INT_MIN, floor / ceil applied on update
    float mnx = (float)INT_MAX, mxx = (float)INT_MIN, mny = (float)INT_MAX, mxy = (float)INT_MIN;
    for (int i = 0; i < 8; ++i) {
        if (mnx > pts[2 * i]) mnx = floorf(pts[2 * i]);
        if (mny > pts[2 * i + 1]) mny = floorf(pts[2 * i + 1]);
        if (mxx < pts[2 * i]) mxx = ceilf(pts[2 * i]);
        if (mxy < pts[2 * i + 1]) mxy = ceilf(pts[2 * i + 1]);
    }
    out->min_x = mnx; out->max_x = mxx; out->min_y = mny; out->max_y = mxy;
    return GFORB_OK;
}

int gforb_distinctive_descriptors(gforb_matcher* m, const uint8_t* desc, const int32_t* set_start, int n, int32_t* best_idx,
                                  int32_t* best_median) {
    if (!m || n < 0 || !set_start || (n > 0 && (!best_idx || !best_median))) { set_error("gforb_distinctive_descriptors: bad arguments"); return GFORB_ERR_ARG; }
    if (n == 0) return GFORB_OK;
    int max_n = 0;
    for (int p = 0; p < n; ++p) {
        if (set_start[p + 1] < set_start[p] || set_start[p] < 0) { set_error("gforb_distinctive_descriptors: set_start must be non-decreasing"); return GFORB_ERR_ARG; }
        max_n = std::max(max_n, set_start[p + 1] - set_start[p]);
    }
    if (max_n > distinctive_max_set()) { set_error("gforb_distinctive_descriptors: more than 1024 descriptors in one set"); return GFORB_ERR_CAPACITY; }
    const size_t total = (size_t)set_start[n];
    if (total > 0 && !desc) { set_error("gforb_distinctive_descriptors: null descriptors"); return GFORB_ERR_ARG; }
    const size_t bd = al16(std::max<size_t>(total, 1) * 32), bs = al16((size_t)(n + 1) * 4), bo = al16((size_t)n * 4);
    int rc = m->reserve(m->io, m->io_bytes, bd + bs + 2 * bo);
    if (rc) return rc;
    uint8_t* base = static_cast<uint8_t*>(m->io);
    int32_t* ds = reinterpret_cast<int32_t*>(base + bd);
    int32_t* di = reinterpret_cast<int32_t*>(base + bd + bs);
    int32_t* dm = reinterpret_cast<int32_t*>(base + bd + bs + bo);
    cudaStream_t s = m->stream;
    m->last = s;
    if (total) M_CUDA(cudaMemcpyAsync(base, desc, total * 32, cudaMemcpyHostToDevice, s));
    M_CUDA(cudaMemcpyAsync(ds, set_start, (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, s));
    M_CUDA(cudaEventRecord(m->ev0, s));
    if (launch_distinctive(base, ds, n, max_n, di, dm, s)) { set_error("gforb_distinctive_descriptors: shared-memory opt-in failed"); return GFORB_ERR_CUDA; }
    M_CUDA(cudaEventRecord(m->ev1, s));
    M_CUDA(cudaGetLastError());
    m->timed = true;
    M_CUDA(cudaMemcpyAsync(best_idx, di, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    M_CUDA(cudaMemcpyAsync(best_median, dm, (size_t)n * 4, cudaMemcpyDeviceToHost, s));
    M_CUDA(cudaStreamSynchronize(s));
    return GFORB_OK;
}

int gforb_device_alloc(void** p, size_t bytes) {
    if (!p) return GFORB_ERR_ARG;
    M_CUDA(cudaMalloc(p, std::max<size_t>(bytes, 16)));
    return GFORB_OK;
}
void gforb_device_free(void* p) { if (p) cudaFree(p); }
int gforb_device_copy(void* dst, const void* src, size_t bytes, int dir) {
    if ((!dst || !src) && bytes) return GFORB_ERR_ARG;
    M_CUDA(cudaMemcpy(dst, src, bytes, dir == 0 ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost));
    return GFORB_OK;
}

// ------------------------------------------------------------------------------------------------- batch
int gforb_batch_create(const gforb_params* p, const gforb_limits* lim, int w, int h, int max_frames, gforb_batch** out) {
    if (!p || !out || w <= 0 || h <= 0 || max_frames < 1 || max_frames > 32767) {
        set_error("gforb_batch_create: bad arguments");
        return GFORB_ERR_ARG;
    }
    std::unique_ptr<gforb_batch> b(new (std::nothrow) gforb_batch);
    if (!b) return GFORB_ERR_ARG;
    int rc = make_tables(*p, b->tab);
    if (rc) return rc;
    if (lim) b->lim = *lim;
    rc = ensure_stream(b->own);
    if (rc) return rc;
    b->stream = b->own;
    b->max_frames = max_frames;
    if (const char* ks = std::getenv("GFORB_KSTREAMS")) b->kstreams = std::min(std::max(std::atoi(ks), 1), 4);   // tuning knob
    b->eng.full_frames = false;      // no entry point of the batch handle exposes the padded pyramid levels
    rc = b->eng.init(b->tab, b->lim, w, h, 2 * max_frames, max_frames);
    if (rc) return rc;
    *out = b.release();
    return GFORB_OK;
}
void gforb_batch_destroy(gforb_batch* b) { delete b; }

int gforb_batch_set_stream(gforb_batch* b, void* s) {
    if (!b) return GFORB_ERR_ARG;
    b->stream = s ? (cudaStream_t)s : b->own;
    return GFORB_OK;
}
int gforb_batch_capacity(const gforb_batch* b, int* max_frames, int* K) {
    if (!b) return GFORB_ERR_ARG;
    if (max_frames) *max_frames = b->max_frames;
    if (K) *K = b->eng.g.K;
    return GFORB_OK;
}
int gforb_batch_device_input(gforb_batch* b, uint8_t** dev_ptr, size_t* image_stride, size_t* pitch) {
    if (!b) return GFORB_ERR_ARG;
    if (dev_ptr) *dev_ptr = b->eng.b.in;
    if (image_stride) *image_stride = (size_t)b->eng.g.W * b->eng.g.H;
    if (pitch) *pitch = (size_t)b->eng.g.W;
    return GFORB_OK;
}
int gforb_batch_set_device_input(gforb_batch* b, const uint8_t* dev_images_or_null) {
    if (!b) return GFORB_ERR_ARG;
    b->eng.b.in_ext = dev_images_or_null;
    return GFORB_OK;
}
int gforb_batch_upload(gforb_batch* b, const uint8_t* left, const uint8_t* right, int n) {
    if (!b || !left || !right || n < 1 || n > b->max_frames) { set_error("gforb_batch_upload: bad arguments"); return GFORB_ERR_ARG; }
    if (b->eng.b.in_ext) { set_error("gforb_batch_upload: a caller-owned device input is set (gforb_batch_set_device_input)"); return GFORB_ERR_ARG; }
    int rc = b->eng.upload_images(left, (size_t)b->eng.g.W, 0, 2, n, b->stream);
    if (rc) return rc;
    return b->eng.upload_images(right, (size_t)b->eng.g.W, 1, 2, n, b->stream);
}

namespace {
// kernels for frames [f0, f0 + n) on stream s
int enqueue_frames(gforb_batch* b, int f0, int n, const gforb_stereo_params* prm, cudaStream_t s, bool timing, bool fork_blur = false) {
    Engine& E = b->eng;
    const int n_cap = b->max_frames;
    int rc = E.run_extract(2 * n, s, timing ? b->ev : nullptr, 2 * f0, fork_blur);
    if (rc) return rc;
    b->last_launches += extract_launches(E.g);
    if (timing && !prm) cudaEventRecord(b->ev[6], s);
    if (prm) {
        const size_t K = (size_t)E.g.K;
        StereoView v{};
        v.kL = E.b.kps; v.kR = E.b.kps + K; v.dL = E.b.desc; v.dR = E.b.desc + K * 32;
        v.nL = E.b.nkp; v.nR = E.b.nkp + 1; v.k_stride = 2 * K; v.n_stride = 2;
        fill_view_levels(v, E, E, 2, 1);
        v.f0 = f0;
        v.min_d = nullptr; v.max_d = nullptr;
        v.uright = E.b.uright; v.depth = E.b.depth; v.best_dist = E.b.st_best_dist; v.best_idx = E.b.st_best_idx;
        v.sad = E.b.st_sad; v.nmatch = E.b.st_nmatch; v.out_stride = K;
        v.cut_dist = E.b.st_cut; v.ticket = E.b.st_ticket; v.proc_count = E.b.st_ticket + n_cap;
        GFORB_RANGE_PUSH("gforb ComputeStereoMatches", 5);
        launch_stereo(v, norm_params(*prm), n, (int)K, s);
        GFORB_RANGE_POP();
        b->last_launches += 1;
        if (timing) cudaEventRecord(b->ev[6], s);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return cuda_fail(e, "stereo launch");
    }
    return GFORB_OK;
}
}  // namespace

namespace {
int ensure_pipeline(gforb_batch* b, int nchunks) {
    cudaError_t e;
#define GB(call, what) do { e = (call); if (e != cudaSuccess) return cuda_fail(e, what); } while (0)
    if (!b->s_up) GB(cudaStreamCreateWithFlags(&b->s_up, cudaStreamNonBlocking), "cudaStreamCreate(upload)");
    if (!b->s_down) GB(cudaStreamCreateWithFlags(&b->s_down, cudaStreamNonBlocking), "cudaStreamCreate(download)");
    for (auto& sa : b->s_alt)
        if (!sa) GB(cudaStreamCreateWithFlags(&sa, cudaStreamNonBlocking), "cudaStreamCreate(kernels)");
    if (!b->ev_start) GB(cudaEventCreateWithFlags(&b->ev_start, cudaEventDisableTiming), "cudaEventCreate");
    if (!b->ev_down) GB(cudaEventCreateWithFlags(&b->ev_down, cudaEventDisableTiming), "cudaEventCreate");
    while ((int)b->ev_up.size() < nchunks) {
        cudaEvent_t a = nullptr, d = nullptr;
        GB(cudaEventCreateWithFlags(&a, cudaEventDisableTiming), "cudaEventCreate");
        b->ev_up.push_back(a);
        GB(cudaEventCreateWithFlags(&d, cudaEventDisableTiming), "cudaEventCreate");
        b->ev_done.push_back(d);
    }
#undef GB
    return GFORB_OK;
}
}  // namespace

// Device-resident inputs: every stage is ONE launch over the whole batch on the batch's stream.  (Measured on
// B200: cutting a resident batch into 32-frame chunks over 4 streams is 4 % slower, 53.0k vs 55.2k frames/s --
// the big launches already fill the machine; chunking only pays when it hides the H2D/D2H copies.)
int gforb_batch_run(gforb_batch* b, int n, const gforb_stereo_params* prm) {
    if (!b || n < 1 || n > b->max_frames) { set_error("gforb_batch_run: bad frame count"); return GFORB_ERR_ARG; }
    b->last_launches = 0;
    static const bool no_fork = getenv("GFORB_NO_FORK") != nullptr;      // A/B switch for measurements
    return enqueue_frames(b, 0, n, prm, b->stream, b->timing, !no_fork);
}

int gforb_batch_stage_timing(gforb_batch* b, int enable) {
    if (!b) return GFORB_ERR_ARG;
    if (enable) {
        for (auto& e : b->ev)
            if (!e) {
                cudaError_t err = cudaEventCreate(&e);
                if (err != cudaSuccess) return cuda_fail(err, "cudaEventCreate");
            }
    }
    b->timing = enable != 0;
    return GFORB_OK;
}

int gforb_batch_stage_times(gforb_batch* b, float* ms6) {
    if (!b || !ms6 || !b->timing) { set_error("gforb_batch_stage_times: timing not enabled"); return GFORB_ERR_STATE; }
    cudaError_t e = cudaEventSynchronize(b->ev[6]);
    if (e != cudaSuccess) return cuda_fail(e, "cudaEventSynchronize");
    for (int i = 0; i < 6; ++i) {
        e = cudaEventElapsedTime(&ms6[i], b->ev[i], b->ev[i + 1]);
        if (e != cudaSuccess) return cuda_fail(e, "cudaEventElapsedTime");
    }
    return GFORB_OK;
}

int gforb_batch_launches_per_run(const gforb_batch* b, int with_stereo) {
    if (!b) return GFORB_ERR_ARG;
    return extract_launches(b->eng.g) + (with_stereo ? 1 : 0);
}
int gforb_batch_last_launches(const gforb_batch* b) { return b ? b->last_launches : GFORB_ERR_ARG; }

int gforb_batch_download(gforb_batch* b, int n, int32_t* counts, gforb_keypoint* kps, uint8_t* desc,
                         float* u_right, float* depth) {
    if (!b || n < 1 || n > b->max_frames) return GFORB_ERR_ARG;
    Engine& E = b->eng;
    const size_t K = (size_t)E.g.K, NI = 2 * (size_t)n;
    cudaError_t e = cudaSuccess;
    if (counts) e = cudaMemcpyAsync(counts, E.b.nkp, NI * 4, cudaMemcpyDeviceToHost, b->stream);
    if (kps && e == cudaSuccess) e = cudaMemcpyAsync(kps, E.b.kps, NI * K * sizeof(gforb_keypoint), cudaMemcpyDeviceToHost, b->stream);
    if (desc && e == cudaSuccess) e = cudaMemcpyAsync(desc, E.b.desc, NI * K * 32, cudaMemcpyDeviceToHost, b->stream);
    if (u_right && e == cudaSuccess) e = cudaMemcpyAsync(u_right, E.b.uright, (size_t)n * K * 4, cudaMemcpyDeviceToHost, b->stream);
    if (depth && e == cudaSuccess) e = cudaMemcpyAsync(depth, E.b.depth, (size_t)n * K * 4, cudaMemcpyDeviceToHost, b->stream);
    return e == cudaSuccess ? GFORB_OK : cuda_fail(e, "cudaMemcpyAsync(batch download)");
}

int gforb_batch_device_results(gforb_batch* b, const int32_t** counts, const gforb_keypoint** kps, const uint8_t** desc,
                               const float** u_right, const float** depth, int* K) {
    if (!b || !b->eng.ready) { set_error("gforb_batch_device_results: batch not initialised"); return GFORB_ERR_ARG; }
    Engine& E = b->eng;
    if (counts) *counts = E.b.nkp;
    if (kps) *kps = E.b.kps;
    if (desc) *desc = E.b.desc;
    if (u_right) *u_right = E.b.uright;
    if (depth) *depth = E.b.depth;
    if (K) *K = E.g.K;
    return GFORB_OK;
}

int gforb_batch_sync(gforb_batch* b) {
    if (!b) return GFORB_ERR_ARG;
    return b->eng.check_errors(b->stream);
}

int gforb_batch_debug_copy_only(gforb_batch* b, int on) {
    if (!b) return GFORB_ERR_ARG;
    b->copy_only = on != 0;
    return gforb_batch_reset_graphs(b);
}
int gforb_batch_reset_graphs(gforb_batch* b) {
    if (!b) return GFORB_ERR_ARG;
    for (auto& gs : b->graphs) if (gs.exec) cudaGraphExecDestroy(gs.exec);
    b->graphs.clear();
    return GFORB_OK;
}
int gforb_batch_set_chunk(gforb_batch* b, int frames) {
    if (!b || frames < 0) return GFORB_ERR_ARG;
    b->chunk_frames = frames;
    return gforb_batch_reset_graphs(b);        // the captured sequences were cut with the old chunk size
}

// Host buffers in, host buffers out.  The batch is cut into chunks; chunk c's H2D copy (upload stream), its
// kernels (the batch's stream) and its D2H copy (download stream) are chained with events, so the copy engines
// and the SMs work on different chunks at any time (kernels of even/odd chunks on two streams).  Results are complete when this returns.
namespace {
int frames_chunks(const gforb_batch* b, int n, std::vector<int>& csz);
int frames_enqueue(gforb_batch* b, const uint8_t* left, const uint8_t* right, int n,
                   const gforb_stereo_params* prm, int32_t* counts, gforb_keypoint* kps,
                   uint8_t* desc, float* u_right, float* depth);
}  // namespace

int gforb_stereo_frames_batch_async(gforb_batch* b, const uint8_t* left, const uint8_t* right, int n,
                                    const gforb_stereo_params* prm, int32_t* counts, gforb_keypoint* kps,
                                    uint8_t* desc, float* u_right, float* depth) {
    if (!b || !left || !right || n < 1 || n > b->max_frames) { set_error("gforb_stereo_frames_batch: bad arguments"); return GFORB_ERR_ARG; }
    if (b->eng.b.in_ext) { set_error("gforb_stereo_frames_batch: a caller-owned device input is set (gforb_batch_set_device_input)"); return GFORB_ERR_ARG; }
    static const bool graphs_off = getenv("GFORB_NO_GRAPH") != nullptr;     // A/B switch for measurements
    std::vector<int> csz;
    if (graphs_off || !b->use_graphs || frames_chunks(b, n, csz) != 1)
        return frames_enqueue(b, left, right, n, prm, counts, kps, desc, u_right, depth);
    const void* args[8] = {left, right, counts, kps, desc, u_right, depth, nullptr};
    auto same_prm = [](const gforb_stereo_params& x, const gforb_stereo_params& y) {      // field by field: no padding, no -0.0 / NaN surprises
        return std::memcmp(&x.bf, &y.bf, 4) == 0 && std::memcmp(&x.b, &y.b, 4) == 0 && x.mode == y.mode && x.th_high == y.th_high &&
               x.th_low == y.th_low && std::memcmp(&x.min_x, &y.min_x, 4) == 0;
    };
    for (auto& gs : b->graphs) {
        if (gs.n == n && gs.has_prm == (prm != nullptr) && std::memcmp(gs.args, args, sizeof args) == 0 &&
            (!prm || same_prm(gs.prm, *prm))) {
            if (!gs.exec)          // negative entry: this argument set could not be captured (pageable buffers): plain launches
                return frames_enqueue(b, left, right, n, prm, counts, kps, desc, u_right, depth);
            cudaError_t e = cudaGraphLaunch(gs.exec, b->stream);
            if (e != cudaSuccess) return cuda_fail(e, "cudaGraphLaunch");
            b->last_launches = gs.launches;
            return GFORB_OK;
        }
    }
    // first call with these arguments: capture it (the streams and events of the pipeline exist before the capture starts)
    { const int rcp = ensure_pipeline(b, 1); if (rcp) return rcp; }
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    bool ok = cudaStreamBeginCapture(b->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
    int rc = GFORB_OK;
    if (ok) {
        rc = frames_enqueue(b, left, right, n, prm, counts, kps, desc, u_right, depth);
        ok = cudaStreamEndCapture(b->stream, &graph) == cudaSuccess && rc == GFORB_OK && graph != nullptr;
    }
    if (ok) ok = cudaGraphInstantiate(&exec, graph, 0) == cudaSuccess;
    if (graph) cudaGraphDestroy(graph);
    if (!ok) {                     // e.g. pageable host buffers cannot be captured: THIS argument set stays on plain launches
        (void)cudaGetLastError();
        if (exec) { cudaGraphExecDestroy(exec); exec = nullptr; }
    }
    if (b->graphs.size() >= 8) { if (b->graphs.front().exec) cudaGraphExecDestroy(b->graphs.front().exec); b->graphs.erase(b->graphs.begin()); }
    gforb_batch::GraphSlot gs{};
    std::memcpy(gs.args, args, sizeof args);
    gs.n = n; gs.has_prm = prm != nullptr; if (prm) gs.prm = *prm;
    gs.exec = exec; gs.launches = b->last_launches;
    b->graphs.push_back(gs);
    if (!exec) return frames_enqueue(b, left, right, n, prm, counts, kps, desc, u_right, depth);
    cudaError_t e = cudaGraphLaunch(exec, b->stream);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGraphLaunch");
    return GFORB_OK;
}

namespace {
// chunk schedule: fixed size when the caller set one; otherwise a short first chunk (the pipeline fills after
// 8 frames' worth of H2D), 32-frame chunks in the middle, and a short tail (its kernels + D2H are exposed)
int frames_chunks(const gforb_batch* b, int n, std::vector<int>& csz) {
    csz.clear();
    if (b->chunk_frames > 0 || n < 64) {
        const int chunk = std::min(b->chunk_frames > 0 ? b->chunk_frames : (n >= 48 ? 16 : n), n);
        for (int f = 0; f < n; f += chunk) csz.push_back(std::min(chunk, n - f));
    } else {
        csz.push_back(8); csz.push_back(24);
        int mid = n - 56;
        for (; mid >= 32; mid -= 32) csz.push_back(32);
        if (mid > 0) csz.push_back(mid);
        csz.push_back(16); csz.push_back(8);
    }
    return (int)csz.size();
}

int frames_enqueue(gforb_batch* b, const uint8_t* left, const uint8_t* right, int n,
                   const gforb_stereo_params* prm, int32_t* counts, gforb_keypoint* kps,
                   uint8_t* desc, float* u_right, float* depth) {
    if (!b || !left || !right || n < 1 || n > b->max_frames) { set_error("gforb_stereo_frames_batch: bad arguments"); return GFORB_ERR_ARG; }
    std::vector<int> csz;
    const int nchunks = frames_chunks(b, n, csz);
    std::vector<int> cf0(nchunks, 0);
    for (int c = 1; c < nchunks; ++c) cf0[c] = cf0[c - 1] + csz[c - 1];
    Engine& E = b->eng;
    cudaError_t e;
#define GB(call, what) do { e = (call); if (e != cudaSuccess) return cuda_fail(e, what); } while (0)
    { const int rcp = ensure_pipeline(b, nchunks); if (rcp) return rcp; }
    const size_t K = (size_t)E.g.K, isz = (size_t)E.g.W * E.g.H;
    b->last_launches = 0;
    // the copy streams start after whatever the caller already queued on the batch's stream
    GB(cudaEventRecord(b->ev_start, b->stream), "cudaEventRecord");
    GB(cudaStreamWaitEvent(b->s_up, b->ev_start, 0), "cudaStreamWaitEvent");
    GB(cudaStreamWaitEvent(b->s_down, b->ev_start, 0), "cudaStreamWaitEvent");
    if (nchunks > 1)     // a single chunk runs on the batch's own stream only (and must not leave unjoined work in a capture)
        for (auto& sa : b->s_alt) GB(cudaStreamWaitEvent(sa, b->ev_start, 0), "cudaStreamWaitEvent");
    for (int c = 0; c < nchunks; ++c) {
        const int f0 = cf0[c], m = csz[c];
        GB(cudaMemcpy2DAsync(E.b.in + isz * (2 * (size_t)f0), 2 * isz, left + isz * f0, isz, isz, m,
                             cudaMemcpyHostToDevice, b->s_up), "cudaMemcpy2DAsync(left)");
        GB(cudaMemcpy2DAsync(E.b.in + isz * (2 * (size_t)f0 + 1), 2 * isz, right + isz * f0, isz, isz, m,
                             cudaMemcpyHostToDevice, b->s_up), "cudaMemcpy2DAsync(right)");
        GB(cudaEventRecord(b->ev_up[c], b->s_up), "cudaEventRecord");
    }
    for (int c = 0; c < nchunks; ++c) {
        const int f0 = cf0[c], m = csz[c];
        // even chunks on the batch's stream, odd chunks on a second one: the tail of one chunk's low-occupancy
        // kernels (quadtree, stereo) is filled by the next chunk's streaming kernels
        const int ks = c % b->kstreams;
        cudaStream_t sk = ks ? b->s_alt[ks - 1] : b->stream;
        GB(cudaStreamWaitEvent(sk, b->ev_up[c], 0), "cudaStreamWaitEvent");
        if (!b->copy_only) {
            int rc = enqueue_frames(b, f0, m, prm, sk, false);
            if (rc) return rc;
        }
        GB(cudaEventRecord(b->ev_done[c], sk), "cudaEventRecord");
        GB(cudaStreamWaitEvent(b->s_down, b->ev_done[c], 0), "cudaStreamWaitEvent");
        const size_t i0 = 2 * (size_t)f0, ni = 2 * (size_t)m;
        if (counts) GB(cudaMemcpyAsync(counts + i0, E.b.nkp + i0, ni * 4, cudaMemcpyDeviceToHost, b->s_down), "D2H counts");
        if (kps) GB(cudaMemcpyAsync(kps + i0 * K, E.b.kps + i0 * K, ni * K * sizeof(gforb_keypoint), cudaMemcpyDeviceToHost, b->s_down), "D2H keypoints");
        if (desc) GB(cudaMemcpyAsync(desc + i0 * K * 32, E.b.desc + i0 * K * 32, ni * K * 32, cudaMemcpyDeviceToHost, b->s_down), "D2H descriptors");
        if (prm && u_right) GB(cudaMemcpyAsync(u_right + (size_t)f0 * K, E.b.uright + (size_t)f0 * K, (size_t)m * K * 4, cudaMemcpyDeviceToHost, b->s_down), "D2H uRight");
        if (prm && depth) GB(cudaMemcpyAsync(depth + (size_t)f0 * K, E.b.depth + (size_t)f0 * K, (size_t)m * K * 4, cudaMemcpyDeviceToHost, b->s_down), "D2H depth");
    }
    // rejoin: work queued on the batch's stream after this call is ordered after the last download
    GB(cudaEventRecord(b->ev_down, b->s_down), "cudaEventRecord");
    GB(cudaStreamWaitEvent(b->stream, b->ev_down, 0), "cudaStreamWaitEvent");
#undef GB
    return GFORB_OK;
}
}  // namespace

int gforb_stereo_frames_batch(gforb_batch* b, const uint8_t* left, const uint8_t* right, int n,
                              const gforb_stereo_params* prm, int32_t* counts, gforb_keypoint* kps,
                              uint8_t* desc, float* u_right, float* depth) {
    const int rc = gforb_stereo_frames_batch_async(b, left, right, n, prm, counts, kps, desc, u_right, depth);
    return rc ? rc : gforb_batch_sync(b);
}

/* ---------------------------------------------------------------------------------------------------------------
 * HASHING::MultiIndexHashing, the descriptor side (include/Hashing.h:171-216, src/Hashing.cc); kernels in k_mih.cu
 * --------------------------------------------------------------------------------------------------------------- */
struct gforb_mih {
    int T = 0, bits = 0, cap = 0, dynamics = 0;
    long long upper = 0;                      // entries handed to insert so far: no table holds more slots than that
    int32_t *head = nullptr, *pool_id = nullptr, *pool_next = nullptr, *fill = nullptr, *err = nullptr;
    cudaStream_t stream = nullptr;
    void* io = nullptr; size_t io_bytes = 0;
    void* pin = nullptr; size_t pin_bytes = 0;
    int reserve(void*& p, size_t& have, size_t want, bool host) {
        if (want <= have) return GFORB_OK;
        if (p) { if (host) cudaFreeHost(p); else cudaFree(p); p = nullptr; have = 0; }
        cudaError_t e = host ? cudaHostAlloc(&p, want, cudaHostAllocDefault) : cudaMalloc(&p, want);
        if (e != cudaSuccess) return cuda_fail(e, "gforb_mih: allocation");
        have = want;
        return GFORB_OK;
    }
    ~gforb_mih() {
        for (void* p : {(void*)head, (void*)pool_id, (void*)pool_next, (void*)fill, (void*)err, io}) if (p) cudaFree(p);
        if (pin) cudaFreeHost(pin);
        if (stream) cudaStreamDestroy(stream);
    }
};

namespace {
// room for `more` further entries per table: pools are [T][cap]; growing re-pitches them (slots keep their numbers)
int mih_grow(gforb_mih* h, long long more) {
    const long long need = h->upper + more;
    if (need <= h->cap) return GFORB_OK;
    long long ncap = std::max<long long>(h->cap * 2LL, 4096);
    while (ncap < need) ncap *= 2;
    if (ncap > (1LL << 30)) { set_error("gforb_mih: more than 2^30 entries"); return GFORB_ERR_CAPACITY; }
    int32_t *nid = nullptr, *nnext = nullptr;
    M_CUDA(cudaMalloc(&nid, (size_t)h->T * ncap * 4));
    cudaError_t e = cudaMalloc(&nnext, (size_t)h->T * ncap * 4);
    if (e != cudaSuccess) { cudaFree(nid); return cuda_fail(e, "cudaMalloc(mih pool)"); }
    if (h->cap > 0) {
        cudaMemcpy2DAsync(nid, (size_t)ncap * 4, h->pool_id, (size_t)h->cap * 4, (size_t)h->cap * 4, h->T, cudaMemcpyDeviceToDevice, h->stream);
        cudaMemcpy2DAsync(nnext, (size_t)ncap * 4, h->pool_next, (size_t)h->cap * 4, (size_t)h->cap * 4, h->T, cudaMemcpyDeviceToDevice, h->stream);
        cudaStreamSynchronize(h->stream);
        cudaFree(h->pool_id); cudaFree(h->pool_next);
    }
    h->pool_id = nid; h->pool_next = nnext; h->cap = (int)ncap;
    M_CUDA(cudaGetLastError());
    return GFORB_OK;
}
int mih_check_err(gforb_mih* h) {
    int32_t bits = 0;
    M_CUDA(cudaMemcpyAsync(&bits, h->err, 4, cudaMemcpyDeviceToHost, h->stream));
    M_CUDA(cudaStreamSynchronize(h->stream));
    if (bits) { cudaMemsetAsync(h->err, 0, 4, h->stream); set_error("gforb_mih: slot pool overflow"); return GFORB_ERR_OVERFLOW; }
    return GFORB_OK;
}
}  // namespace

int gforb_mih_create(int feature_length, int table_num, gforb_mih** out) {
    if (!out || feature_length != 256 || table_num < 16 || table_num > 32) {
        set_error("gforb_mih_create: 256-bit descriptors and 16..32 tables (8..16-bit substrings; MapPoint::mvbHashed holds 32 flags)");
        return GFORB_ERR_ARG;
    }
    std::unique_ptr<gforb_mih> h(new (std::nothrow) gforb_mih);
    if (!h) return GFORB_ERR_ARG;
    h->T = table_num; h->bits = feature_length / table_num;               // bits_per_substring (src/Hashing.cc:492)
    int rc = ensure_stream(h->stream);
    if (rc) return rc;
    const size_t heads = (size_t)h->T << h->bits;
    M_CUDA(cudaMalloc(&h->head, heads * 4));
    M_CUDA(cudaMalloc(&h->fill, 32 * 4));
    M_CUDA(cudaMalloc(&h->err, 4));
    M_CUDA(cudaMemsetAsync(h->head, 0xFF, heads * 4, h->stream));         // every bucket empty (initialize, :522-531)
    M_CUDA(cudaMemsetAsync(h->fill, 0, 32 * 4, h->stream));
    M_CUDA(cudaMemsetAsync(h->err, 0, 4, h->stream));
    M_CUDA(cudaStreamSynchronize(h->stream));
    *out = h.release();
    return GFORB_OK;
}
void gforb_mih_destroy(gforb_mih* h) { delete h; }
int gforb_mih_substring_len(const gforb_mih* h) { return h ? h->bits : GFORB_ERR_ARG; }
int gforb_mih_dynamics(const gforb_mih* h) { return h ? h->dynamics : 0; }

int gforb_mih_update_dynamics(gforb_mih* h, double e) {                   // updateDynamics (:957-971); host-side policy
    if (!h) return 0;
    if (std::fabs(e) < 0.001) return h->dynamics;                         // MIN_TIME_DIFF
    if (e > 0 && h->dynamics + 20 > 1) h->dynamics--;                     // NUM_CANDIDATE_FEATURES, MIN_BUCKET_SIZE
    if (e < 0 && h->dynamics + 20 < 20) h->dynamics++;                    // MAX_BUCKET_SIZE
    return h->dynamics;
}

int gforb_mih_hash(gforb_mih* h, const uint8_t* desc, int n, uint32_t* keys) {
    if (!h || n < 0 || (n > 0 && (!desc || !keys))) { set_error("gforb_mih_hash: bad arguments"); return GFORB_ERR_ARG; }
    if (n == 0) return GFORB_OK;
    const size_t in = al16((size_t)n * 32), outb = (size_t)n * h->T * 4;
    int rc = h->reserve(h->io, h->io_bytes, in + outb, false);
    if (rc) return rc;
    uint8_t* dev = static_cast<uint8_t*>(h->io);
    M_CUDA(cudaMemcpyAsync(dev, desc, (size_t)n * 32, cudaMemcpyHostToDevice, h->stream));
    launch_mih_hash(dev, n, h->T, h->bits, reinterpret_cast<uint32_t*>(dev + in), h->stream);
    M_CUDA(cudaMemcpyAsync(keys, dev + in, outb, cudaMemcpyDeviceToHost, h->stream));
    M_CUDA(cudaStreamSynchronize(h->stream));
    return GFORB_OK;
}

int gforb_mih_insert_device(gforb_mih* h, const uint8_t* dev_desc, const int32_t* dev_entry, int n, uint32_t* dev_hashed,
                            uint8_t* dev_success, void* cuda_stream) {
    if (!h || n < 0 || (n > 0 && (!dev_desc || !dev_entry || !dev_hashed || !dev_success))) { set_error("gforb_mih_insert: bad arguments"); return GFORB_ERR_ARG; }
    if ((uintptr_t)dev_desc & 3) { set_error("gforb_mih_insert: descriptors must be 4-byte aligned"); return GFORB_ERR_ARG; }
    if (n == 0) return GFORB_OK;
    cudaStream_t s = cuda_stream ? (cudaStream_t)cuda_stream : h->stream;
    int rc = mih_grow(h, n);
    if (rc) return rc;
    M_CUDA(cudaMemsetAsync(dev_success, 0, (size_t)n, s));
    for (int i0 = 0; i0 < n; i0 += mih_max_batch()) {                      // chunks keep the reference's insertion order
        const int c = std::min(n - i0, mih_max_batch());
        if (launch_mih_insert(dev_desc + 32 * (size_t)i0, dev_entry + i0, c, h->T, h->bits, h->cap, h->head, h->pool_id, h->pool_next,
                              h->fill, dev_hashed + i0, dev_success + i0, h->err, s)) {
            set_error("gforb_mih_insert: shared-memory opt-in failed");
            return GFORB_ERR_CUDA;
        }
    }
    h->upper += n;
    M_CUDA(cudaGetLastError());
    return GFORB_OK;
}

int gforb_mih_insert(gforb_mih* h, const uint8_t* desc, const int32_t* entry, int n, uint32_t* hashed, uint8_t* success) {
    if (!h || n < 0 || (n > 0 && (!desc || !entry))) { set_error("gforb_mih_insert: bad arguments"); return GFORB_ERR_ARG; }
    if (n == 0) return GFORB_OK;
    const size_t od = 0, oe = al16((size_t)n * 32), oh = oe + al16((size_t)n * 4), os = oh + al16((size_t)n * 4), total = os + al16((size_t)n);
    int rc = h->reserve(h->io, h->io_bytes, total, false);
    if (rc) return rc;
    rc = h->reserve(h->pin, h->pin_bytes, total, true);
    if (rc) return rc;
    uint8_t* dev = static_cast<uint8_t*>(h->io);
    uint8_t* pin = static_cast<uint8_t*>(h->pin);
    std::memcpy(pin + od, desc, (size_t)n * 32);
    std::memcpy(pin + oe, entry, (size_t)n * 4);
    if (hashed) std::memcpy(pin + oh, hashed, (size_t)n * 4); else std::memset(pin + oh, 0, (size_t)n * 4);
    M_CUDA(cudaMemcpyAsync(dev, pin, os, cudaMemcpyHostToDevice, h->stream));
    rc = gforb_mih_insert_device(h, dev + od, reinterpret_cast<int32_t*>(dev + oe), n, reinterpret_cast<uint32_t*>(dev + oh), dev + os, nullptr);
    if (rc) return rc;
    M_CUDA(cudaMemcpyAsync(pin + oh, dev + oh, total - oh, cudaMemcpyDeviceToHost, h->stream));
    rc = mih_check_err(h);
    if (rc) return rc;
    if (hashed) std::memcpy(hashed, pin + oh, (size_t)n * 4);
    if (success) std::memcpy(success, pin + os, (size_t)n);
    return GFORB_OK;
}

namespace {
int mih_query_impl(gforb_mih* h, const uint8_t* qdesc, int m, const int* take, int keep, int need, int32_t* cand, int32_t* count,
                   int32_t* addi, int32_t* addi_count, int stride, bool device, cudaStream_t s, const char* fn) {
    if (stride < need) { set_error(std::string(fn) + ": stride smaller than the largest possible candidate count"); return GFORB_ERR_CAPACITY; }
    if (m == 0) return GFORB_OK;
    if (device) {
        launch_mih_query(qdesc, m, h->T, h->bits, h->cap, h->head, h->pool_id, h->pool_next, take, keep, cand, count, addi, addi_count, stride, s);
        M_CUDA(cudaGetLastError());
        return GFORB_OK;
    }
    const size_t oq = 0, oc = al16((size_t)m * 32), on = oc + al16((size_t)m * stride * 4), oa = on + al16((size_t)m * 4),
                 oan = oa + (addi ? al16((size_t)m * stride * 4) : 0), total = oan + (addi ? al16((size_t)m * 4) : 0);
    int rc = h->reserve(h->io, h->io_bytes, total, false);
    if (rc) return rc;
    uint8_t* dev = static_cast<uint8_t*>(h->io);
    M_CUDA(cudaMemcpyAsync(dev + oq, qdesc, (size_t)m * 32, cudaMemcpyHostToDevice, h->stream));
    M_CUDA(cudaMemsetAsync(dev + oc, 0xFF, on - oc, h->stream));           // unused tail of every row reads -1
    if (addi) M_CUDA(cudaMemsetAsync(dev + oa, 0xFF, oan - oa, h->stream));
    launch_mih_query(dev + oq, m, h->T, h->bits, h->cap, h->head, h->pool_id, h->pool_next, take, keep, reinterpret_cast<int32_t*>(dev + oc),
                     reinterpret_cast<int32_t*>(dev + on), addi ? reinterpret_cast<int32_t*>(dev + oa) : nullptr,
                     addi ? reinterpret_cast<int32_t*>(dev + oan) : nullptr, stride, h->stream);
    M_CUDA(cudaMemcpyAsync(cand, dev + oc, (size_t)m * stride * 4, cudaMemcpyDeviceToHost, h->stream));
    M_CUDA(cudaMemcpyAsync(count, dev + on, (size_t)m * 4, cudaMemcpyDeviceToHost, h->stream));
    if (addi) {
        M_CUDA(cudaMemcpyAsync(addi, dev + oa, (size_t)m * stride * 4, cudaMemcpyDeviceToHost, h->stream));
        M_CUDA(cudaMemcpyAsync(addi_count, dev + oan, (size_t)m * 4, cudaMemcpyDeviceToHost, h->stream));
    }
    M_CUDA(cudaStreamSynchronize(h->stream));
    return GFORB_OK;
}
// Bucket::back's n per table for the plain form (:699-846): tables 0 .. nactive-1, nretrieves + dynamics each
int mih_take_plain(const gforb_mih* h, int nactive, int nretrieves, int* take) {
    if (nactive == 0) nactive = h->T;                                      // nactive_tables_top == NULL
    int need = 0;
    for (int k = 0; k < 32; ++k) {
        take[k] = k < nactive ? std::max(nretrieves + h->dynamics, 0) : 0;  // n < 0 is undefined in the reference: nothing
        need += take[k];
    }
    return need;
}
int mih_take_split(const gforb_mih* h, const int32_t* nfeatures, int* take) {
    int need = 0;
    for (int k = 0; k < 32; ++k) {
        take[k] = (k < h->T && nfeatures[k] > 0) ? std::max(nfeatures[k] + h->dynamics, 0) : 0;     // :868-872
        need += take[k];
    }
    return need;
}
}  // namespace

int gforb_mih_query(gforb_mih* h, const uint8_t* qdesc, int m, int nactive_tables, int nretrieves, int32_t* cand, int stride, int32_t* count) {
    if (!h || m < 0 || nactive_tables < 0 || nactive_tables > (h ? h->T : 0) || nretrieves < 0 || (m > 0 && (!qdesc || !cand || !count))) {
        set_error("gforb_mih_query: bad arguments");
        return GFORB_ERR_ARG;
    }
    int take[32];
    const int need = mih_take_plain(h, nactive_tables, nretrieves, take);
    return mih_query_impl(h, qdesc, m, take, INT_MAX, need, cand, count, nullptr, nullptr, stride, false, h->stream, "gforb_mih_query");
}
int gforb_mih_query_device(gforb_mih* h, const uint8_t* dev_qdesc, int m, int nactive_tables, int nretrieves, int32_t* dev_cand, int stride,
                           int32_t* dev_count, void* cuda_stream) {
    if (!h || m < 0 || nactive_tables < 0 || nactive_tables > (h ? h->T : 0) || nretrieves < 0 || (m > 0 && (!dev_qdesc || !dev_cand || !dev_count)) ||
        ((uintptr_t)dev_qdesc & 3)) {
        set_error("gforb_mih_query: bad arguments");
        return GFORB_ERR_ARG;
    }
    int take[32];
    const int need = mih_take_plain(h, nactive_tables, nretrieves, take);
    return mih_query_impl(h, dev_qdesc, m, take, INT_MAX, need, dev_cand, dev_count, nullptr, nullptr, stride, true,
                          cuda_stream ? (cudaStream_t)cuda_stream : h->stream, "gforb_mih_query");
}
int gforb_mih_query_split(gforb_mih* h, const uint8_t* qdesc, int m, const int32_t* nfeatures, int32_t* cand, int32_t* count,
                          int32_t* addi, int32_t* addi_count, int stride) {
    if (!h || m < 0 || !nfeatures || (m > 0 && (!qdesc || !cand || !count || !addi || !addi_count))) {
        set_error("gforb_mih_query_split: bad arguments");
        return GFORB_ERR_ARG;
    }
    int take[32];
    const int need = mih_take_split(h, nfeatures, take);
    return mih_query_impl(h, qdesc, m, take, 20 + h->dynamics, need, cand, count, addi, addi_count, stride, false, h->stream,
                          "gforb_mih_query_split");
}
int gforb_mih_sync(gforb_mih* h) {
    if (!h) return GFORB_ERR_ARG;
    return mih_check_err(h);
}

}  // extern "C"
