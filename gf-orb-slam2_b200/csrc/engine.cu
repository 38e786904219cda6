// Host engine: constructor tables, level geometry, arenas, stage sequencing.
#include <cstdlib>

#include "engine.hpp"
#include "tma.cuh"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstring>

namespace gforb {

namespace {
thread_local std::string g_err = "";
inline int cv_round(float v) { return (int)lrintf(v); }     // round-half-even, like cvRound
inline int cv_round(double v) { return (int)lrint(v); }
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
}  // namespace

void set_error(const std::string& msg) { g_err = msg; }
const char* last_error() { return g_err.c_str(); }
int cuda_fail(cudaError_t e, const char* what) {
    set_error(std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")");
    return GFORB_ERR_CUDA;
}

// reference src/ORBextractor.cc:409-469.  scaleFactor is a double member initialised from the float argument.
int make_tables(const gforb_params& p, Tables& t) {
    if (p.nlevels < 1 || p.nlevels > kMaxLevels || p.nfeatures < 0 || !(p.scale_factor > 1.0f) ||
        p.nfeatures > 60000) {
        set_error("gforb: bad extractor parameters (need 1<=nlevels<=12, scale_factor>1, 0<=nfeatures<=60000)");
        return GFORB_ERR_ARG;
    }
    t.prm = p;
    const int nl = p.nlevels;
    const double sf = (double)p.scale_factor;
    t.scale.assign(nl, 1.f); t.sigma2.assign(nl, 1.f);
    for (int i = 1; i < nl; ++i) {
        t.scale[i] = (float)(t.scale[i - 1] * sf);
        t.sigma2[i] = t.scale[i] * t.scale[i];
    }
    t.inv_scale.resize(nl); t.inv_sigma2.resize(nl);
    for (int i = 0; i < nl; ++i) { t.inv_scale[i] = 1.0f / t.scale[i]; t.inv_sigma2[i] = 1.0f / t.sigma2[i]; }

    t.quota.assign(nl, 0);
    const float factor = (float)(1.0f / sf);
    float want = p.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nl));
    int sum = 0;
    for (int l = 0; l < nl - 1; ++l) {
        t.quota[l] = cv_round(want);
        sum += t.quota[l];
        want *= factor;
    }
    t.quota[nl - 1] = std::max(p.nfeatures - sum, 0);

    // rows of the radius-15 disc (:453-468)
    std::memset(t.umax, 0, sizeof(t.umax));
    const int vmax = (int)std::floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
    const int vmin = (int)std::ceil(kHalfPatch * std::sqrt(2.f) / 2);
    for (int v = 0; v <= vmax; ++v) t.umax[v] = cv_round(std::sqrt((double)(kHalfPatch * kHalfPatch) - v * v));
    for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
        while (t.umax[v0] == t.umax[v0 + 1]) ++v0;
        t.umax[v] = v0;
        ++v0;
    }
    return GFORB_OK;
}

namespace {
// cv::resize INTER_LINEAR per-axis table: source offset and the two 11-bit weights, packed for the kernel as
// {ofs, (w1 << 16) | w0}.  Same arithmetic as OpenCV's resize(): fx = (float)((d+0.5)*scale-0.5) with
// scale = 1/((double)dst/src); weights = saturate_cast<short>(f*2048) (round-half-even).
void axis_table(int slen, int dlen, std::vector<int32_t>& out) {
    const double scale = 1. / ((double)dlen / slen);
    for (int d = 0; d < dlen; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = (int)std::floor(f);
        f -= (float)s;
        if (s < 0) { f = 0.f; s = 0; }
        if (s >= slen - 1) { f = 0.f; s = slen - 1; }
        const int w0 = std::min(std::max(cv_round((1.f - f) * 2048.f), -32768), 32767);
        const int w1 = std::min(std::max(cv_round(f * 2048.f), -32768), 32767);
        out.push_back(s);
        out.push_back((int32_t)(((uint32_t)w1 << 16) | ((uint32_t)w0 & 0xFFFFu)));
    }
}

// 4-pixel-group table of the fast resize kernel (k_pyramid.cu): per group of 4 adjacent output pixels
// {aligned source byte offset wb, PRMT selector for pixels 0-1 out of source words (w0,w1), two selectors for
// pixels 2-3 out of (w0,w1) then w2, 4 x packed weights}.  Returns false when a group's taps do not fit the
// 12-byte window (scale factors above ~2.3): that level then runs the generic kernel.
bool group_table(const std::vector<int32_t>& tabs, int tab_x, int dlen, std::vector<int32_t>& out) {
    const int ng = (dlen + 3) / 4;
    for (int gi = 0; gi < ng; ++gi) {
        int sx[4], cf[4];
        for (int k = 0; k < 4; ++k) {
            const int x = std::min(4 * gi + k, dlen - 1);
            sx[k] = tabs[tab_x + 2 * x];
            cf[k] = tabs[tab_x + 2 * x + 1];
            if ((cf[k] & 0xFFFF) > 2048 || ((uint32_t)cf[k] >> 16) > 2048) return false;   // weights must be plain u16
        }
        const int wb = sx[0] & ~3;
        const int o[4] = {sx[0] - wb, sx[1] - wb, sx[2] - wb, sx[3] - wb};
        if (o[1] < o[0] || o[1] + 1 > 7 || o[2] < 0 || o[3] < o[2] || o[3] + 1 > 11) return false;
        const uint32_t selA = (uint32_t)o[0] | ((uint32_t)(o[0] + 1) << 4) | ((uint32_t)o[1] << 8) | ((uint32_t)(o[1] + 1) << 12);
        const int q[4] = {o[2], o[2] + 1, o[3], o[3] + 1};
        uint32_t s1 = 0, s2 = 0;
        for (int i = 0; i < 4; ++i) {
            s1 |= (uint32_t)(q[i] <= 7 ? q[i] : 0) << (4 * i);
            s2 |= (uint32_t)(q[i] <= 7 ? i : 4 + (q[i] - 8)) << (4 * i);
        }
        out.push_back(wb); out.push_back((int32_t)selA); out.push_back((int32_t)s1); out.push_back((int32_t)s2);
        for (int k = 0; k < 4; ++k) out.push_back(cf[k]);
    }
    return true;
}
}  // namespace

int make_geometry(const Tables& t, const gforb_limits& lim, int W, int H, int nimg, Geom& g,
                  std::vector<int32_t>& tabs, size_t& pyr_bytes, size_t& blur_bytes) {
    std::memset(&g, 0, sizeof(g));
    tabs.clear();
    const int nl = t.prm.nlevels;
    g.nlevels = nl; g.W = W; g.H = H; g.nimg = nimg; g.img0 = 0;
    g.ini_th = t.prm.ini_th_fast; g.min_th = t.prm.min_th_fast;
    if (W > 4096 + 32 || H > 4096 + 32 || nimg < 1 || nimg > 65535) {
        set_error("gforb: image larger than 4128 px or bad batch size");
        return GFORB_ERR_ARG;
    }
    size_t poff = 0, boff = 0;
    int cell_base = 0, kbase = 0, btile = 0, cell_cap = 1, max_kcap = 0;
    for (int l = 0; l < nl; ++l) {
        LevelGeom& L = g.lv[l];
        L.w = cv_round((float)W * t.inv_scale[l]);      // :1180-1181
        L.h = cv_round((float)H * t.inv_scale[l]);
        if (L.w < 40 || L.h < 40) {
            char msg[160];
            snprintf(msg, sizeof msg, "gforb: %dx%d image too small for %d pyramid levels (level %d is %dx%d, need >= 40)",
                     W, H, nl, l, L.w, L.h);
            set_error(msg);
            return GFORB_ERR_ARG;
        }
        L.pitch = (int)align_up((size_t)kXPad + L.w + kEdge, 64);
        L.rows = L.h + 2 * kEdge;
        L.img_stride = (size_t)L.pitch * L.rows;
        L.off = poff;
        poff += L.img_stride * nimg;
        L.bpitch = (int)align_up((size_t)L.w, 64);
        L.bimg_stride = (size_t)L.bpitch * L.h;
        L.boff = boff;
        boff += L.bimg_stride * nimg;

        // cell grid, float arithmetic as written (:771-783)
        L.maxBX = L.w - kEdge + 3; L.maxBY = L.h - kEdge + 3;
        const float width = (float)(L.maxBX - kMinBorder), height = (float)(L.maxBY - kMinBorder);
        L.nCols = (int)(width / 30.f); L.nRows = (int)(height / 30.f);
        if (L.nCols < 1 || L.nRows < 1) { L.nCols = L.nRows = 0; L.wCell = L.hCell = 1; }
        else { L.wCell = (int)std::ceil(width / L.nCols); L.hCell = (int)std::ceil(height / L.nRows); }
        L.cell_base = cell_base; L.ncells = L.nCols * L.nRows;
        cell_base += L.ncells;
        cell_cap = std::max(cell_cap, ((L.wCell + 1) / 2) * ((L.hCell + 1) / 2));
        L.fast_bw = (int)align_up((size_t)L.wCell + 6 + 15, 16);   // TMA box: 16-byte multiple, start rounded down to 16
        L.fast_bh = L.hCell + 6;

        // quadtree (:543-546)
        L.quota = t.quota[l];
        L.nIni = (int)std::round((float)(L.maxBX - kMinBorder) / (float)(L.maxBY - kMinBorder));
        L.hX = L.nIni > 0 ? (float)(L.maxBX - kMinBorder) / L.nIni : 1.f;
        L.kcap = std::max(L.quota + 3, 4 * std::max(L.nIni, 1)) + 1;
        L.kbase = kbase;
        kbase += L.kcap;
        max_kcap = std::max(max_kcap, L.kcap);

        L.scale = t.scale[l]; L.inv_scale = t.inv_scale[l];
        L.patch_size = (float)(int)(31 * t.scale[l]);    // :839

        // blur: a warp task = 128 columns x one strip; strips are balanced so that none exceeds 56 rows
        L.btile_base = btile; L.btiles_x = (L.w + 127) / 128;
        const int nstrips = (L.h + 55) / 56;
        L.bstrip_rows = (L.h + nstrips - 1) / nstrips;
        btile += L.btiles_x * nstrips;

        if (l > 0) {
            const LevelGeom& P = g.lv[l - 1];
            const double sx = 1. / ((double)L.w / P.w), sy = 1. / ((double)L.h / P.h);
            L.area2x = (std::abs(sx - cv_round(sx)) < DBL_EPSILON && std::abs(sy - cv_round(sy)) < DBL_EPSILON &&
                        cv_round(sx) == 2 && cv_round(sy) == 2) ? 1 : 0;
            L.tab_x = (int)tabs.size(); axis_table(P.w, L.w, tabs);
            L.tab_y = (int)tabs.size(); axis_table(P.h, L.h, tabs);
            while (tabs.size() % 4) tabs.push_back(0);           // group entries are read as 2 x uint4
            std::vector<int32_t> grp;
            L.tab_g = -1;
            if (!L.area2x && group_table(tabs, L.tab_x, L.w, grp)) {
                L.tab_g = (int)tabs.size();
                tabs.insert(tabs.end(), grp.begin(), grp.end());
                // TMA box of the source level that covers any 128-pixel x kResizeTileRows-row output tile
                const int ng = (L.w + 3) / 4;
                int bw = 0, bh = 0;
                for (int bx = 0; bx * 32 < ng; ++bx) {
                    const int g0 = std::min(bx * 32, std::max(ng - 32, 0));     // last warp column is shifted left
                    const int gl = std::min(g0 + 31, ng - 1);
                    const int xb = grp[8 * g0] & ~15;
                    bw = std::max(bw, (int)align_up((size_t)(grp[8 * gl] + 12 - xb), 16));
                }
                for (int y0 = 0; y0 < L.h; y0 += kResizeTileRows) {
                    const int yl = std::min(y0 + kResizeTileRows - 1, L.h - 1);
                    bh = std::max(bh, tabs[L.tab_y + 2 * yl] + 2 - tabs[L.tab_y + 2 * y0]);
                }
                if (bw <= 256 && bh <= 256) { L.rs_bw = bw; L.rs_bh = bh; }
            }
        }
    }
    g.cells_total = cell_base;
    g.cell_cap = cell_cap;
    g.kstage = kbase;
    g.K = (int)align_up((size_t)kbase, 8);
    g.btiles_total = btile;
    g.max_cand = lim.max_candidates_per_level > 0 ? lim.max_candidates_per_level : 12288;
    g.max_cand = std::min(std::max(g.max_cand, 256), 65535);
    g.max_nodes = std::max(256, max_kcap + 4);
    if (g.K > 65535) { set_error("gforb: too many keypoints per image"); return GFORB_ERR_ARG; }
    pyr_bytes = poff; blur_bytes = boff;
    return GFORB_OK;
}

// ------------------------------------------------------------------------------------------------
int encode_tile_map_u8(CUtensorMap* map, const void* base, uint64_t pitch, uint64_t rows, uint64_t nimg,
                       uint32_t box_w, uint32_t box_h) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) return -1;
        fn = (EncodeFn)p;
    }
    const cuuint64_t dims[3] = {pitch, rows, nimg};
    const cuuint64_t strides[2] = {pitch, pitch * rows};
    const cuuint32_t box[3] = {box_w, box_h, 1};
    const cuuint32_t estr[3] = {1, 1, 1};
    return (int)fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
}

Engine::~Engine() { release(); }

void Engine::release() {
    auto fr = [](auto*& p) { if (p) cudaFree(p); p = nullptr; };
    fr(b.in); fr(b.cell_tab); fr(b.pyr); fr(b.blur); fr(b.tabs); fr(b.cell_cnt); fr(b.cell_pts); fr(b.lvl_cnt); fr(b.oct_retry); fr(b.lvl_pts);
    fr(b.kps); fr(b.desc); fr(b.nkp); fr(b.uright); fr(b.depth); fr(b.st_best_dist); fr(b.st_best_idx);
    fr(b.st_sad); fr(b.st_nmatch); fr(b.st_cut); fr(b.st_ticket); fr(b.err);
    if (h_small) cudaFreeHost(h_small);
    h_small = nullptr;
    if (s_aux) { cudaStreamDestroy(s_aux); s_aux = nullptr; }
    if (ev_fork) { cudaEventDestroy(ev_fork); ev_fork = nullptr; }
    if (ev_join) { cudaEventDestroy(ev_join); ev_join = nullptr; }
    ready = false; pyramid_valid = false;
}

#define GFORB_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { release(); return cuda_fail(e_, #call); } } while (0)

int Engine::init(const Tables& t, const gforb_limits& lim, int W, int H, int nimg, int nframes_stereo) {
    release();
    std::vector<int32_t> tabs;
    size_t pyr_bytes = 0, blur_bytes = 0;
    int rc = make_geometry(t, lim, W, H, nimg, g, tabs, pyr_bytes, blur_bytes);
    if (rc != GFORB_OK) return rc;
    nframes_cap = nframes_stereo;
    const size_t K = (size_t)g.K, NI = (size_t)nimg, NF = (size_t)std::max(nframes_stereo, 1);

    GFORB_CUDA(cudaMalloc(&b.in, NI * W * H));
    pyr_bytes += 64;    // the right frame word of a row may run 2 bytes past the pitch (into row padding)
    GFORB_CUDA(cudaMalloc(&b.pyr, pyr_bytes));
    // alignment bytes outside the 19-px frame are never written, nor (without full_frames) its outer 16 px: nothing reads them.
    // GFORB_POISON_PYR=1 fills the arena with 0xCD instead, so that a test run proves it (tools / parity suites).
    GFORB_CUDA(cudaMemset(b.pyr, getenv("GFORB_POISON_PYR") ? 0xCD : 0, pyr_bytes));
    GFORB_CUDA(cudaMalloc(&b.blur, blur_bytes));
    GFORB_CUDA(cudaMalloc(&b.tabs, std::max<size_t>(tabs.size(), 2) * 4));
    GFORB_CUDA(cudaMalloc(&b.cell_cnt, NI * g.cells_total * 4));
    {
        std::vector<uint32_t> ct;
        build_fast_cell_table(g, ct);
        GFORB_CUDA(cudaMalloc(&b.cell_tab, std::max<size_t>(ct.size(), 8) * 4));
        if (!ct.empty()) GFORB_CUDA(cudaMemcpy(b.cell_tab, ct.data(), ct.size() * 4, cudaMemcpyHostToDevice));
    }
    GFORB_CUDA(cudaMalloc(&b.cell_pts, NI * g.cells_total * (size_t)g.cell_cap * 4));
    GFORB_CUDA(cudaMalloc(&b.lvl_cnt, NI * g.nlevels * 4));
    GFORB_CUDA(cudaMalloc(&b.oct_retry, NI * g.nlevels * 4));
    GFORB_CUDA(cudaMemset(b.oct_retry, 0, NI * g.nlevels * 4));
    GFORB_CUDA(cudaMalloc(&b.lvl_pts, NI * g.kstage * 4));
    GFORB_CUDA(cudaMalloc(&b.kps, NI * K * sizeof(gforb_keypoint)));
    GFORB_CUDA(cudaMalloc(&b.desc, NI * K * 32));
    GFORB_CUDA(cudaMalloc(&b.nkp, NI * 4));
    GFORB_CUDA(cudaMalloc(&b.uright, NF * K * 4));
    GFORB_CUDA(cudaMalloc(&b.depth, NF * K * 4));
    GFORB_CUDA(cudaMalloc(&b.st_best_dist, NF * K * 4));
    GFORB_CUDA(cudaMalloc(&b.st_best_idx, NF * K * 4));
    GFORB_CUDA(cudaMalloc(&b.st_sad, NF * K * 4));
    GFORB_CUDA(cudaMalloc(&b.st_nmatch, NF * 4));
    GFORB_CUDA(cudaMalloc(&b.st_cut, NF * K * 4));
    GFORB_CUDA(cudaMalloc(&b.st_ticket, 2 * NF * 4));
    GFORB_CUDA(cudaMemset(b.st_ticket, 0, 2 * NF * 4));
    GFORB_CUDA(cudaMalloc(&b.err, 16));
    GFORB_CUDA(cudaMemset(b.err, 0, 16));
    GFORB_CUDA(cudaMemset(b.nkp, 0, NI * 4));
    GFORB_CUDA(cudaMemset(b.lvl_cnt, 0, NI * g.nlevels * 4));
    if (!tabs.empty()) GFORB_CUDA(cudaMemcpy(b.tabs, tabs.data(), tabs.size() * 4, cudaMemcpyHostToDevice));
    h_small_ints = NI * (size_t)(g.nlevels + 1) + NF + 16;
    GFORB_CUDA(cudaHostAlloc(&h_small, h_small_ints * 4, cudaHostAllocDefault));

    const size_t osm = octree_smem_bytes(g);
    if (osm > 227 * 1024) {
        release();
        set_error("gforb: quadtree working set exceeds 227 KB of shared memory; lower max_candidates_per_level or nfeatures");
        return GFORB_ERR_ARG;
    }
    GFORB_CUDA((cudaError_t)octree_configure(g));
    GFORB_CUDA((cudaError_t)fast_configure(g));
    // TMA descriptors: one per level for the FAST cell tiles
    for (int l = 0; l < g.nlevels; ++l) {
        const LevelGeom& L = g.lv[l];
        const int rcu = encode_tile_map_u8(&maps.fast[l], b.pyr + L.off, (uint64_t)L.pitch, (uint64_t)L.rows,
                                           (uint64_t)nimg, (uint32_t)L.fast_bw, (uint32_t)L.fast_bh);
        if (rcu != 0) {
            release();
            set_error("gforb: cuTensorMapEncodeTiled failed (" + std::to_string(rcu) + ")");
            return GFORB_ERR_CUDA;
        }
    }
    for (int l = 1; l < g.nlevels; ++l) {
        const LevelGeom& L = g.lv[l];
        const LevelGeom& P = g.lv[l - 1];
        if (!L.rs_bw) continue;
        const int rcu = encode_tile_map_u8(&rsmaps[l], b.pyr + P.off, (uint64_t)P.pitch, (uint64_t)P.rows,
                                           (uint64_t)nimg, (uint32_t)L.rs_bw, (uint32_t)L.rs_bh);
        if (rcu != 0) {
            release();
            set_error("gforb: cuTensorMapEncodeTiled(resize) failed (" + std::to_string(rcu) + ")");
            return GFORB_ERR_CUDA;
        }
    }
    ready = true;
    return GFORB_OK;
}

int Engine::upload_images(const uint8_t* host, size_t pitch, int first_img, int step, int count, cudaStream_t s) {
    const size_t isz = (size_t)g.W * g.H;
    if (step == 1 && pitch == (size_t)g.W) {
        cudaError_t e = cudaMemcpyAsync(b.in + isz * first_img, host, isz * count, cudaMemcpyHostToDevice, s);
        if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpyAsync(images)");
        return GFORB_OK;
    }
    // strided destination (left/right interleave): one 2-D copy whose "rows" are whole images when tight
    if (pitch == (size_t)g.W) {
        cudaError_t e = cudaMemcpy2DAsync(b.in + isz * first_img, isz * step, host, isz, isz, count,
                                          cudaMemcpyHostToDevice, s);
        if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpy2DAsync(images)");
        return GFORB_OK;
    }
    for (int i = 0; i < count; ++i) {
        cudaError_t e = cudaMemcpy2DAsync(b.in + isz * (first_img + (size_t)i * step), g.W,
                                          host + (size_t)i * pitch * g.H, pitch, g.W, g.H, cudaMemcpyHostToDevice, s);
        if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpy2DAsync(image)");
    }
    return GFORB_OK;
}

int Engine::run_pyramid(int nimg, cudaStream_t s) {
    g.img0 = 0;
    launch_pyramid(g, b, nimg, s, rsmaps, full_frames);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "pyramid launch");
    pyramid_valid = true;
    return GFORB_OK;
}

int Engine::run_extract(int nimg, cudaStream_t s, cudaEvent_t* ev, int img0, bool fork_blur) {
    g.img0 = img0;   // Geom travels by value as a kernel parameter
    // Dependencies: pyramid -> FAST -> quadtree -> describe, pyramid -> blur -> describe.  The quadtree is latency-bound
    // (38 % of the issue slots, ncu) and the blur a streaming kernel that needs only the pyramid, so with fork_blur the
    // blur runs on a side stream next to the quadtree and joins before describe.  Stage timing keeps the serial order.
    const bool fork = fork_blur && !ev;
    if (fork && !s_aux) {
        if (cudaStreamCreateWithFlags(&s_aux, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming) != cudaSuccess ||
            cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming) != cudaSuccess)
            return cuda_fail(cudaGetLastError(), "side stream");
    }
    if (ev) cudaEventRecord(ev[0], s);
    GFORB_RANGE_PUSH("gforb ComputePyramid", 0);
    launch_pyramid(g, b, nimg, s, rsmaps, full_frames);
    GFORB_RANGE_POP();
    if (ev) cudaEventRecord(ev[1], s);
    GFORB_RANGE_PUSH("gforb FAST cells", 1);
    launch_fast(g, maps, b, nimg, s);
    GFORB_RANGE_POP();
    if (ev) cudaEventRecord(ev[2], s);
    if (fork) {
        cudaEventRecord(ev_fork, s);
        cudaStreamWaitEvent(s_aux, ev_fork, 0);
        GFORB_RANGE_PUSH("gforb DistributeOctTree", 2);
        launch_octree(g, b, nimg, s);
        GFORB_RANGE_POP();
        GFORB_RANGE_PUSH("gforb GaussianBlur (side stream)", 3);
        launch_blur(g, b, nimg, s_aux);
        GFORB_RANGE_POP();
        cudaEventRecord(ev_join, s_aux);
        cudaStreamWaitEvent(s, ev_join, 0);
    } else {
        GFORB_RANGE_PUSH("gforb DistributeOctTree", 2);
        launch_octree(g, b, nimg, s);
        GFORB_RANGE_POP();
        if (ev) cudaEventRecord(ev[3], s);
        GFORB_RANGE_PUSH("gforb GaussianBlur", 3);
        launch_blur(g, b, nimg, s);
        GFORB_RANGE_POP();
    }
    if (ev) cudaEventRecord(ev[4], s);
    GFORB_RANGE_PUSH("gforb IC_Angle + computeOrbDescriptor", 4);
    launch_describe(g, b, nimg, s);
    GFORB_RANGE_POP();
    if (ev) cudaEventRecord(ev[5], s);
    g.img0 = 0;
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return cuda_fail(e, "extract launch");
    pyramid_valid = true;
    return GFORB_OK;
}

int Engine::check_errors(cudaStream_t s) {
    cudaError_t e = cudaMemcpyAsync(h_small, b.err, 4, cudaMemcpyDeviceToHost, s);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpyAsync(err)");
    e = cudaStreamSynchronize(s);
    if (e != cudaSuccess) return cuda_fail(e, "cudaStreamSynchronize");
    const int bits = h_small[0];
    if (bits) {
        cudaMemsetAsync(b.err, 0, 4, s);
        char msg[200];
        snprintf(msg, sizeof msg, "gforb: device working set overflow (flags 0x%x:%s%s%s%s)", bits,
                 bits & kErrCandOverflow ? " candidates-per-level" : "", bits & kErrCellOverflow ? " cell" : "",
                 bits & kErrNodeOverflow ? " quadtree-nodes" : "", bits & kErrKpOverflow ? " keypoints" : "");
        set_error(msg);
        return GFORB_ERR_OVERFLOW;
    }
    return GFORB_OK;
}

}  // namespace gforb
