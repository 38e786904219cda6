// Host-side engine: geometry tables (ORBextractor constructor arithmetic), device arenas, stage launches.
#pragma once
#include "gforb_internal.cuh"

namespace gforb {

// thread-local error text behind gforb_last_error_string()
void set_error(const std::string& msg);
const char* last_error();
int cuda_fail(cudaError_t e, const char* what);   // records the message, returns GFORB_ERR_CUDA

// ORBextractor::ORBextractor tables (reference src/ORBextractor.cc:409-469)
struct Tables {
    gforb_params prm;
    std::vector<float> scale, inv_scale, sigma2, inv_sigma2;
    std::vector<int> quota;
    int umax[16];
};
int make_tables(const gforb_params& p, Tables& t);

// Fill a Geom for (W, H, nimg); host_tabs receives the resize tables to upload.  GFORB_ERR_ARG if the image
// cannot carry the configured pyramid.
int make_geometry(const Tables& t, const gforb_limits& lim, int W, int H, int nimg, Geom& g,
                  std::vector<int32_t>& host_tabs, size_t& pyr_bytes, size_t& blur_bytes);

class Engine {
public:
    Engine() = default;
    ~Engine();
    Engine(const Engine&) = delete;
    Engine& operator=(const Engine&) = delete;

    int init(const Tables& t, const gforb_limits& lim, int W, int H, int nimg, int nframes_stereo);
    void release();

    int upload_images(const uint8_t* host, size_t pitch, int first_img, int step, int count, cudaStream_t s);
    // pyramid only / full extraction for images [0, nimg)
    int run_pyramid(int nimg, cudaStream_t s);
    // images [img0, img0 + nimg); 6 stage events when given
    int run_extract(int nimg, cudaStream_t s, cudaEvent_t* stage_events = nullptr, int img0 = 0, bool fork_blur = false);
    // read + clear the sticky device error word (synchronises `s`)
    int check_errors(cudaStream_t s);

    Geom g{};
    TileMaps maps{};
    CUtensorMap rsmaps[kMaxLevels]{};   // resize source tiles (k_pyramid.cu)
    Buffers b{};
    int nframes_cap = 0;
    bool ready = false;
    bool pyramid_valid = false;
    bool full_frames = true;               // write all 19 px of every level's frame (handles that expose the padded levels); the
                                           // batch path clears it: its kernels read the ROI and 3 px of frame (k_pyramid.cu)
    cudaStream_t s_aux = nullptr;          // side stream of run_extract(fork_blur)
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    int32_t* h_small = nullptr;   // pinned: counts / flags read-back
    size_t h_small_ints = 0;
};

}  // namespace gforb
