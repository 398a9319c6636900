/* Extractor state shared by the liborbx translation units (extractor.cu owns it; the stereo
 * matcher in matcher.cu reads the device-resident pyramids / keypoints / descriptors).
 * Not part of the ABI. */
#ifndef ORBX_EXTRACTOR_STATE_H
#define ORBX_EXTRACTOR_STATE_H
#include <vector>
#include "orbx_internal.h"

/* ------------------------------------------------------------------ device-side level table */
struct LevelDev {
    int w, h, pitch, rows;
    unsigned long long frame_bytes;
    uint8_t* base;          /* bordered pyramid, frame-major */
    int bpitch;
    unsigned long long bframe_bytes;
    uint8_t* blur;          /* blurred level, w x h */
    const short4* xtab;     /* resize: per dx {sx, a0, a1, 0} */
    const short4* ytab;     /* resize: per dy {sy, b0, b1, 0} */
    const uint4* xcoef;     /* word-wise resize tables over the padded row (k_resize_w4) */
    const uint32_t* xmeta;
    const uint32_t* xsel;   /* per word: 4 PRMT selectors (source byte offsets off, off+1 of each pixel) */
    const short4* ypad;     /* per padded row {sy, b0, b1, 0} */
    int fast_resize;
    int band_smem;          /* k_resize_band: bytes of the tallest band's source rows */
    int cell_begin, cell_end;
    int fast_box_rows;      /* tallest FAST strip of the level = rows of its TMA box */
    int N, nIni;
    float hX;
    int cand_off, cand_cap; /* into the per-frame candidate arrays */
    int sel_off, sel_cap;   /* into the per-frame selected-keypoint arrays */
    float scale;
    float patch_size;
};

struct Cell {
    short level, x0, y0, cw, ch, offx, offy, pad;
    int slot_off;           /* into the per-frame candidate-slot array */
    unsigned rcp_dw;        /* floor(2^32/(cw-6))+1: exact i/(cw-6) by __umulhi for i < 2^16 */
};

#define ORBX_MAX_LANES 8
struct Tables {
    int nlevels;
    int l0;                 /* first level of a quadtree launch: level = blockIdx.x + l0 (per-level launches of the latency path) */
    int f0;                 /* first frame of the launch: frame = blockIdx + f0 (a batch runs as chunks on several streams) */
    LevelDev lv[ORBX_MAX_LEVELS];
};

__device__ __forceinline__ int reflect101(int p, int n) {
    /* |p| excursions are < n here (border 19 < smallest supported level) but loop anyway */
    while (p < 0 || p >= n) p = (p < 0) ? -p : 2 * (n - 1) - p;
    return p;
}

__device__ __forceinline__ uint8_t* lv_ptr(const LevelDev& L, int f, int x, int y) {
    return L.base + (size_t)f * L.frame_bytes + (size_t)(y + ORBX_EDGE) * L.pitch + ORBX_XOFF + x;
}

struct orbx_extractor {
    orbx_params P;
    int nlevels;
    std::vector<float> sf, inv_sf, sigma2, inv_sigma2;
    std::vector<int> nfeat, umax;
    Tables T;
    std::vector<LevelDev> lv;
    std::vector<Cell> cells;
    int ncells = 0, max_cells_level = 0;
    size_t slots_per_frame = 0, cand_per_frame = 0, sel_per_frame = 0;
    int kcap = 0, maxn = 0;
    size_t oct_smem = 0;
    int nblur_tiles = 0;
    int fast_tp = 8, fast_rows = 8, fast_list_cap = 2;   /* tile pitch of the short-strip launch group (of every unit when there is one group) */
    int fast_tp_tall = 8;                                /* tile pitch of the tall-strip launch group */
    int nunits = 0;
    int nunits_short = 0, fast_rows_short = 8;   /* units [0, nunits_short) are launched with the smaller tile */
    int fast_cl_cap = 1024;         /* FAST corner-list entries in use (<= FAST_CL) */
    void* d_units = nullptr;        /* FAST units (extractor.cu: struct Unit) */
    /* device-side Frame grid of the last batch (matcher.cu: orbx_extractor_assign_grid) */
    int *d_grid_off = nullptr, *d_grid_idx = nullptr;
    void *d_grid_probs = nullptr, *h_grid_probs = nullptr;
    int grid_cells = 0, grid_cells_cap = 0;
    void* d_odmaps = nullptr;       /* TMA tensor maps of the describe stage (extractor.cu: struct ODMaps) */
    cudaStream_t stream = nullptr;
    cudaStream_t side = nullptr;    /* the blur runs here beside FAST / quadtree / order (it needs the pyramid only) */
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_join2 = nullptr;
    /* CUDA graph of the launch chain of a small batch (one stream + side stream), replayed while the
     * input pointer / layout / batch / lapping stay the same (ORBX_GRAPH=0 disables) */
    bool use_graph = true;
    bool dataflow = true;           /* per-level branches inside the graph (ORBX_DATAFLOW=0: one chain + side stream) */
    /* per-level dataflow of the latency path: branch stream l runs FAST + quadtree of level l as soon as
     * level l of the pyramid exists */
    cudaStream_t branch[ORBX_MAX_LEVELS] = {};
    cudaEvent_t ev_lvl[ORBX_MAX_LEVELS] = {}, ev_oct[ORBX_MAX_LEVELS] = {};
    int unit_lo[ORBX_MAX_LEVELS] = {}, unit_hi[ORBX_MAX_LEVELS] = {};   /* FAST units of a level (contiguous after the sort) */
    cudaGraphExec_t graph_exec = nullptr;
    const uint8_t* g_img = nullptr;
    size_t g_fs = 0;
    int g_rs = 0, g_batch = 0, g_launches = 0;
    bool g_lap = false;
    const uint8_t* s_img = nullptr;  /* key of the previous call */
    size_t s_fs = 0;
    int s_rs = 0, s_batch = 0;
    bool s_lap = false;
    int od_pad = 0;                 /* extra dynamic shared memory of the describe kernel: caps its resident CTAs per SM (tuning, ORBX_OD_PAD) */
    int lanes = 2;                  /* a batch of >= 32 frames runs as up to `lanes` chunks on as many streams (ORBX_LANES); 2 and 4 give the same device-resident rate, 2 the better end-to-end rate */
    cudaStream_t lane[ORBX_MAX_LANES] = {};
    cudaEvent_t ev_lane[ORBX_MAX_LANES] = {};
    /* device buffers */
    uint8_t* d_pyr = nullptr;
    uint8_t* d_blur = nullptr;
    uint8_t* d_input = nullptr;      /* staging for orbx_extract_host */
    short4* d_tabs = nullptr;
    uint8_t* d_wtabs = nullptr;
    Cell* d_cells = nullptr;
    int4* d_blur_tiles = nullptr;
    uint32_t* d_slots = nullptr;
    int* d_cell_count = nullptr;
    uint32_t* d_cand = nullptr;
    uint32_t* d_pnode = nullptr;
    int* d_cand_count = nullptr;
    uint32_t* d_sel = nullptr;
    int* d_sel_count = nullptr;
    int* d_dest = nullptr;
    int* d_lapping = nullptr;
    orbx_keypoint* d_kps = nullptr;
    uint8_t* d_desc = nullptr;
    int* d_nk = nullptr;           /* [2*max_batch]: counts then mono indices */
    /* pinned host mirrors */
    orbx_keypoint* h_kps = nullptr;
    uint8_t* h_desc = nullptr;
    int* h_nk = nullptr;
    int* h_lapping = nullptr;
    int last_batch = 0, last_launches = 0;
    bool synced = false;
    bool profiling = false;
    int fork_blur = 2;             /* blur on the side stream: 2 = beside quadtree + order, 1 = also beside FAST, 0 = one stream (ORBX_FORK_BLUR) */
    cudaEvent_t ev[7] = {};
    /* Frame::ComputeStereoMatches state (stereo.cu), allocated on first use */
    float* d_uright = nullptr;     /* [max_batch * kcap] mvuRight of the left frames */
    float* d_depth = nullptr;
    float* h_uright = nullptr;
    float* h_depth = nullptr;
    int* d_pairs = nullptr;        /* [2*max_batch] left / right frame index per pair */
    int* h_pairs = nullptr;
    unsigned short* d_rowent = nullptr;   /* per pair: right-keypoint indices per image row (vRowIndices) */
    size_t rowent_per_pair = 0;
    uint32_t* d_rowstart = nullptr;       /* per pair: CSR offsets of d_rowent per image row */
    int* d_sad = nullptr;                 /* per pair: accepted bestSAD per left keypoint or -1 */
    int* d_stdone = nullptr;              /* per pair: matching CTAs finished (the last one applies the median cut) */
    cudaEvent_t ev_stereo = nullptr;
    int last_pairs = 0;
};

#endif
