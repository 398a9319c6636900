/*
 * orbx.h -- C ABI of the B200-native ORB front-end (liborbx.so).
 *
 * Drop-in boundary for the hot path of geoeo/ORB_SLAM3_V1.0.  The
 * reference has no FFI layer: the path is C++ classes inside libORB_SLAM3.so, so every
 * entry point below names the reference interface it stands behind (file:line under
 * /root/reference).  The C++ shim in include/orbx/ re-creates those class interfaces on
 * top of this ABI; INTEGRATION.md shows the binding a maintainer would add.
 *
 * Conventions: plain pointers and sizes only; every function returns an orbx_status
 * (never exits, never throws); device work is enqueued on the handle's stream; there
 * is NO CPU fallback -- without a CUDA device every call fails with ORBX_ERR_CUDA.
 */
#ifndef ORBX_H
#define ORBX_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    ORBX_OK = 0,
    ORBX_ERR_INVALID = 1,   /* bad argument / unsupported size */
    ORBX_ERR_CUDA = 2,      /* CUDA runtime error or no device; see orbx_last_error() */
    ORBX_ERR_CAPACITY = 3,  /* caller buffer too small */
    ORBX_ERR_EMPTY = 4      /* empty image (reference operator() returns -1) */
} orbx_status;

const char* orbx_last_error(void);
int orbx_version(void);

/* cv::KeyPoint memory layout (28 bytes): what ORBextractor::operator() fills
 * (upstream form; the fork's 24-byte ORB_SLAM3::KeyPoint is include/KeyPoint.h:7-18). */
typedef struct {
    float x, y;        /* pt, level-0 pixel coordinates */
    float size;        /* PATCH_SIZE * mvScaleFactor[octave], truncated to int */
    float angle;       /* degrees [0,360), cv::fastAtan2 of the intensity centroid */
    float response;    /* FAST score */
    int32_t octave;
    int32_t class_id;  /* -1 */
} orbx_keypoint;

/* ---------------------------------------------------------------- extractor
 * ORBextractor::ORBextractor        src/ORBextractor.cc:82-149   (5-arg upstream form)
 * ORBextractor::operator() / extractFeatures   src/ORBextractor.cc:543-585
 * ComputePyramid :607-623, ComputeKeyPointsOctTree :433-541, DistributeOctTree :226-431,
 * IC_Angle src/cuda/Angle_gpu.cu:26-80, computeOrbDescriptor src/cuda/Orb_gpu.cu:311-350 */
typedef struct {
    int nfeatures;
    float scale_factor;
    int nlevels;
    int ini_th_fast;
    int min_th_fast;
    int width, height;   /* image size this handle is built for */
    int max_batch;       /* frames per call (device buffers are sized for this) */
    int device;          /* CUDA ordinal */
} orbx_params;

typedef struct orbx_extractor orbx_extractor;

orbx_status orbx_extractor_create(const orbx_params* p, orbx_extractor** out);
void orbx_extractor_destroy(orbx_extractor* e);

/* mvScaleFactor / mvInvScaleFactor / mvLevelSigma2 / mvInvLevelSigma2 /
 * mnFeaturesPerLevel / umax  (src/ORBextractor.cc:92-143). Any pointer may be NULL. */
orbx_status orbx_extractor_tables(const orbx_extractor* e, float* scale, float* inv_scale,
                                  float* sigma2, float* inv_sigma2, int* feat_per_level,
                                  int* umax16);
/* capacity (keypoints per frame) a caller buffer needs for orbx_extractor_fetch */
int orbx_extractor_capacity(const orbx_extractor* e);

/* Run the extractor on `batch` frames that are ALREADY in device memory
 * (u8, row-major; frame f starts at d_images + f*frame_stride, rows row_stride apart).
 * lapping: host array of 2*batch ints {x0,x1} per frame (vLappingArea), or NULL = {0,0}.
 * Asynchronous: returns after enqueueing on the handle's stream. */
orbx_status orbx_extract_device(orbx_extractor* e, const uint8_t* d_images, size_t frame_stride,
                                int row_stride, int batch, const int* lapping);
/* Same, from host memory (pinned memory makes the H2D copy asynchronous). */
orbx_status orbx_extract_host(orbx_extractor* e, const uint8_t* h_images, size_t frame_stride,
                              int row_stride, int batch, const int* lapping);
/* Enqueue the device->pinned-host copy of counts + keypoints + descriptors of the last
 * extract and wait for it.  After this, orbx_extractor_fetch is a plain memcpy. */
orbx_status orbx_extractor_sync(orbx_extractor* e);
/* Copy frame `f` of the last batch to caller buffers (cap keypoints).  *n = number of
 * keypoints, *mono_index = return value of the reference operator(). */
orbx_status orbx_extractor_fetch(orbx_extractor* e, int f, orbx_keypoint* kps, uint8_t* desc,
                                 int cap, int* n, int* mono_index);
/* Per-frame counts only (host arrays of `batch` ints), after orbx_extractor_sync. */
orbx_status orbx_extractor_counts(orbx_extractor* e, int* n, int* mono_index);

/* mvImagePyramid[level] of frame f: size query and download of the bordered
 * (w+38)x(h+38) buffer, densely packed (upstream layout: ROI at (19,19)). */
orbx_status orbx_extractor_level_size(const orbx_extractor* e, int level, int* w, int* h);
orbx_status orbx_extractor_level_download(orbx_extractor* e, int f, int level, uint8_t* out);
/* 7x7 sigma-2 blurred level (w x h, no border), for tests. */
orbx_status orbx_extractor_blurred_download(orbx_extractor* e, int f, int level, uint8_t* out);
/* FAST candidates handed to DistributeOctTree for (frame, level): coordinates relative to
 * minBorder (16,16), reference order (cell raster, then raster inside the cell). */
orbx_status orbx_extractor_candidates(orbx_extractor* e, int f, int level, int* xs, int* ys,
                                      int* resp, int cap, int* n);

/* the stream all work of this handle is enqueued on (a cudaStream_t) */
void* orbx_extractor_stream(orbx_extractor* e);
/* number of kernel launches issued by the last orbx_extract_* call */
int orbx_extractor_last_launches(const orbx_extractor* e);
/* device time of each stage of the last extract (ms), measured with CUDA events when
 * profiling was enabled with orbx_extractor_set_profiling(e, 1). stages: 0 pyramid,
 * 1 fast, 2 octree, 3 order, 4 blur, 5 orient+describe.  */
orbx_status orbx_extractor_set_profiling(orbx_extractor* e, int on);
orbx_status orbx_extractor_stage_ms(orbx_extractor* e, float* ms6);

/* pinned host memory helpers for callers without a CUDA runtime of their own */
orbx_status orbx_host_alloc(void** p, size_t bytes);
void orbx_host_free(void* p);
orbx_status orbx_device_alloc(int device, void** p, size_t bytes);
void orbx_device_free(void* p);
orbx_status orbx_memcpy_h2d(void* d, const void* h, size_t bytes);
orbx_status orbx_memcpy_d2h(void* h, const void* d, size_t bytes);
orbx_status orbx_device_synchronize(void);

/* ---------------------------------------------------------------- matcher
 * ORBmatcher::DescriptorDistance   src/ORBmatcher.cc:1375-1391
 * n pairs of 32-byte descriptors (host pointers) -> n ints. */
orbx_status orbx_hamming_pairs(const uint8_t* a, const uint8_t* b, int n, int* out);

#ifdef __cplusplus
}
#endif
#endif
