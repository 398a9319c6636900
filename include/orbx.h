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
 * Threads: handles are independent; any number of host threads may create, use and destroy
 * their own handles concurrently (the reference runs the left / right extractors and the
 * Tracking / LocalMapping / LoopClosing matchers on separate threads).  ONE handle is not
 * re-entrant: one calling thread at a time.  orbx_last_error() is per thread.  A handle
 * returned by a create function is ready: its tables are in device memory.
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
/* The same level IN PLACE on the device, for callers that chain their own kernels behind the extraction (the
 * reference's mvImagePyramid[level] is a cv::cuda::HostMem a kernel can read, include/ORBextractor.h:94): *d_roi = ROI
 * pixel (0,0) of frame 0, rows `pitch` bytes apart, frame f at + f * frame_stride; the 19-px reflect-101 border lies
 * around the ROI (d_roi[-19 * pitch - 19] is valid).  Contents belong to the last extract enqueued on the handle's
 * stream: order readers after orbx_extractor_stream(e).  Any output pointer may be NULL. */
orbx_status orbx_extractor_level_device(orbx_extractor* e, int level, const uint8_t** d_roi, int* w, int* h, int* pitch,
                                        size_t* frame_stride);
/* Keypoints, descriptors and counts of the last extract in place on the device: frame f's records start at
 * d_kps + f * capacity and d_desc + 32 * f * capacity, already in their final vLappingArea order; d_counts[f] = number of
 * keypoints, d_counts[max_batch + f] = monoIndex.  Same ordering rule as above. */
orbx_status orbx_extractor_results_device(orbx_extractor* e, const orbx_keypoint** d_kps, const uint8_t** d_desc,
                                          const int** d_counts, int* capacity);
/* 7x7 sigma-2 blurred level (w x h, no border), for tests. */
orbx_status orbx_extractor_blurred_download(orbx_extractor* e, int f, int level, uint8_t* out);
/* FAST candidates handed to DistributeOctTree for (frame, level): coordinates relative to
 * minBorder (16,16), reference order (cell raster, then raster inside the cell). */
orbx_status orbx_extractor_candidates(orbx_extractor* e, int f, int level, int* xs, int* ys,
                                      int* resp, int cap, int* n);

/* the stream all work of this handle is enqueued on (a cudaStream_t) */
void* orbx_extractor_stream(orbx_extractor* e);
/* internal streams a batch of `batch` frames is spread over (chunks of frames run concurrently;
 * 1 for small batches and in profiling mode).  Everything is ordered with the handle's stream. */
int orbx_extractor_streams(const orbx_extractor* e, int batch);
/* number of kernel launches issued by the last orbx_extract_* call */
int orbx_extractor_last_launches(const orbx_extractor* e);
/* device time of each stage of the last extract (ms), measured with CUDA events when
 * profiling was enabled with orbx_extractor_set_profiling(e, 1). stages: 0 pyramid,
 * 1 fast, 2 octree, 3 order, 4 blur, 5 orient+describe.  */
orbx_status orbx_extractor_set_profiling(orbx_extractor* e, int on);
/* Tuning / test knob: capacity (entries, 1..1024, default 1024 = all of it) of the per-region corner list the
 * FAST kernel keeps for its non-maximum-suppression pass; regions with more corners take the
 * score-map scan path.  Results do not depend on it. */
orbx_status orbx_extractor_set_fast_corner_list(orbx_extractor* e, int entries);
orbx_status orbx_extractor_stage_ms(orbx_extractor* e, float* ms6);

/* pinned host memory helpers for callers without a CUDA runtime of their own */
orbx_status orbx_host_alloc(void** p, size_t bytes);
/* page-locked, write-combined: for staging buffers the host only WRITES (the image grabber's frame
 * queue); CPU reads of such memory are slow */
orbx_status orbx_host_alloc_wc(void** p, size_t bytes);
void orbx_host_free(void* p);
orbx_status orbx_device_alloc(int device, void** p, size_t bytes);
void orbx_device_free(void* p);
orbx_status orbx_memcpy_h2d(void* d, const void* h, size_t bytes);
orbx_status orbx_memcpy_d2h(void* h, const void* d, size_t bytes);
orbx_status orbx_device_synchronize(void);

/* ---------------------------------------------------------------- matcher
 * All matcher entry points take HOST pointers, are batched over `nprob` independent problems
 * (one CTA each; nprob = 1 is the reference's single call) and return after the results are
 * in the caller's buffers.  A matcher handle owns one stream and its staging buffers. */
typedef struct orbx_matcher orbx_matcher;
orbx_status orbx_matcher_create(int device, orbx_matcher** out);
void orbx_matcher_destroy(orbx_matcher* m);
void* orbx_matcher_stream(orbx_matcher* m);
/* kernel launches issued by the last matcher call */
int orbx_matcher_last_launches(const orbx_matcher* m);

/* ORBmatcher::DescriptorDistance   src/ORBmatcher.cc:1375-1391
 * n pairs of 32-byte descriptors (host pointers) -> n ints. */
orbx_status orbx_hamming_pairs(const uint8_t* a, const uint8_t* b, int n, int* out);

/* ORBmatcher::ComputeThreeMaxima   src/ORBmatcher.cc:1328-1370, on bin counts (host; the
 * device kernels run the same code inside their rotation-histogram step). */
orbx_status orbx_three_maxima(const int* counts, int L, int* ind1, int* ind2, int* ind3);

/* The part of Frame the projection searches read (src/Frame.cc:157-176 AssignFeaturesToGrid,
 * :404-468 GetFeaturesInArea, :470-480 PosInGrid), flattened.  The grid itself is rebuilt on
 * the device from (kx, ky). */
typedef struct {
    int n;                          /* mNumKeypoints */
    const float* kx; const float* ky; const int32_t* koct;   /* mvKeysUn: pt.x, pt.y, octave */
    const float* kangle;            /* mvKeysUn angle; may be NULL unless checkOrientation */
    const uint8_t* desc;            /* mDescriptors, n x 32 */
    float minx, miny;               /* mnMinX, mnMinY */
    float gw_inv, gh_inv;           /* mfGridElementWidthInv / HeightInv */
    int gcols, grows;               /* mFrameGridCols / Rows */
    const float* scale;             /* mvScaleFactors */
    int nlevels;
    const float* uright;            /* mvuRight of a stereo / RGB-D frame (F->Nleft == -1), e.g. the output of
                                     * orbx_stereo_matches; NULL = monocular (all -1).  Read by
                                     * orbx_search_by_projection only (src/ORBmatcher.cc:81-86). */
} orbx_frame_view;

/* MapPoint fields SearchByProjection reads (src/ORBmatcher.cc:37-70) */
typedef struct {
    int n;
    const uint8_t* inview;          /* mbTrackInView && !isBad() */
    const float* x; const float* y; /* mTrackProjX / mTrackProjY */
    const int32_t* level;           /* mnTrackScaleLevel */
    const float* viewcos;           /* mTrackViewCos */
    const float* depth;             /* mTrackDepth (bFarPoints gate) */
    const uint8_t* desc;            /* GetDescriptor(), n x 32 */
    const float* proj_xr;           /* mTrackProjXR (src/Frame.cc:325); required when the frame view has uright */
    const uint8_t* has_obs;         /* Observations() > 0 per map point; NULL = all.  Only such a point blocks the
                                     * keypoint it was assigned to for later points (src/ORBmatcher.cc:77-79) */
} orbx_mappoints;

/* Map points handed to Fuse, after the host-side projection of src/ORBmatcher.cc:702-760 (:880-917
 * for the Sim3 overload): valid[i] = 0 for points the reference skips before the search (null,
 * bad, already in the KeyFrame, behind the camera, outside the image or the distance range). */
typedef struct {
    int32_t n;
    const uint8_t* valid;
    const float* u; const float* v;     /* projection uv */
    const float* ur;                    /* u - bf/z (first overload; may be NULL = no stereo gate) */
    const int32_t* level;               /* pMP->PredictScale(dist3D, pKF) */
    const uint8_t* desc;                /* pMP->GetDescriptor(), n x 32 */
} orbx_fuse_points;

/* Search core of ORBmatcher::Fuse(pKF, vpMapPoints, th, bRight)   src/ORBmatcher.cc:678-862
 * (variant 0) and Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)   :864-975 (variant 1), nprob
 * KeyFrames per call.  KF = the KeyFrame's mvKeysUn/descriptors/grid/mvScaleFactors; uright[p] =
 * mvuRight (NULL: monocular), inv_sigma2[p] = mvInvLevelSigma2 (variant 0).  best_idx[p][i] =
 * the keypoint the reference selects for map point i if bestDist <= TH_LOW, else -1; best_dist
 * (may be NULL) its distance.  The graph updates (:826-848 / :958-968: Replace / AddObservation /
 * AddMapPoint, nFused++) are applied by the caller in map-point order. */
orbx_status orbx_fuse_search(orbx_matcher* m, int nprob, const orbx_frame_view* KF, const float* const* uright,
                             const float* const* inv_sigma2, const orbx_fuse_points* MP, float th, int variant,
                             int* const* best_idx, int* const* best_dist);

/* ORBmatcher::SearchBySim3(pKF1, pKF2, vpMatches12, S12, th)   src/ORBmatcher.cc:977-1200.
 * MP1in2[p] holds one entry per KeyFrame-1 keypoint: its map point projected into KeyFrame 2 with
 * S21 (valid = 0 if there is none, it is bad or already matched, or it fails the checks of
 * :1022-1050); MP2in1[p] the converse (:1084-1113).  match12[p][i1] = keypoint of KeyFrame 2 whose
 * map point the reference stores in vpMatches12[i1] (both directed searches agree), else -1;
 * nfound[p] = the return value. */
orbx_status orbx_search_by_sim3(orbx_matcher* m, int nprob, const orbx_frame_view* KF1, const orbx_frame_view* KF2,
                                const orbx_fuse_points* MP1in2, const orbx_fuse_points* MP2in1, float th,
                                int* const* match12, int* nfound);

/* ORBmatcher::SearchForInitialization(F1, F2, windowSize, nnRatio, checkOrientation)
 * src/ORBmatcher.cc:329-439, nprob frame pairs per call.  Of F1 the keypoints (mvKeysUn x, y,
 * octave, angle) and descriptors are read, of F2 also the grid.  matches12[p] (F1[p].n entries) =
 * vnMatches12, nmatches[p] = the returned count. */
orbx_status orbx_search_for_initialization(orbx_matcher* m, int nprob, const orbx_frame_view* F1, const orbx_frame_view* F2,
                                           int windowSize, float nnRatio, int checkOrientation,
                                           int* const* matches12, int* nmatches);

/* MapPoint::ComputeDistinctiveDescriptors   src/MapPoint.cc:343-412, batched over nsets map points:
 * desc rows set_off[s] .. set_off[s+1] are the observation descriptors of map point s (set_off[0]
 * = 0); best[s] = row (within the set) with the least median Hamming distance to the set, first
 * on ties, -1 for an empty set; median (may be NULL) receives that median. */
orbx_status orbx_distinctive_descriptors(orbx_matcher* m, int nsets, const int* set_off, const uint8_t* desc,
                                         int* best, int* median);

/* DBoW2 FORB::meanValue   Thirdparty/DBoW2/src/DBoW2/FORB.cpp:28-79, batched over nsets descriptor sets (same CSR
 * layout as above): mean[s*32 .. s*32+32) = the per-bit majority descriptor of set s (bit set iff at least
 * N/2 + N%2 of its N descriptors have it; N == 1: the descriptor itself; N == 0: 32 zero bytes, where the reference
 * releases the Mat).  (FORB::distance, FORB.cpp:81-101, is orbx_hamming_pairs.) */
orbx_status orbx_mean_value(orbx_matcher* m, int nsets, const int* set_off, const uint8_t* desc, uint8_t* mean);

/* Frame::AssignFeaturesToGrid + PosInGrid   src/Frame.cc:157-176, 470-480, for nframes frames:
 * cell_off[p][gcols*grows + 1] / cell_idx[p][F[p].n] = CSR over the cells in the reference's linear
 * order (y*cols + x), keypoints in push_back order inside a cell; the unused tail of cell_idx
 * (keypoints outside the grid) is -1.  Only n, kx, ky and the grid fields of F are read. */
orbx_status orbx_assign_features_to_grid(orbx_matcher* m, int nframes, const orbx_frame_view* F,
                                         int* const* cell_off, int* const* cell_idx);
/* Frame::GetFeaturesInArea   src/Frame.cc:404-468, nq queries against one frame: out_count[q] =
 * vIndices.size(), out_idx[q*cap .. ) = the first min(count, cap) indices in the reference's order. */
orbx_status orbx_features_in_area(orbx_matcher* m, const orbx_frame_view* F, int nq, const float* x, const float* y,
                                  const float* r, const int* minLevel, const int* maxLevel, int* out_count,
                                  int* out_idx, int cap);
/* The same grid for every frame of an extractor's last batch, built on the device from its
 * resident keypoints on its stream (mvKeysUn == mvKeys: rectified / distortion-free cameras), and
 * its read-back (cell_idx receives min(idx_cap, capacity) entries). */
orbx_status orbx_extractor_assign_grid(orbx_extractor* e, float minx, float miny, float gw_inv, float gh_inv,
                                       int gcols, int grows);
orbx_status orbx_extractor_grid_fetch(orbx_extractor* e, int f, int* cell_off, int* cell_idx, int idx_cap);

/* ORBmatcher::SearchByProjection(Frame, vpMapPoints, th, bFarPoints, thFarPoints, nnRatio,
 * checkOrientation)   src/ORBmatcher.cc:31-123.
 * claimed[p][i] != 0: frame keypoint i already holds a MapPoint with observations (:77-79);
 * NULL or claimed[p] == NULL = none.  match_mp[p][i] = index of the map point frame keypoint i
 * holds after the call or -1; nmatches[p] = the function's return value (every assignment
 * counts, also one that a later map point overwrites).  The reference's sequential rule -- a
 * keypoint assigned to a map point WITH observations is skipped by every later map point, one
 * assigned to a point without observations can be taken over (:77-79, MP.has_obs) -- and the
 * stereo gate |mTrackProjXR - mvuRight[idx]| <= r for keypoints with mvuRight > 0 (:81-86,
 * F.uright + MP.proj_xr) are reproduced exactly. */
orbx_status orbx_search_by_projection(orbx_matcher* m, int nprob, const orbx_frame_view* F,
                                      const orbx_mappoints* MP, const uint8_t* const* claimed,
                                      float th, int bFarPoints, float thFarPoints, float nnRatio,
                                      int* const* match_mp, int* nmatches);

/* KeyFrame map points already projected into the frame by the host shim
 * (src/ORBmatcher.cc:1225-1262: pose, frustum and distance gates, PredictScale). */
typedef struct {
    int n;
    const uint8_t* valid;           /* passed the gates, not in sAlreadyFound */
    const float* u; const float* v; /* projection */
    const int32_t* level;           /* nPredictedLevel */
    const float* angle;             /* pKF->mvKeysUn[i].angle */
    const uint8_t* desc;
} orbx_kf_points;

/* ORBmatcher::SearchByProjection(Frame, KeyFrame, sAlreadyFound, th, checkOrientation)
 * src/ORBmatcher.cc:1202-1326 (rotation histogram :1290-1322). */
orbx_status orbx_search_by_projection_kf(orbx_matcher* m, int nprob, const orbx_frame_view* F,
                                         const orbx_kf_points* MP, const uint8_t* const* occupied,
                                         float th, int checkOrientation, int* const* match_mp,
                                         int* nmatches);

/* Device-resident forms of the two searches above, for the frames an extractor has just processed (no host round trip
 * of the Frame side, no host synchronisation between extraction and search): problem p searches frame frames[p] of
 * e's last batch -- mvKeysUn == the extracted keypoints (rectified / distortion-free camera), mDescriptors and
 * mNumKeypoints are read in place from e's device buffers; stereo_pairs[p] >= 0 (may be NULL = mono) names the pair of
 * e's last orbx_stereo_matches_async call whose mvuRight the stereo gate of :81-86 reads.  Of `grid` only the grid
 * constants (minx .. grows), scale and nlevels are read.  claimed[p] / occupied[p] / match_mp[p] hold
 * orbx_extractor_capacity(e) entries (entries beyond the frame's keypoint count come back as -1).  The matcher's
 * stream waits for e's stream through an event; the call returns with the results on the host. */
orbx_status orbx_search_by_projection_resident(orbx_matcher* m, orbx_extractor* e, int nprob, const int* frames,
                                               const int* stereo_pairs, const orbx_frame_view* grid,
                                               const orbx_mappoints* MP, const uint8_t* const* claimed, float th,
                                               int bFarPoints, float thFarPoints, float nnRatio, int* const* match_mp,
                                               int* nmatches);
orbx_status orbx_search_by_projection_kf_resident(orbx_matcher* m, orbx_extractor* e, int nprob, const int* frames,
                                                  const orbx_frame_view* grid, const orbx_kf_points* MP,
                                                  const uint8_t* const* occupied, float th, int checkOrientation,
                                                  int* const* match_mp, int* nmatches);

/* KeyFrame fields SearchForTriangulation reads (src/ORBmatcher.cc:441-676); the DBoW2
 * FeatureVector is flattened to sorted node ids + CSR index lists. */
typedef struct {
    int n;
    const float* kx; const float* ky; const int32_t* koct; const float* kangle;
    const uint8_t* desc;
    const uint8_t* has_mp;          /* GetMapPoint(idx) != nullptr; may be NULL */
    const float* uright;            /* mvuRight; may be NULL (mono) */
    int nnodes; const int32_t* node_id; const int32_t* node_off; const int32_t* node_idx;
} orbx_kf_view;

/* ORBmatcher::SearchForTriangulation   src/ORBmatcher.cc:441-676.  F12 (row-major 3x3, as
 * Pinhole::epipolarConstrain builds it, src/CameraModels/Pinhole.cpp:104-110) and the epipole
 * `ep` (:453-456) are computed once by the host shim: 9 / 2 floats per problem.  scale2[p] =
 * pKF2->mvScaleFactors.  pairs[p] receives 2*npairs[p] ints (idx1, idx2) in idx1 order
 * (capacity 2*k1[p].n); nmatches[p] = the function's return value. */
orbx_status orbx_search_for_triangulation(orbx_matcher* m, int nprob, const orbx_kf_view* k1,
                                          const orbx_kf_view* k2, const float* F12,
                                          const float* ep, const float* const* scale2,
                                          int bOnlyStereo, int bCoarse, int checkOrientation,
                                          int* const* pairs, int* npairs, int* nmatches);

/* ORBmatcher::SearchByBoW(pKF, F, vpMapPointMatches, nnRatio, checkOrientation)
 * src/ORBmatcher.cc:133-327, nprob (KeyFrame, Frame) problems per call.  Both sides are flattened
 * as orbx_kf_view (of F only n, kangle, desc and the feature vector are read; KF.has_mp[i] =
 * pMP && !pMP->isBad()).  nleft[p] = F->Nleft (NULL or -1: one camera; otherwise keypoints >=
 * Nleft belong to the right camera and take the fork's second branch).  match_kf[p][i] = KeyFrame
 * keypoint whose MapPoint the reference stores in vpMapPointMatches[i], or -1. */
orbx_status orbx_search_by_bow(orbx_matcher* m, int nprob, const orbx_kf_view* KF, const orbx_kf_view* F,
                               const int* nleft, float nnRatio, int checkOrientation, int* const* match_kf,
                               int* nmatches);


/* Frame::ComputeStereoMatches (upstream ORB-SLAM3 v1.0; absent from the fork:
 * include/Frame.h:53-62).  Works on the DEVICE-RESIDENT results of the
 * last orbx_extract_* call of each handle (keypoints, descriptors, mvImagePyramid): pair p
 * matches frame left_frames[p] of `left` against frame right_frames[p] of `right` (`left` and
 * `right` may be the same handle).  uRight/depth: npairs x stride floats (stride >=
 * orbx_extractor_capacity(left)), entry i = mvuRight[i] / mvDepth[i] of the left keypoint i. */
orbx_status orbx_stereo_matches(orbx_extractor* left, orbx_extractor* right, int npairs,
                                const int* left_frames, const int* right_frames, float bf, float b,
                                float* uRight, float* depth, int stride);
/* Same, asynchronous and device-resident: enqueues on left's stream (after right's work) and
 * leaves mvuRight / mvDepth in the handle; orbx_stereo_fetch copies them out. */
orbx_status orbx_stereo_matches_async(orbx_extractor* left, orbx_extractor* right, int npairs,
                                      const int* left_frames, const int* right_frames, float bf,
                                      float b);
orbx_status orbx_stereo_fetch(orbx_extractor* left, int npairs, float* uRight, float* depth, int stride);

/* Brute-force 2-nearest-neighbour Hamming search + ratio test (BASELINE.json config 5; the
 * fork only declares an unused cv::BFMatcher, src/Frame.cc:45).  Ties go to the lowest
 * database index.  Host pointers; d1/i1/d2/i2: nq ints (257 / -1 when the database has fewer
 * than 1 / 2 rows); pass (may be NULL): d1 < ratio * d2. */
orbx_status orbx_knn2(orbx_matcher* m, const uint8_t* q, int nq, const uint8_t* db, int ndb,
                      float ratio, int* d1, int* i1, int* d2, int* i2, uint8_t* pass);
/* Device-resident building blocks (multi-GPU database sharding): local top-2 of every query
 * against db rows [0, ndb) reported with global indices index_base + row, as two packed keys
 * per query (dist << 32 | index; ~0 = none) in d_keys[2*nq]; asynchronous on m's stream. */
orbx_status orbx_knn2_device(orbx_matcher* m, const void* d_q, int nq, const void* d_db, int ndb,
                             long long index_base, unsigned long long* d_keys);
/* Merge `nparts` key sets (d_parts[(p*nq + q)*2 + {0,1}], e.g. the all-gathered per-rank
 * results) into final d1/i1/d2/i2/pass device arrays (any may be NULL). */
orbx_status orbx_knn2_merge_device(orbx_matcher* m, const unsigned long long* d_parts, int nparts,
                                   int nq, float ratio, int* d_d1, int* d_i1, int* d_d2,
                                   int* d_i2, uint8_t* d_pass);
/* orbx_knn2_device without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  Same arguments and results as
 * orbx_knn2_device; the per-slice partial keys live in a static 8 MiB device array inside
 * liborbx.so (orbx_knn2_noalloc_scratch_bytes(); held by every GPU that loads the library until
 * exit) instead of the handle's grow-on-demand scratch.  The call needs
 * slices * nq * 16 bytes of it (slices = how many grid.y slices the database is cut into: about
 * 444 / ceil(nq / 1024) for large databases, 1 -- no scratch at all -- when nq >= ~450 k and
 * ndb <= 4 M); more returns ORBX_ERR_CAPACITY with the numbers in orbx_last_error().  Calls no
 * cudaMalloc / cudaFree.  All calls on a device share the array: calls that overlap on the GPU
 * (two handles, two streams) race -- serialise them or use orbx_knn2_device. */
orbx_status orbx_knn2_device_noalloc(orbx_matcher* m, const void* d_q, int nq, const void* d_db, int ndb,
                                     long long index_base, unsigned long long* d_keys);
size_t orbx_knn2_noalloc_scratch_bytes(void);
/* tuning knob for the kNN kernel: 0 = 8 POPC per pair, 1 = carry-save adder tree (4 POPC) */
orbx_status orbx_matcher_set_knn_variant(orbx_matcher* m, int variant);
orbx_status orbx_matcher_synchronize(orbx_matcher* m);

/* ------------------------------------------------------------------------------------------
 * ROS-side image pre-processing
 *   ImageGrabber::ImageGrabber         ros2_ws/src/mono-inertial/include/image_grabber.hpp:38-49
 *                                      (holds m_undistortion_map_1/2 and the output size m_width x m_height; the maps
 *                                      are made by cv::fisheye::initUndistortRectifyMap for a 2048x1536 BGR camera at
 *                                      ros2_ws/src/mono-inertial/src/mono_inertial_node.cpp:54-71)
 *   ImageGrabber::ConvertImageToGPU    ros2_ws/src/mono-inertial/include/image_grabber.hpp:96-110
 *       cv::cuda::remap(INTER_CUBIC, BORDER_CONSTANT 0) -> cv::cuda::resize(INTER_LINEAR) -> cvtColor(BGR2GRAY)
 * One handle = one camera.  map_x / map_y are the two CV_32FC1 planes of initUndistortRectifyMap (map_w x map_h,
 * dense, host memory; copied).  Input: 8-bit BGR interleaved, src_w x src_h; output: 8-bit gray, dst_w x dst_h, dense.
 * The arithmetic is that of the CPU cv::remap / cv::resize / cv::cvtColor, bit for bit (csrc/preproc.cu header).
 * Calls on one handle share its staging buffers: whatever streams they are enqueued on, each call is ordered after the
 * previous one on the device (an event), so one handle can feed several rotating extractor handles; the HOST side of
 * a handle is not re-entrant (one calling thread per handle). */
typedef struct orbx_preproc orbx_preproc;
orbx_status orbx_preproc_create(int src_w, int src_h, const float* map_x, const float* map_y, int map_w, int map_h,
                                int dst_w, int dst_h, int max_batch, int device, orbx_preproc** out);
void orbx_preproc_destroy(orbx_preproc* p);
void* orbx_preproc_stream(orbx_preproc* p);
/* `batch` BGR frames in device memory (frame f at d_bgr + f*frame_stride, rows row_stride BYTES apart) -> d_gray
 * (batch x dst_h x dst_w, dense).  Asynchronous on `stream` (a cudaStream_t; NULL = the handle's stream). */
orbx_status orbx_preproc_device(orbx_preproc* p, const uint8_t* d_bgr, size_t frame_stride, int row_stride, int batch,
                                uint8_t* d_gray, void* stream);
/* the same from / to host memory; returns with h_gray filled */
orbx_status orbx_preproc_host(orbx_preproc* p, const uint8_t* h_bgr, size_t frame_stride, int row_stride, int batch,
                              uint8_t* h_gray);
/* ConvertImageToGPU -> System::TrackMonocular -> ORBextractor::operator() chained on the device: the BGR frames are
 * copied in, pre-processed and extracted on e's stream; the gray images never visit the host.  e must have been made
 * for dst_w x dst_h on the same device.  Asynchronous like orbx_extract_host; results through orbx_extractor_sync /
 * orbx_extractor_fetch. */
orbx_status orbx_preproc_extract_host(orbx_preproc* p, orbx_extractor* e, const uint8_t* h_bgr, size_t frame_stride,
                                      int row_stride, int batch, const int* lapping);

#ifdef __cplusplus
}
#endif
#endif
