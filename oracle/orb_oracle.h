/*
 * orb_oracle.h -- CPU ORACLE for the ORB front-end hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product library (liborbx.so, the
 * C++ shim in include/orbx/, the python package) may link, import or call
 * anything declared here; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs do.
 *
 * Parity status: the reference ships NO tests or golden vectors for this path
 * and cannot be compiled in this image (needs OpenCV-CUDA, Eigen, Pangolin ...;
 * SURVEY.md F2/F3), so the oracle is pinned against Python cv2 4.13.0 for every
 * OpenCV primitive it restates (resize, copyMakeBorder, GaussianBlur, FAST,
 * fastAtan2, ORB descriptor sampling) -- tests/test_oracle_cv2.py and the
 * committed fixtures in tests/golden/.  DistributeOctTree / matcher logic
 * follows the cited reference lines.  "parity unpinned by the reference".
 *
 * All file:line citations are into /root/reference (geoeo/ORB_SLAM3_V1.0).
 */
#ifndef ORB_ORACLE_H
#define ORB_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* cv::KeyPoint memory layout (28 bytes) -- target ABI, SURVEY 8(a) a16 */
typedef struct {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orc_keypoint;

/* ---- OpenCV primitive restatements ---- */
void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride,
                          uint8_t* dst, int dw, int dh, int dstride);
void orc_border_reflect101(uint8_t* buf, int w, int h, int stride, int border);
/* ROS-side pre-processing, ros2_ws/src/mono-inertial/include/image_grabber.hpp:96-110 (CPU arithmetic of cv::remap
 * INTER_CUBIC / BORDER_CONSTANT 0, cv::resize INTER_LINEAR, cv::cvtColor BGR2GRAY) */
void orc_remap_cubic_table(short* itab /* [32*32][16] */);
void orc_remap_cubic_bgr(const uint8_t* src, int sw, int sh, const float* mapx, const float* mapy, int mw, int mh, uint8_t* dst);
void orc_preprocess_bgr(const uint8_t* src, int sw, int sh, const float* mapx, const float* mapy, int mw, int mh, int dw, int dh,
                        uint8_t* gray);
void orc_gaussian7_s2_u8(const uint8_t* src, int w, int h, int sstride,
                         uint8_t* dst, int dstride);
/* cv::FAST(img, kps, th, nonmax=true, TYPE_9_16) on a w x h sub-image.
 * Returns the number of keypoints; (x,y,score) in raster order. */
int orc_fast9_16(const uint8_t* img, int w, int h, int stride, int threshold,
                 int* xs, int* ys, int* scores, int max_out);
float orc_fast_atan2(float y, float x);

/* ---- extractor (upstream ORBextractor semantics) ---- */
typedef struct orc_extractor orc_extractor;
orc_extractor* orc_extractor_create(int nfeatures, float scaleFactor, int nlevels,
                                    int iniThFAST, int minThFAST);
void orc_extractor_destroy(orc_extractor*);
/* tables: each array has nlevels entries; umax has 16 */
void orc_extractor_tables(const orc_extractor*, float* scale, float* inv_scale,
                          float* sigma2, float* inv_sigma2, int* feat_per_level,
                          int* umax);
/* operator(): returns number of keypoints (or -1 on empty image); *mono_index
 * receives the return value of the reference operator(). kps/desc may be NULL
 * to query the count. lapping = {x0,x1}. */
int orc_extract(orc_extractor*, const uint8_t* image, int width, int height,
                int stride, const int* lapping, orc_keypoint* kps, uint8_t* desc,
                int max_kps, int* mono_index);
/* state left by the last orc_extract */
int orc_level_size(const orc_extractor*, int level, int* w, int* h);
/* copies the bordered (w+38)x(h+38) level buffer, densely packed */
void orc_level_bordered(const orc_extractor*, int level, uint8_t* out);
/* FAST candidates handed to DistributeOctTree at `level` (coords relative to
 * minBorder); returns count */
int orc_level_candidates(const orc_extractor*, int level, int* xs, int* ys,
                         int* resp, int max_out);
/* per-level keypoints after the quadtree (level coords, incl. border offset
 * removed: x,y are ROI coordinates), angle filled */
int orc_level_keypoints(const orc_extractor*, int level, orc_keypoint* out, int max_out);

/* DistributeOctTree + best-per-node (src/ORBextractor.cc:226-431, 514-533).
 * Inputs are candidate coordinates relative to (minX,minY). Writes the index
 * (into the input arrays) of each selected keypoint, in list order. */
int orc_distribute_octtree(int n, const int* xs, const int* ys, const int* resp,
                           int minX, int maxX, int minY, int maxY, int N,
                           int* out_idx, int max_out);

/* IC_Angle on a bordered level image: img points at ROI (0,0) */
float orc_ic_angle(const uint8_t* roi, int stride, int x, int y);
/* rBRIEF on a blurred image (ROI origin), keypoint at integer (x,y), angle deg */
void orc_orb_descriptor(const uint8_t* blurred_roi, int stride, int x, int y,
                        float angle_deg, uint8_t* desc32);

/* ---- matcher ---- */
int orc_descriptor_distance(const uint8_t* a, const uint8_t* b); /* ORBmatcher.cc:1375-1391 */
void orc_hamming_pairs(const uint8_t* a, const uint8_t* b, int n, int* out);

/* ORBmatcher::ComputeThreeMaxima (ORBmatcher.cc:1328-1370) on bin counts */
void orc_three_maxima(const int* counts, int L, int* ind1, int* ind2, int* ind3);

/* Frame grid (Frame.cc:157-176, 404-480) flattened */
typedef struct {
    int n;                    /* frame keypoints */
    const float* kx; const float* ky; const int* koct;   /* mvKeysUn */
    const uint8_t* desc;      /* n x 32 */
    const float* uright;      /* mvuRight, may be NULL (treated as -1) */
    float minx, miny;         /* mnMinX, mnMinY */
    float gw_inv, gh_inv;     /* mfGridElementWidthInv / HeightInv */
    int gcols, grows;         /* mFrameGridCols / Rows */
    const float* scale;       /* mvScaleFactors */
    int nlevels;
} orc_frame_view;

/* MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:343-412) on N observation descriptors:
 * index of the descriptor with the least median distance to the rest (first on ties), -1 if N==0. */
int orc_distinctive_descriptor(const uint8_t* desc, int N, int* median_out);

/* DBoW2 FORB::meanValue (Thirdparty/DBoW2/src/DBoW2/FORB.cpp:28-79): per-bit majority of N descriptors -> 32 bytes */
void orc_forb_mean_value(const uint8_t* desc, int N, uint8_t* mean);

/* Frame::AssignFeaturesToGrid + PosInGrid (Frame.cc:157-176, 470-480) as a CSR over the cells
 * (linear index y*cols + x, push_back order inside a cell); cell_idx has F->n entries, the unused
 * tail (keypoints outside the grid) is -1. */
void orc_assign_features_to_grid(const orc_frame_view* F, int* cell_off, int* cell_idx);
/* Frame::GetFeaturesInArea (Frame.cc:404-468) for nq queries: out_count[q] = vIndices.size(),
 * out_idx[q][0..min(count,cap)) = vIndices in the reference's order. */
void orc_features_in_area(const orc_frame_view* F, int nq, const float* x, const float* y, const float* r,
                          const int* minLevel, const int* maxLevel, int* out_count, int* out_idx, int cap);

/* SearchByProjection(Frame, MapPoints, th, ...) ORBmatcher.cc:31-123.
 * Map points are flattened: in-view flag, projected x,y, predicted level,
 * view cos, descriptor. `claimed[i]!=0` marks frame keypoints that already
 * hold a MapPoint with observations. mp_proj_xr = mTrackProjXR (read when
 * F->uright != NULL: the stereo gate of :81-86); mp_has_obs[i] = Observations() > 0
 * of map point i (NULL = all): only such a point blocks a keypoint it was assigned
 * to (:77-79). match_mp[i] receives the map-point index finally held by frame
 * keypoint i (or -1). Returns nmatches (every assignment counts). */
int orc_search_by_projection(const orc_frame_view* F, int nmp, const uint8_t* mp_inview,
                             const float* mp_x, const float* mp_y, const int* mp_level,
                             const float* mp_viewcos, const float* mp_depth,
                             const uint8_t* mp_desc, const float* mp_proj_xr,
                             const uint8_t* mp_has_obs, const uint8_t* claimed_in,
                             float th, int bFarPoints, float thFarPoints, float nnRatio,
                             int* match_mp);

/* SearchByProjection(Frame, KeyFrame, ...) ORBmatcher.cc:1202-1326 with the
 * projection already done by the host shim: per KF map point i (valid flag,
 * u,v, predicted level, KF keypoint angle, descriptor). */
int orc_search_by_projection_kf(const orc_frame_view* F, const float* fangle, int nmp,
                                const uint8_t* mp_valid, const float* mp_u, const float* mp_v,
                                const int* mp_level, const float* mp_angle,
                                const uint8_t* mp_desc, const uint8_t* occupied_in,
                                float th, int checkOrientation, int* match_mp);

/* SearchForTriangulation ORBmatcher.cc:441-676 flattened: two keyframes,
 * feature vectors as sorted (node id, CSR offsets, indices). F12 = fundamental
 * matrix as computed by Pinhole::epipolarConstrain (Pinhole.cpp:104-131),
 * row-major 3x3; ep = epipole in image 2. */
typedef struct {
    int n;
    const float* kx; const float* ky; const int* koct; const float* kangle;
    const uint8_t* desc;
    const uint8_t* has_mp;       /* GetMapPoint(idx)!=null */
    const float* uright;         /* may be NULL */
    int nnodes; const int* node_id; const int* node_off; const int* node_idx;
} orc_kf_view;
/* Search core of Fuse (ORBmatcher.cc:678-862 variant 0, :864-975 variant 1) on projected map
 * points (valid, u, v, ur, predicted level, descriptor): best_idx[i] = selected keypoint if
 * bestDist <= TH_LOW else -1; best_dist[i] = its distance (-1 if no candidate). */
void orc_fuse_search(const orc_frame_view* KF, const float* uright, const float* inv_sigma2, int nmp,
                     const uint8_t* valid, const float* mu, const float* mv, const float* mur, const int* mlevel,
                     const uint8_t* mdesc, float th, int variant, int* best_idx, int* best_dist);

/* SearchBySim3 (ORBmatcher.cc:977-1200) after the host-side projections: per KF1 keypoint its map
 * point projected into KF2 (valid, u, v, predicted level, descriptor) and the converse;
 * match12[i1] = keypoint of KF2 on agreement, else -1; returns nFound. */
int orc_search_by_sim3(const orc_frame_view* KF1, const orc_frame_view* KF2,
                       const uint8_t* valid1, const float* u1in2, const float* v1in2, const int* level1, const uint8_t* desc1,
                       const uint8_t* valid2, const float* u2in1, const float* v2in1, const int* level2, const uint8_t* desc2,
                       float th, int* match12);

/* SearchForInitialization ORBmatcher.cc:329-439: vnMatches12 (F1->n entries), returns nmatches. */
int orc_search_for_initialization(const orc_frame_view* F1, const float* angle1, const orc_frame_view* F2,
                                  const float* angle2, int windowSize, float nnRatio, int checkOrientation,
                                  int* vnMatches12);

/* SearchByBoW(KeyFrame, Frame, ...) ORBmatcher.cc:133-327: F is flattened like a keyframe (angles,
 * descriptors, feature vector); match_kf[i] (F->n entries) = KeyFrame keypoint whose MapPoint goes
 * to Frame keypoint i, or -1.  Nleft = F->Nleft (-1: one camera). */
int orc_search_by_bow(const orc_kf_view* KF, const orc_kf_view* F, int Nleft, float nnRatio,
                      int checkOrientation, int* match_kf);

int orc_search_for_triangulation(const orc_kf_view* k1, const orc_kf_view* k2,
                                 const float* F12, float epx, float epy,
                                 const float* scale2 /* pKF2->mvScaleFactors */,
                                 int bOnlyStereo, int bCoarse, int checkOrientation,
                                 int* pairs /* 2*n1 */, int* npairs);

/* Frame::ComputeStereoMatches (upstream v1.0). Pyramids are
 * given as per-level ROI pointers + strides (bordered buffers behind them). */
typedef struct {
    int n; const float* kx; const float* ky; const int* koct; const uint8_t* desc;
    int nlevels; const uint8_t* const* lvl; const int* lvl_stride;
    const int* lvl_w; const int* lvl_h;
} orc_stereo_side;
void orc_stereo_matches(const orc_stereo_side* L, const orc_stereo_side* R,
                        const float* scale, const float* inv_scale,
                        float bf, float b, float* uRight, float* depth);

/* brute-force kNN k=2: ties -> lowest index. out arrays have nq entries. */
void orc_knn2(const uint8_t* q, int nq, const uint8_t* db, int ndb,
              int* d1, int* i1, int* d2, int* i2, int nthreads);

/* timing helper for bench.py: runs `iters` extractions on `nthreads` threads
 * (independent frames), returns seconds of wall time. */
double orc_bench_extract(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh,
                         const uint8_t* images, int nimages, int width, int height,
                         int iters_per_thread, int nthreads, long* total_kps);

#ifdef __cplusplus
}
#endif
#endif
