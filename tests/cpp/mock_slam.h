// Minimal stand-ins for the reference's Frame / KeyFrame / MapPoint / camera / pose types, exposing exactly the
// members the bodies of /root/reference src/ORBmatcher.cc read (see the per-function lists in
// include/orbx/ORBmatcher_ref.h).  Geometry is a tiny float implementation of the Sophus / Eigen subset used.
// Test infrastructure only.
#ifndef ORBX_TEST_MOCK_SLAM_H
#define ORBX_TEST_MOCK_SLAM_H
#include <cmath>
#include <cstdint>
#include <map>
#include <memory>
#include <set>
#include <tuple>
#include <vector>

namespace mock {

struct Vec2 { float v[2]; float operator()(int i) const { return v[i]; } };
struct Vec3 {
    float v[3];
    float operator()(int i) const { return v[i]; }
    Vec3 operator-(const Vec3& o) const { return Vec3{{v[0] - o.v[0], v[1] - o.v[1], v[2] - o.v[2]}}; }
    Vec3 operator/(float s) const { return Vec3{{v[0] / s, v[1] / s, v[2] / s}}; }
    float norm() const { return std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }
};
struct Mat3 { float m[9]; float operator()(int i, int j) const { return m[3 * i + j]; } };
struct SE3 {            /* x -> R x + t */
    Mat3 R{{1, 0, 0, 0, 1, 0, 0, 0, 1}};
    Vec3 t{{0, 0, 0}};
    SE3() {}
    SE3(const Mat3& R_, const Vec3& t_) : R(R_), t(t_) {}
    Vec3 operator*(const Vec3& x) const {
        Vec3 o;
        for (int i = 0; i < 3; ++i) o.v[i] = R.m[3 * i] * x.v[0] + R.m[3 * i + 1] * x.v[1] + R.m[3 * i + 2] * x.v[2] + t.v[i];
        return o;
    }
    SE3 operator*(const SE3& b) const {
        SE3 o;
        for (int i = 0; i < 3; ++i)
            for (int j = 0; j < 3; ++j) o.R.m[3 * i + j] = R.m[3 * i] * b.R.m[j] + R.m[3 * i + 1] * b.R.m[3 + j] + R.m[3 * i + 2] * b.R.m[6 + j];
        o.t = (*this) * b.t;
        return o;
    }
    SE3 inverse() const {
        SE3 o;
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) o.R.m[3 * i + j] = R.m[3 * j + i];
        for (int i = 0; i < 3; ++i) o.t.v[i] = -(o.R.m[3 * i] * t.v[0] + o.R.m[3 * i + 1] * t.v[1] + o.R.m[3 * i + 2] * t.v[2]);
        return o;
    }
    Mat3 rotationMatrix() const { return R; }
    Vec3 translation() const { return t; }
};
struct Sim3 {           /* x -> s R x + t */
    float s = 1;
    Mat3 R{{1, 0, 0, 0, 1, 0, 0, 0, 1}};
    Vec3 t{{0, 0, 0}};
    Vec3 operator*(const Vec3& x) const {
        Vec3 o;
        for (int i = 0; i < 3; ++i) o.v[i] = s * (R.m[3 * i] * x.v[0] + R.m[3 * i + 1] * x.v[1] + R.m[3 * i + 2] * x.v[2]) + t.v[i];
        return o;
    }
    Sim3 inverse() const {
        Sim3 o;
        o.s = 1.f / s;
        for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) o.R.m[3 * i + j] = R.m[3 * j + i];
        for (int i = 0; i < 3; ++i) o.t.v[i] = -o.s * (o.R.m[3 * i] * t.v[0] + o.R.m[3 * i + 1] * t.v[1] + o.R.m[3 * i + 2] * t.v[2]);
        return o;
    }
    Mat3 rotationMatrix() const { return R; }
    Vec3 translation() const { return t; }
    float scale() const { return s; }
};
struct Pinhole {
    float fx = 1, fy = 1, cx = 0, cy = 0;
    Vec2 project(const Vec3& p) const { return Vec2{{fx * p.v[0] / p.v[2] + cx, fy * p.v[1] / p.v[2] + cy}}; }   /* Pinhole.cpp:31-37 */
    Mat3 toK_() const { return Mat3{{fx, 0, cx, 0, fy, cy, 0, 0, 1}}; }
};
struct Pt { float x, y; };
struct KeyPoint { Pt pt; float size = 31, angle = 0, response = 0; int octave = 0, class_id = -1; };
struct Desc { const uint8_t* data; };     /* cv::Mat stand-in: row of 32 bytes / matrix start */
typedef std::map<unsigned, std::vector<unsigned>> FeatureVector;

struct Frame;
struct KeyFrame;
struct MapPoint {
    bool mbTrackInView = false, mbTrackInViewR = false, bad = false;
    float mTrackProjX = 0, mTrackProjY = 0, mTrackProjXR = 0, mTrackViewCos = 1, mTrackDepth = 1;
    int mnTrackScaleLevel = 0, nobs = 1, level = 0, id = -1;
    Vec3 pos{{0, 0, 1}};
    uint8_t desc[32] = {0};
    std::vector<std::pair<const KeyFrame*, int>> obs;
    std::vector<int> events;              /* test log: what Fuse did to this point */
    bool isBad() const { return bad; }
    int Observations() const { return nobs; }
    Desc GetDescriptor() const { return Desc{desc}; }
    Vec3 GetWorldPos() const { return pos; }
    float GetMaxDistanceInvariance() const { return 1e9f; }
    float GetMinDistanceInvariance() const { return 0.f; }
    template <class T> int PredictScale(float, const T&) const { return level; }
    std::tuple<int, int> GetIndexInKeyFrame(const std::shared_ptr<KeyFrame>& kf) const {
        for (auto& o : obs) if (o.first == kf.get()) return std::make_tuple(o.second, -1);
        return std::make_tuple(-1, -1);
    }
    bool IsInKeyFrame(const std::shared_ptr<KeyFrame>& kf) const { for (auto& o : obs) if (o.first == kf.get()) return true; return false; }
    void AddObservation(const std::shared_ptr<KeyFrame>& kf, int idx) { obs.emplace_back(kf.get(), idx); nobs++; }
    void Replace(const std::shared_ptr<MapPoint>& by) { bad = true; replaced_by = by.get(); }
    const MapPoint* replaced_by = nullptr;
};

struct FrameBase {
    std::shared_ptr<std::vector<KeyPoint>> mvKeysUn = std::make_shared<std::vector<KeyPoint>>();
    Desc mDescriptors{nullptr};
    std::vector<float> mvuRight, mvScaleFactors, mvInvLevelSigma2;
    FeatureVector mFeatVec;
    float mnMinX = 0, mnMinY = 0, mnMaxX = 1e9f, mnMaxY = 1e9f, mfGridElementWidthInv = 0, mfGridElementHeightInv = 0;
    std::shared_ptr<Pinhole> mpCamera = std::make_shared<Pinhole>(), mpCamera2;
    SE3 Tcw;
    SE3 GetPose() const { return Tcw; }
    SE3 GetPoseInverse() const { return Tcw.inverse(); }
};
struct Frame : FrameBase {
    int mNumKeypoints = 0, Nleft = -1, gcols = 64, grows = 48;
    std::vector<std::shared_ptr<MapPoint>> mvpMapPoints;
    int getFrameGridCols() const { return gcols; }
    int getFrameGridRows() const { return grows; }
};
struct KeyFrame : FrameBase {
    int N = 0, mnGridCols = 64, mnGridRows = 48;
    float mbf = 0, fx = 1, fy = 1, cx = 0, cy = 0;
    std::set<std::shared_ptr<MapPoint>> GetMapPoints() const {
        std::set<std::shared_ptr<MapPoint>> s;
        for (auto& p : mvpMapPoints) if (p && !p->bad) s.insert(p);
        return s;
    }
    Vec3 camera_center{{0, 0, 0}};
    std::vector<std::shared_ptr<MapPoint>> mvpMapPoints;
    std::vector<std::shared_ptr<MapPoint>> GetMapPointMatches() const { return mvpMapPoints; }
    std::shared_ptr<MapPoint> GetMapPoint(int idx) const { return mvpMapPoints[idx]; }
    void AddMapPoint(const std::shared_ptr<MapPoint>& p, int idx) { mvpMapPoints[idx] = p; }
    Vec3 GetTranslationInverse() const { return camera_center; }
    bool IsInImage(float x, float y) const { return x >= mnMinX && x < mnMaxX && y >= mnMinY && y < mnMaxY; }
};

}  // namespace mock
#endif
