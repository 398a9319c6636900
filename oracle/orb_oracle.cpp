/*
 * orb_oracle.cpp -- CPU ORACLE (test infrastructure only; see orb_oracle.h).
 *
 * A plain C++17 restatement of the upstream ORB-SLAM3 v1.0 front-end semantics
 * that BASELINE.json's north_star names, built from
 *   - the reference code that is citable under /root/reference (quadtree,
 *     tables, pattern, matcher) and
 *   - OpenCV CPU arithmetic restated from SURVEY.md Appendix A/B.
 * Build: oracle/Makefile (g++ -O2 -ffp-contract=off).  No OpenCV dependency.
 * Parity status: "parity unpinned by the reference" -- it ships no tests and cannot be built here.  What pins this
 * file instead (DESIGN.md section 4 has the table):
 *   - Python cv2 4.13.0 for everything OpenCV defines: resize, border, blur, FAST, fastAtan2, IC_Angle (through
 *     cv2.ORB's angles), rBRIEF (through cv2.ORB's descriptors), remap / resize / cvtColor of the ROS front, and the
 *     whole operator() re-composed from cv2 primitives (tests/test_oracle_golden.py, tests/test_oracle_cv2_live.py);
 *   - second restatements in straight Python, written separately from the same source text, for what is ORB-SLAM's
 *     own logic: quadtree (tests/cv2_compose.py), every matcher search (tests/test_oracle_matcher.py,
 *     tests/test_oracle_matcher_py.py), ComputeStereoMatches (tests/test_oracle_stereo_py.py; its specification,
 *     SURVEY App. A5, is itself a recall of upstream and cannot be checked in this tree).
 */
#include "orb_oracle.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <climits>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <list>
#include <thread>
#include <utility>
#include <vector>

namespace {

const int PATCH_SIZE = 31;       /* src/ORBextractor.cc:78 */
const int HALF_PATCH_SIZE = 15;  /* src/ORBextractor.cc:79 */
const int EDGE_THRESHOLD = 19;   /* upstream value (fork has 5, :80); SURVEY App. A */

static const int8_t kPattern[1024] = {
#include "../include/orbx_pattern31.inc"
};

inline int cvRoundf(float v) { return (int)lrintf(v); }   /* round-half-even */
inline int cvRoundd(double v) { return (int)lrint(v); }

/* ------------------------------------------------------------------ B1 */
void resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride,
                      uint8_t* dst, int dw, int dh, int dstride) {
    const double scale_x = (double)sw / dw, scale_y = (double)sh / dh;
    std::vector<int> xofs(dw), yofs(dh);
    std::vector<short> ax(2 * dw), ay(2 * dh);
    for (int dx = 0; dx < dw; ++dx) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = (int)floorf(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        xofs[dx] = sx;
        ax[2 * dx] = (short)cvRoundf((1.f - fx) * 2048);
        ax[2 * dx + 1] = (short)cvRoundf(fx * 2048);
    }
    for (int dy = 0; dy < dh; ++dy) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = (int)floorf(fy);
        fy -= sy;
        yofs[dy] = sy;   /* rows are NOT clamped in the table (unlike columns): the two source rows are clipped when
                          * fetched, so an enlarging resize blends row 0 (or sh-1) with itself with both weights */
        ay[2 * dy] = (short)cvRoundf((1.f - fy) * 2048);
        ay[2 * dy + 1] = (short)cvRoundf(fy * 2048);
    }
    std::vector<int> r0(dw), r1(dw);
    for (int dy = 0; dy < dh; ++dy) {
        const int y0 = std::min(std::max(yofs[dy], 0), sh - 1), y1 = std::min(std::max(yofs[dy] + 1, 0), sh - 1);
        const uint8_t* s0 = src + (size_t)y0 * sstride;
        const uint8_t* s1 = src + (size_t)y1 * sstride;
        for (int dx = 0; dx < dw; ++dx) {
            const int x0 = xofs[dx], x1 = std::min(x0 + 1, sw - 1);
            r0[dx] = s0[x0] * ax[2 * dx] + s0[x1] * ax[2 * dx + 1];
            r1[dx] = s1[x0] * ax[2 * dx] + s1[x1] * ax[2 * dx + 1];
        }
        const int b0 = ay[2 * dy], b1 = ay[2 * dy + 1];
        uint8_t* d = dst + (size_t)dy * dstride;
        for (int dx = 0; dx < dw; ++dx)
            d[dx] = (uint8_t)((((b0 * (r0[dx] >> 4)) >> 16) + ((b1 * (r1[dx] >> 4)) >> 16) + 2) >> 2);
    }
}

/* ------------------------------------------------------------------ B5 */
inline int reflect101(int p, int n) {
    if (n == 1) return 0;
    while (p < 0 || p >= n) {
        if (p < 0) p = -p;
        else p = 2 * (n - 1) - p;
    }
    return p;
}

/* buf points at the top-left of the (w+2b)x(h+2b) bordered buffer whose
 * interior [b,b+w)x[b,b+h) is already filled. */
void border_reflect101(uint8_t* buf, int w, int h, int stride, int b) {
    for (int y = 0; y < h; ++y) {
        uint8_t* row = buf + (size_t)(y + b) * stride;
        for (int x = 0; x < b; ++x) {
            row[x] = row[b + reflect101(x - b, w)];
            row[b + w + x] = row[b + reflect101(w + x, w)];
        }
    }
    for (int y = 0; y < b; ++y) {
        memcpy(buf + (size_t)y * stride, buf + (size_t)(b + reflect101(y - b, h)) * stride, w + 2 * b);
        memcpy(buf + (size_t)(b + h + y) * stride, buf + (size_t)(b + reflect101(h + y, h)) * stride, w + 2 * b);
    }
}

/* ------------------------------------------------------------------ B2 */
void gaussian7(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride) {
    static const int t[7] = {18, 34, 48, 56, 48, 34, 18};
    std::vector<int> tmp((size_t)w * h);
    for (int y = 0; y < h; ++y) {
        const uint8_t* s = src + (size_t)y * sstride;
        for (int x = 0; x < w; ++x) {
            int acc = 0;
            for (int k = -3; k <= 3; ++k) acc += t[k + 3] * s[reflect101(x + k, w)];
            tmp[(size_t)y * w + x] = acc;
        }
    }
    for (int y = 0; y < h; ++y) {
        uint8_t* d = dst + (size_t)y * dstride;
        for (int x = 0; x < w; ++x) {
            int acc = 0;
            for (int k = -3; k <= 3; ++k) acc += t[k + 3] * tmp[(size_t)reflect101(y + k, h) * w + x];
            int v = (acc + (1 << 15)) >> 16;
            d[x] = (uint8_t)std::min(std::max(v, 0), 255);
        }
    }
}

/* ------------------------------------------------------------------ B4 */
/* Bresenham ring r=3, same layout as src/cuda/Fast_gpu.cu:226-254 / cv::FAST */
static const int kRingX[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingY[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

/* A = max over the 16 arcs of 9 contiguous ring pixels of min(ring - v) (bright)
 * or min(v - ring) (dark).  Corner at threshold t  <=>  A > t.
 * cornerScore (src/cuda/Fast_gpu.cu:193-216) = largest t that is still a corner = A-1. */
inline int fast_arc_strength(const uint8_t* p, int stride) {
    int d[16];
    const int v = p[0];
    for (int k = 0; k < 16; ++k) d[k] = (int)p[kRingY[k] * stride + kRingX[k]] - v;
    int best = INT_MIN;
    for (int s = 0; s < 16; ++s) {
        int mn = INT_MAX, mx = INT_MIN;
        for (int k = 0; k < 9; ++k) {
            const int e = d[(s + k) & 15];
            mn = std::min(mn, e);
            mx = std::max(mx, e);
        }
        best = std::max(best, std::max(mn, -mx));
    }
    return best;
}

/* every 9-arc of the 16-ring contains one pixel of each opposite pair (k, k+8):
 * if both pixels of a pair are within +-t of the centre there is no corner. */
inline bool fast_quick_reject(const uint8_t* p, int stride, int t) {
    const int v = p[0], lo = v - t, hi = v + t;
    for (int k = 0; k < 8; k += 2) {
        const int a = p[kRingY[k] * stride + kRingX[k]], b = p[kRingY[k + 8] * stride + kRingX[k + 8]];
        if (a >= lo && a <= hi && b >= lo && b <= hi) return true;
    }
    return false;
}

int fast9_16(const uint8_t* img, int w, int h, int stride, int threshold,
             int* xs, int* ys, int* scores, int max_out) {
    if (w < 7 || h < 7) return 0;
    threshold = std::min(std::max(threshold, 0), 255);
    /* score map: 0 for non-corners (cv::FAST keeps uchar score rows, zero-filled) */
    std::vector<uint8_t> sc((size_t)w * h, 0);
    for (int y = 3; y < h - 3; ++y)
        for (int x = 3; x < w - 3; ++x) {
            const uint8_t* p = img + (size_t)y * stride + x;
            if (fast_quick_reject(p, stride, threshold)) continue;
            const int A = fast_arc_strength(p, stride);
            if (A > threshold) sc[(size_t)y * w + x] = (uint8_t)(A - 1);
        }
    int n = 0;
    for (int y = 3; y < h - 3; ++y)
        for (int x = 3; x < w - 3; ++x) {
            const uint8_t* c = &sc[(size_t)y * w + x];
            const int s = c[0];
            if (s == 0) continue;   /* non-corner, or a score-0 corner which can never beat a neighbour */
            if (s > c[-1] && s > c[1] && s > c[-w - 1] && s > c[-w] && s > c[-w + 1] &&
                s > c[w - 1] && s > c[w] && s > c[w + 1]) {
                if (n < max_out) { xs[n] = x; ys[n] = y; scores[n] = s; }
                ++n;
            }
        }
    return n;
}

/* ------------------------------------------------------------------ B3 */
float fast_atan2(float y, float x) {
    const float scale = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * scale, p3 = -0.3258083974640975f * scale;
    const float p5 = 0.1555786518463281f * scale, p7 = -0.04432655554792128f * scale;
    const float eps = (float)2.2204460492503131e-16;
    float ax = fabsf(x), ay = fabsf(y), a, c, c2;
    if (ax >= ay) {
        c = ay / (ax + eps);
        c2 = c * c;
        a = (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    } else {
        c = ax / (ay + eps);
        c2 = c * c;
        a = 90.f - (((p7 * c2 + p5) * c2 + p3) * c2 + p1) * c;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}

/* ------------------------------------------------------- quadtree (a4) */
struct Cand { float x, y, resp; int idx; };

struct Node {               /* include/ORBextractor.h:39-50 */
    std::vector<Cand> vKeys;
    int ULx, ULy, URx, URy, BLx, BLy, BRx, BRy;
    std::list<Node>::iterator lit;
    bool bNoMore = false;
};

/* src/ORBextractor.cc:151-207 */
void DivideNode(const Node& p, Node& n1, Node& n2, Node& n3, Node& n4) {
    const int halfX = (int)ceilf((float)(p.URx - p.ULx) / 2);
    const int halfY = (int)ceilf((float)(p.BRy - p.ULy) / 2);
    n1.ULx = p.ULx; n1.ULy = p.ULy;
    n1.URx = p.ULx + halfX; n1.URy = p.ULy;
    n1.BLx = p.ULx; n1.BLy = p.ULy + halfY;
    n1.BRx = p.ULx + halfX; n1.BRy = p.ULy + halfY;
    n2.ULx = n1.URx; n2.ULy = n1.URy;
    n2.URx = p.URx; n2.URy = p.URy;
    n2.BLx = n1.BRx; n2.BLy = n1.BRy;
    n2.BRx = p.URx; n2.BRy = p.ULy + halfY;
    n3.ULx = n1.BLx; n3.ULy = n1.BLy;
    n3.URx = n1.BRx; n3.URy = n1.BRy;
    n3.BLx = p.BLx; n3.BLy = p.BLy;
    n3.BRx = n1.BRx; n3.BRy = p.BLy;
    n4.ULx = n3.URx; n4.ULy = n3.URy;
    n4.URx = n2.BRx; n4.URy = n2.BRy;
    n4.BLx = n3.BRx; n4.BLy = n3.BRy;
    n4.BRx = p.BRx; n4.BRy = p.BRy;
    for (const Cand& kp : p.vKeys) {
        if (kp.x < n1.URx) {
            if (kp.y < n1.BRy) n1.vKeys.push_back(kp);
            else n3.vKeys.push_back(kp);
        } else if (kp.y < n1.BRy) n2.vKeys.push_back(kp);
        else n4.vKeys.push_back(kp);
    }
    if (n1.vKeys.size() == 1) n1.bNoMore = true;
    if (n2.vKeys.size() == 1) n2.bNoMore = true;
    if (n3.vKeys.size() == 1) n3.bNoMore = true;
    if (n4.vKeys.size() == 1) n4.bNoMore = true;
}

/* src/ORBextractor.cc:209-224 -- NOT a total order (ties on size and UL.x) */
bool compareNodes(const std::pair<int, Node*>& e1, const std::pair<int, Node*>& e2) {
    if (e1.first < e2.first) return true;
    else if (e1.first > e2.first) return false;
    else return e1.second->ULx < e2.second->ULx;
}

/* src/ORBextractor.cc:226-431 + best-per-node :514-533. Returns selected
 * candidates in list order. */
std::vector<Cand> DistributeOctTree(const std::vector<Cand>& vToDistributeKeys, int minX, int maxX,
                                    int minY, int maxY, int N) {
    std::vector<Cand> result;
    if (vToDistributeKeys.empty()) return result;
    const int nIni = (int)roundf((float)(maxX - minX) / (maxY - minY));
    if (nIni < 1) return result; /* reference divides by zero here; defined as "no keypoints" */
    const float hX = (float)(maxX - minX) / nIni;

    std::list<Node> lNodes;
    std::vector<Node*> vpIniNodes(nIni);
    for (int i = 0; i < nIni; i++) {
        Node ni;
        ni.ULx = (int)(hX * (float)i); ni.ULy = 0;
        ni.URx = (int)(hX * (float)(i + 1)); ni.URy = 0;
        ni.BLx = ni.ULx; ni.BLy = maxY - minY;
        ni.BRx = ni.URx; ni.BRy = maxY - minY;
        lNodes.push_back(ni);
        vpIniNodes[i] = &lNodes.back();
    }
    for (const Cand& kp : vToDistributeKeys) vpIniNodes[(int)(kp.x / hX)]->vKeys.push_back(kp);

    auto lit = lNodes.begin();
    while (lit != lNodes.end()) {
        if (lit->vKeys.size() == 1) { lit->bNoMore = true; lit++; }
        else if (lit->vKeys.empty()) lit = lNodes.erase(lit);
        else lit++;
    }

    bool bFinish = false;
    std::vector<std::pair<int, Node*>> vSizeAndPointerToNode;

    auto addChild = [&](Node& n, int* nToExpand) {
        if (n.vKeys.size() > 0) {
            lNodes.push_front(n);
            if (n.vKeys.size() > 1) {
                if (nToExpand) (*nToExpand)++;
                vSizeAndPointerToNode.push_back(std::make_pair((int)n.vKeys.size(), &lNodes.front()));
                lNodes.front().lit = lNodes.begin();
            }
        }
    };

    while (!bFinish) {
        int prevSize = (int)lNodes.size();
        lit = lNodes.begin();
        int nToExpand = 0;
        vSizeAndPointerToNode.clear();
        while (lit != lNodes.end()) {
            if (lit->bNoMore) { lit++; continue; }
            Node n1, n2, n3, n4;
            DivideNode(*lit, n1, n2, n3, n4);
            addChild(n1, &nToExpand); addChild(n2, &nToExpand);
            addChild(n3, &nToExpand); addChild(n4, &nToExpand);
            lit = lNodes.erase(lit);
        }
        if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) {
            bFinish = true;
        } else if (((int)lNodes.size() + nToExpand * 3) > N) {
            while (!bFinish) {
                prevSize = (int)lNodes.size();
                std::vector<std::pair<int, Node*>> vPrev = vSizeAndPointerToNode;
                vSizeAndPointerToNode.clear();
                std::sort(vPrev.begin(), vPrev.end(), compareNodes);
                for (int j = (int)vPrev.size() - 1; j >= 0; j--) {
                    Node n1, n2, n3, n4;
                    DivideNode(*vPrev[j].second, n1, n2, n3, n4);
                    addChild(n1, nullptr); addChild(n2, nullptr);
                    addChild(n3, nullptr); addChild(n4, nullptr);
                    lNodes.erase(vPrev[j].second->lit);
                    if ((int)lNodes.size() >= N) break;
                }
                if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) bFinish = true;
            }
        }
    }

    result.reserve(lNodes.size());
    for (auto it = lNodes.begin(); it != lNodes.end(); it++) {
        const std::vector<Cand>& v = it->vKeys;
        const Cand* pKP = &v[0];
        float maxResponse = pKP->resp;
        for (size_t k = 1; k < v.size(); k++)
            if (v[k].resp > maxResponse) { pKP = &v[k]; maxResponse = v[k].resp; }
        result.push_back(*pKP);
    }
    return result;
}

/* ------------------------------------------------------- angle (a5) */
float ic_angle(const uint8_t* center, int step, const int* umax) {
    int m_01 = 0, m_10 = 0;
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
        int v_sum = 0;
        const int d = umax[v];
        for (int u = -d; u <= d; ++u) {
            const int val_plus = center[u + v * step], val_minus = center[u - v * step];
            v_sum += (val_plus - val_minus);
            m_10 += u * (val_plus + val_minus);
        }
        m_01 += v * v_sum;
    }
    return fast_atan2((float)m_01, (float)m_10);
}

/* ------------------------------------------------------- rBRIEF (a7) */
void orb_descriptor(const uint8_t* center, int step, float angle_deg, uint8_t* desc) {
    const float factorPI = (float)(3.14159265358979323846 / 180.f);
    const float angle = angle_deg * factorPI;
    const float a = cosf(angle), b = sinf(angle);
    const int8_t* pat = kPattern;
    for (int i = 0; i < 32; ++i, pat += 32) {
        int val = 0;
        for (int j = 0; j < 8; ++j) {
            const float x0 = pat[4 * j], y0 = pat[4 * j + 1], x1 = pat[4 * j + 2], y1 = pat[4 * j + 3];
            const int t0 = center[cvRoundf(x0 * b + y0 * a) * step + cvRoundf(x0 * a - y0 * b)];
            const int t1 = center[cvRoundf(x1 * b + y1 * a) * step + cvRoundf(x1 * a - y1 * b)];
            val |= (t0 < t1) << j;
        }
        desc[i] = (uint8_t)val;
    }
}

void compute_umax(std::vector<int>& umax) { /* src/ORBextractor.cc:126-143 */
    umax.assign(HALF_PATCH_SIZE + 1, 0);
    int v, v0, vmax = (int)floor(HALF_PATCH_SIZE * sqrt(2.f) / 2 + 1);
    int vmin = (int)ceil(HALF_PATCH_SIZE * sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (v = 0; v <= vmax; ++v) umax[v] = cvRoundd(sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
        while (umax[v0] == umax[v0 + 1]) ++v0;
        umax[v] = v0;
        ++v0;
    }
}

struct Level {
    int w = 0, h = 0, stride = 0;
    std::vector<uint8_t> buf;               /* (w+38)x(h+38) */
    uint8_t* roi() { return buf.data() + (size_t)EDGE_THRESHOLD * stride + EDGE_THRESHOLD; }
    const uint8_t* roi() const { return buf.data() + (size_t)EDGE_THRESHOLD * stride + EDGE_THRESHOLD; }
    std::vector<Cand> cands;                /* vToDistributeKeys */
    std::vector<orc_keypoint> kps;          /* level coords */
};

} // namespace

struct orc_extractor {
    int nfeatures; double scaleFactor; int nlevels, iniTh, minTh;
    std::vector<float> sf, inv_sf, sigma2, inv_sigma2;
    std::vector<int> nfeat, umax;
    std::vector<Level> lv;
};

extern "C" {

void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw,
                          int dh, int dstride) {
    resize_linear_u8(src, sw, sh, sstride, dst, dw, dh, dstride);
}
void orc_border_reflect101(uint8_t* buf, int w, int h, int stride, int border) {
    border_reflect101(buf, w, h, stride, border);
}
void orc_gaussian7_s2_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride) {
    gaussian7(src, w, h, sstride, dst, dstride);
}
int orc_fast9_16(const uint8_t* img, int w, int h, int stride, int threshold, int* xs, int* ys,
                 int* scores, int max_out) {
    return fast9_16(img, w, h, stride, threshold, xs, ys, scores, max_out);
}
float orc_fast_atan2(float y, float x) { return fast_atan2(y, x); }

/* ------------------------------------------------------------------ ROS-side pre-processing
 * ImageGrabber::ConvertImageToGPU, ros2_ws/src/mono-inertial/include/image_grabber.hpp:96-110:
 *   remap(BGR, map1, map2, INTER_CUBIC, BORDER_CONSTANT 0) -> resize(new size, INTER_LINEAR) -> cvtColor(BGR2GRAY).
 * The fork runs the cv::cuda versions (float arithmetic, not runnable here); this restates the arithmetic of the CPU
 * functions of the same names, pinned bit for bit against Python cv2 (tests/test_oracle_golden.py):
 *   remap: map -> fixed point with 5 fractional bits (cvRound(x * 32)), 4x4 taps from a 32 x 32 table of cubic
 *          weights (A = -0.75) scaled to 2^15 whose sum is forced to 2^15 by adjusting the largest / smallest of the
 *          four central weights, (sum + 2^14) >> 15, taps outside the image contribute the border value 0;
 *   resize: the 11-bit fixed-point bilinear of App. B1, per channel;  gray: (3735 B + 19235 G + 9798 R + 2^14) >> 15. */
static void cubic_coeffs(float x, float* c) {
    const float A = -0.75f;
    c[0] = ((A * (x + 1) - 5 * A) * (x + 1) + 8 * A) * (x + 1) - 4 * A;
    c[1] = ((A + 2) * x - (A + 3)) * x * x + 1;
    c[2] = ((A + 2) * (1 - x) - (A + 3)) * (1 - x) * (1 - x) + 1;
    c[3] = 1.f - c[0] - c[1] - c[2];
}
void orc_remap_cubic_table(short* itab /* [32*32][16] */) {
    float tab1[32][4];
    for (int i = 0; i < 32; ++i) cubic_coeffs((float)i / 32.f, tab1[i]);
    for (int i = 0; i < 32; ++i)
        for (int j = 0; j < 32; ++j) {
            short* w = itab + (size_t)(i * 32 + j) * 16;
            int isum = 0;
            for (int k1 = 0; k1 < 4; ++k1)
                for (int k2 = 0; k2 < 4; ++k2) {
                    const float v = tab1[i][k1] * tab1[j][k2];
                    int iv = cvRoundf(v * 32768.f);
                    iv = iv < -32768 ? -32768 : (iv > 32767 ? 32767 : iv);
                    w[k1 * 4 + k2] = (short)iv;
                    isum += iv;
                }
            if (isum != 32768) {
                const int diff = isum - 32768;
                int Mk1 = 2, Mk2 = 2, mk1 = 2, mk2 = 2;
                for (int k1 = 2; k1 < 4; ++k1)
                    for (int k2 = 2; k2 < 4; ++k2) {
                        if (w[k1 * 4 + k2] < w[mk1 * 4 + mk2]) { mk1 = k1; mk2 = k2; }
                        else if (w[k1 * 4 + k2] > w[Mk1 * 4 + Mk2]) { Mk1 = k1; Mk2 = k2; }
                    }
                if (diff < 0) w[Mk1 * 4 + Mk2] = (short)(w[Mk1 * 4 + Mk2] - diff);
                else w[mk1 * 4 + mk2] = (short)(w[mk1 * 4 + mk2] - diff);
            }
        }
}
void orc_remap_cubic_bgr(const uint8_t* src, int sw, int sh, const float* mapx, const float* mapy, int mw, int mh, uint8_t* dst) {
    std::vector<short> itab(32 * 32 * 16);
    orc_remap_cubic_table(itab.data());
    for (int y = 0; y < mh; ++y)
        for (int x = 0; x < mw; ++x) {
            const int sx = cvRoundf(mapx[(size_t)y * mw + x] * 32.f), sy = cvRoundf(mapy[(size_t)y * mw + x] * 32.f);
            const int ix = sx >> 5, iy = sy >> 5;
            const short* w = itab.data() + (size_t)((sy & 31) * 32 + (sx & 31)) * 16;
            int acc[3] = {0, 0, 0};
            for (int k1 = 0; k1 < 4; ++k1) {
                const int yy = iy - 1 + k1;
                if (yy < 0 || yy >= sh) continue;
                for (int k2 = 0; k2 < 4; ++k2) {
                    const int xx = ix - 1 + k2;
                    if (xx < 0 || xx >= sw) continue;
                    const uint8_t* p = src + ((size_t)yy * sw + xx) * 3;
                    for (int c = 0; c < 3; ++c) acc[c] += w[k1 * 4 + k2] * p[c];
                }
            }
            for (int c = 0; c < 3; ++c) {
                const int v = (acc[c] + (1 << 14)) >> 15;
                dst[((size_t)y * mw + x) * 3 + c] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
            }
        }
}
void orc_preprocess_bgr(const uint8_t* src, int sw, int sh, const float* mapx, const float* mapy, int mw, int mh, int dw, int dh,
                        uint8_t* gray) {
    std::vector<uint8_t> und((size_t)mw * mh * 3), plane((size_t)mw * mh), rs((size_t)dw * dh * 3), rp((size_t)dw * dh);
    orc_remap_cubic_bgr(src, sw, sh, mapx, mapy, mw, mh, und.data());
    for (int c = 0; c < 3; ++c) {
        for (size_t i = 0; i < (size_t)mw * mh; ++i) plane[i] = und[3 * i + c];
        if (dw == mw && dh == mh) rp = plane;
        else resize_linear_u8(plane.data(), mw, mh, mw, rp.data(), dw, dh, dw);
        for (size_t i = 0; i < (size_t)dw * dh; ++i) rs[3 * i + c] = rp[i];
    }
    for (size_t i = 0; i < (size_t)dw * dh; ++i)
        gray[i] = (uint8_t)((3735 * rs[3 * i] + 19235 * rs[3 * i + 1] + 9798 * rs[3 * i + 2] + (1 << 14)) >> 15);
}

orc_extractor* orc_extractor_create(int nfeatures, float scaleFactor_, int nlevels, int iniTh, int minTh) {
    orc_extractor* e = new orc_extractor;
    e->nfeatures = nfeatures; e->scaleFactor = scaleFactor_; e->nlevels = nlevels;
    e->iniTh = iniTh; e->minTh = minTh;
    /* src/ORBextractor.cc:92-124 */
    e->sf.resize(nlevels); e->sigma2.resize(nlevels);
    e->sf[0] = 1.0f; e->sigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
        e->sf[i] = e->sf[i - 1] * e->scaleFactor;
        e->sigma2[i] = e->sf[i] * e->sf[i];
    }
    e->inv_sf.resize(nlevels); e->inv_sigma2.resize(nlevels);
    for (int i = 0; i < nlevels; i++) {
        e->inv_sf[i] = 1.0f / e->sf[i];
        e->inv_sigma2[i] = 1.0f / e->sigma2[i];
    }
    e->nfeat.resize(nlevels);
    float factor = 1.0f / e->scaleFactor;
    float nDesired = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
    int sum = 0;
    for (int level = 0; level < nlevels - 1; level++) {
        e->nfeat[level] = cvRoundf(nDesired);
        sum += e->nfeat[level];
        nDesired *= factor;
    }
    e->nfeat[nlevels - 1] = std::max(nfeatures - sum, 0);
    compute_umax(e->umax);
    e->lv.resize(nlevels);
    return e;
}
void orc_extractor_destroy(orc_extractor* e) { delete e; }

void orc_extractor_tables(const orc_extractor* e, float* scale, float* inv_scale, float* sigma2,
                          float* inv_sigma2, int* feat_per_level, int* umax) {
    for (int i = 0; i < e->nlevels; ++i) {
        if (scale) scale[i] = e->sf[i];
        if (inv_scale) inv_scale[i] = e->inv_sf[i];
        if (sigma2) sigma2[i] = e->sigma2[i];
        if (inv_sigma2) inv_sigma2[i] = e->inv_sigma2[i];
        if (feat_per_level) feat_per_level[i] = e->nfeat[i];
    }
    if (umax) for (int i = 0; i < 16; ++i) umax[i] = e->umax[i];
}

int orc_extract(orc_extractor* e, const uint8_t* image, int width, int height, int stride,
                const int* lapping, orc_keypoint* kps, uint8_t* desc, int max_kps, int* mono_index) {
    if (!image || width <= 0 || height <= 0) { if (mono_index) *mono_index = -1; return -1; }
    const int B = EDGE_THRESHOLD;
    /* ---- ComputePyramid (fork sizes :596-599) ---- */
    for (int l = 0; l < e->nlevels; ++l) {
        Level& L = e->lv[l];
        const float scale = e->inv_sf[l];
        L.w = cvRoundf((float)width * scale);
        L.h = cvRoundf((float)height * scale);
        L.stride = L.w + 2 * B;
        L.buf.assign((size_t)L.stride * (L.h + 2 * B), 0);
        if (l == 0) {
            for (int y = 0; y < height; ++y) memcpy(L.roi() + (size_t)y * L.stride, image + (size_t)y * stride, width);
        } else {
            Level& P = e->lv[l - 1];
            resize_linear_u8(P.roi(), P.w, P.h, P.stride, L.roi(), L.w, L.h, L.stride);
        }
        border_reflect101(L.buf.data(), L.w, L.h, L.stride, B);
    }
    /* ---- ComputeKeyPointsOctTree ---- */
    const float W = 35;
    int nkeypoints = 0;
    for (int l = 0; l < e->nlevels; ++l) {
        Level& L = e->lv[l];
        L.cands.clear(); L.kps.clear();
        const int minBorderX = B - 3, minBorderY = minBorderX;
        const int maxBorderX = L.w - B + 3, maxBorderY = L.h - B + 3;
        const float width_f = (float)(maxBorderX - minBorderX), height_f = (float)(maxBorderY - minBorderY);
        const int nCols = (int)(width_f / W), nRows = (int)(height_f / W);
        if (nCols >= 1 && nRows >= 1) {
            const int wCell = (int)ceilf(width_f / nCols), hCell = (int)ceilf(height_f / nRows);
            std::vector<int> xs(4096), ys(4096), sc(4096);
            for (int i = 0; i < nRows; i++) {
                const float iniY = minBorderY + i * hCell;
                float maxY = iniY + hCell + 6;
                if (iniY >= maxBorderY - 3) continue;
                if (maxY > maxBorderY) maxY = maxBorderY;
                for (int j = 0; j < nCols; j++) {
                    const float iniX = minBorderX + j * wCell;
                    float maxX = iniX + wCell + 6;
                    if (iniX >= maxBorderX - 6) continue;
                    if (maxX > maxBorderX) maxX = maxBorderX;
                    const uint8_t* cell = L.roi() + (size_t)(int)iniY * L.stride + (int)iniX;
                    const int cw = (int)maxX - (int)iniX, ch = (int)maxY - (int)iniY;
                    int n = fast9_16(cell, cw, ch, L.stride, e->iniTh, xs.data(), ys.data(), sc.data(), 4096);
                    if (n == 0) n = fast9_16(cell, cw, ch, L.stride, e->minTh, xs.data(), ys.data(), sc.data(), 4096);
                    for (int k = 0; k < n && k < 4096; ++k) {
                        Cand c;
                        c.x = (float)xs[k] + j * wCell;
                        c.y = (float)ys[k] + i * hCell;
                        c.resp = (float)sc[k];
                        c.idx = (int)L.cands.size();
                        L.cands.push_back(c);
                    }
                }
            }
        }
        std::vector<Cand> sel = DistributeOctTree(L.cands, minBorderX, maxBorderX, minBorderY, maxBorderY, e->nfeat[l]);
        const int scaledPatchSize = (int)(PATCH_SIZE * e->sf[l]);
        for (const Cand& c : sel) {
            orc_keypoint k;
            k.x = c.x + minBorderX; k.y = c.y + minBorderY;
            k.size = (float)scaledPatchSize; k.angle = -1; k.response = c.resp;
            k.octave = l; k.class_id = -1;
            L.kps.push_back(k);
        }
        nkeypoints += (int)L.kps.size();
    }
    for (int l = 0; l < e->nlevels; ++l) {
        Level& L = e->lv[l];
        for (orc_keypoint& k : L.kps)
            k.angle = ic_angle(L.roi() + (size_t)cvRoundf(k.y) * L.stride + cvRoundf(k.x), L.stride, e->umax.data());
    }
    /* ---- operator() tail ---- */
    int monoIndex = 0, stereoIndex = nkeypoints - 1;
    const int lap0 = lapping ? lapping[0] : 0, lap1 = lapping ? lapping[1] : 0;
    const bool write = kps && desc && max_kps >= nkeypoints;
    std::vector<uint8_t> blurred;
    for (int l = 0; l < e->nlevels; ++l) {
        Level& L = e->lv[l];
        if (L.kps.empty()) continue;
        blurred.resize((size_t)L.w * L.h);
        gaussian7(L.roi(), L.w, L.h, L.stride, blurred.data(), L.w);
        const float scale = e->sf[l];
        for (orc_keypoint k : L.kps) {
            uint8_t d[32];
            orb_descriptor(blurred.data() + (size_t)cvRoundf(k.y) * L.w + cvRoundf(k.x), L.w, k.angle, d);
            if (l != 0) { k.x *= scale; k.y *= scale; }
            int dst;
            if (k.x >= lap0 && k.x <= lap1) dst = stereoIndex--;
            else dst = monoIndex++;
            if (write) { kps[dst] = k; memcpy(desc + (size_t)dst * 32, d, 32); }
        }
    }
    if (mono_index) *mono_index = monoIndex;
    return nkeypoints;
}

int orc_level_size(const orc_extractor* e, int level, int* w, int* h) {
    if (level < 0 || level >= e->nlevels) return -1;
    *w = e->lv[level].w; *h = e->lv[level].h;
    return 0;
}
void orc_level_bordered(const orc_extractor* e, int level, uint8_t* out) {
    const Level& L = e->lv[level];
    memcpy(out, L.buf.data(), L.buf.size());
}
int orc_level_candidates(const orc_extractor* e, int level, int* xs, int* ys, int* resp, int max_out) {
    const Level& L = e->lv[level];
    for (size_t i = 0; i < L.cands.size() && (int)i < max_out; ++i) {
        xs[i] = (int)L.cands[i].x; ys[i] = (int)L.cands[i].y; resp[i] = (int)L.cands[i].resp;
    }
    return (int)L.cands.size();
}
int orc_level_keypoints(const orc_extractor* e, int level, orc_keypoint* out, int max_out) {
    const Level& L = e->lv[level];
    for (size_t i = 0; i < L.kps.size() && (int)i < max_out; ++i) out[i] = L.kps[i];
    return (int)L.kps.size();
}

int orc_distribute_octtree(int n, const int* xs, const int* ys, const int* resp, int minX, int maxX,
                           int minY, int maxY, int N, int* out_idx, int max_out) {
    std::vector<Cand> v(n);
    for (int i = 0; i < n; ++i) { v[i].x = (float)xs[i]; v[i].y = (float)ys[i]; v[i].resp = (float)resp[i]; v[i].idx = i; }
    std::vector<Cand> sel = DistributeOctTree(v, minX, maxX, minY, maxY, N);
    for (size_t i = 0; i < sel.size() && (int)i < max_out; ++i) out_idx[i] = sel[i].idx;
    return (int)sel.size();
}

float orc_ic_angle(const uint8_t* roi, int stride, int x, int y) {
    std::vector<int> umax;
    compute_umax(umax);
    return ic_angle(roi + (size_t)y * stride + x, stride, umax.data());
}
void orc_orb_descriptor(const uint8_t* blurred_roi, int stride, int x, int y, float angle_deg, uint8_t* desc32) {
    orb_descriptor(blurred_roi + (size_t)y * stride + x, stride, angle_deg, desc32);
}

/* =============================== matcher =============================== */
static const int TH_LOW = 30, TH_HIGH = 100, HISTO_LENGTH = 30; /* include/ORBmatcher.h:73-75 */

int orc_descriptor_distance(const uint8_t* a, const uint8_t* b) { /* src/ORBmatcher.cc:1375-1391 */
    int32_t pa[8], pb[8];
    memcpy(pa, a, 32); memcpy(pb, b, 32);
    int dist = 0;
    for (int i = 0; i < 8; i++) {
        unsigned int v = pa[i] ^ pb[i];
        v = v - ((v >> 1) & 0x55555555);
        v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
        dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
    }
    return dist;
}
void orc_hamming_pairs(const uint8_t* a, const uint8_t* b, int n, int* out) {
    for (int i = 0; i < n; ++i) out[i] = orc_descriptor_distance(a + (size_t)i * 32, b + (size_t)i * 32);
}

void orc_three_maxima(const int* counts, int L, int* ind1, int* ind2, int* ind3) { /* :1328-1370 */
    int max1 = 0, max2 = 0, max3 = 0;
    *ind1 = *ind2 = *ind3 = -1;
    for (int i = 0; i < L; i++) {
        const int s = counts[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; *ind3 = *ind2; *ind2 = *ind1; *ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; *ind3 = *ind2; *ind2 = i; }
        else if (s > max3) { max3 = s; *ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { *ind2 = -1; *ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) { *ind3 = -1; }
}

namespace {
/* Frame::AssignFeaturesToGrid + PosInGrid, src/Frame.cc:157-176, 470-480 */
std::vector<std::vector<int>> build_grid(const orc_frame_view* F) {
    std::vector<std::vector<int>> grid((size_t)F->gcols * F->grows);
    for (int i = 0; i < F->n; ++i) {
        const int px = (int)roundf((F->kx[i] - F->minx) * F->gw_inv);
        const int py = (int)roundf((F->ky[i] - F->miny) * F->gh_inv);
        const int lin = py * F->gcols + px;   /* computeLinearGridIndex: row-major, include/Frame.h */
        if (lin >= 0 && lin < F->gcols * F->grows) grid[lin].push_back(i);
    }
    return grid;
}
/* Frame::GetFeaturesInArea, src/Frame.cc:404-468 */
void features_in_area(const orc_frame_view* F, const std::vector<std::vector<int>>& grid, float x,
                      float y, float r, int minLevel, int maxLevel, std::vector<int>& out) {
    out.clear();
    const int nMinCellX = std::max(0, (int)floorf((x - F->minx - r) * F->gw_inv));
    if (nMinCellX >= F->gcols) return;
    const int nMaxCellX = std::min(F->gcols - 1, (int)ceilf((x - F->minx + r) * F->gw_inv));
    if (nMaxCellX < 0) return;
    const int nMinCellY = std::max(0, (int)floorf((y - F->miny - r) * F->gh_inv));
    if (nMinCellY >= F->grows) return;
    const int nMaxCellY = std::min(F->grows - 1, (int)ceilf((y - F->miny + r) * F->gh_inv));
    if (nMaxCellY < 0) return;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
        for (int iy = nMinCellY; iy <= nMaxCellY; iy++) {
            const std::vector<int>& cell = grid[(size_t)iy * F->gcols + ix];
            for (int idx : cell) {
                if (bCheckLevels)
                    if (F->koct[idx] < minLevel || (maxLevel >= 0 && F->koct[idx] > maxLevel)) continue;
                const float distx = F->kx[idx] - x, disty = F->ky[idx] - y;
                if (fabsf(distx) < r && fabsf(disty) < r) out.push_back(idx);
            }
        }
}
} // namespace

/* MapPoint::ComputeDistinctiveDescriptors, src/MapPoint.cc:343-412, for one descriptor set */
int orc_distinctive_descriptor(const uint8_t* desc, int N, int* median_out) {
    if (N <= 0) { if (median_out) *median_out = -1; return -1; }
    std::vector<float> Distances((size_t)N * N);
    for (int i = 0; i < N; i++) {
        Distances[(size_t)i * N + i] = 0;
        for (int j = i + 1; j < N; j++) {
            const int distij = orc_descriptor_distance(desc + (size_t)i * 32, desc + (size_t)j * 32);
            Distances[(size_t)i * N + j] = (float)distij;
            Distances[(size_t)j * N + i] = (float)distij;
        }
    }
    int BestMedian = INT_MAX, BestIdx = 0;
    for (int i = 0; i < N; i++) {
        std::vector<int> vDists(Distances.begin() + (size_t)i * N, Distances.begin() + (size_t)(i + 1) * N);
        std::sort(vDists.begin(), vDists.end());
        const int median = vDists[(size_t)(0.5 * (N - 1))];
        if (median < BestMedian) { BestMedian = median; BestIdx = i; }
    }
    if (median_out) *median_out = BestMedian;
    return BestIdx;
}

/* DBoW2 FORB::meanValue, Thirdparty/DBoW2/src/DBoW2/FORB.cpp:28-79, on N descriptors of 32 bytes; mean: 32 bytes
 * (all zero for N == 0, where the reference releases the Mat) */
void orc_forb_mean_value(const uint8_t* desc, int N, uint8_t* mean) {
    memset(mean, 0, 32);
    if (N <= 0) return;
    if (N == 1) { memcpy(mean, desc, 32); return; }
    std::vector<int> sum(32 * 8, 0);
    for (int i = 0; i < N; ++i) {
        const unsigned char* p = desc + (size_t)i * 32;
        for (int j = 0; j < 32; ++j, ++p) {
            if (*p & (1 << 7)) ++sum[j * 8];
            if (*p & (1 << 6)) ++sum[j * 8 + 1];
            if (*p & (1 << 5)) ++sum[j * 8 + 2];
            if (*p & (1 << 4)) ++sum[j * 8 + 3];
            if (*p & (1 << 3)) ++sum[j * 8 + 4];
            if (*p & (1 << 2)) ++sum[j * 8 + 5];
            if (*p & (1 << 1)) ++sum[j * 8 + 6];
            if (*p & (1)) ++sum[j * 8 + 7];
        }
    }
    unsigned char* p = mean;
    const int N2 = N / 2 + N % 2;
    for (size_t i = 0; i < sum.size(); ++i) {
        if (sum[i] >= N2) *p |= 1 << (7 - (i % 8));
        if (i % 8 == 7) ++p;
    }
}

/* Frame::AssignFeaturesToGrid as CSR (cells in linear order, push_back order inside a cell) */
void orc_assign_features_to_grid(const orc_frame_view* F, int* cell_off, int* cell_idx) {
    const std::vector<std::vector<int>> grid = build_grid(F);
    int o = 0;
    for (size_t c = 0; c < grid.size(); ++c) {
        cell_off[c] = o;
        for (int i : grid[c]) cell_idx[o++] = i;
    }
    cell_off[grid.size()] = o;
    for (; o < F->n; ++o) cell_idx[o] = -1;
}

/* Frame::GetFeaturesInArea for nq queries; out_idx is [nq][cap], out_count the full vIndices sizes */
void orc_features_in_area(const orc_frame_view* F, int nq, const float* x, const float* y, const float* r,
                          const int* minLevel, const int* maxLevel, int* out_count, int* out_idx, int cap) {
    const std::vector<std::vector<int>> grid = build_grid(F);
    std::vector<int> v;
    for (int q = 0; q < nq; ++q) {
        features_in_area(F, grid, x[q], y[q], r[q], minLevel[q], maxLevel[q], v);
        out_count[q] = (int)v.size();
        for (int k = 0; k < (int)v.size() && k < cap; ++k) out_idx[(size_t)q * cap + k] = v[k];
    }
}

int orc_search_by_projection(const orc_frame_view* F, int nmp, const uint8_t* mp_inview,
                             const float* mp_x, const float* mp_y, const int* mp_level,
                             const float* mp_viewcos, const float* mp_depth, const uint8_t* mp_desc,
                             const float* mp_proj_xr, const uint8_t* mp_has_obs,
                             const uint8_t* claimed_in, float th, int bFarPoints, float thFarPoints,
                             float nnRatio, int* match_mp) {
    /* src/ORBmatcher.cc:31-123.  match_mp[idx] plays F->mvpMapPoints[idx]: a keypoint is skipped when
     * it held a MapPoint with observations before the call (claimed_in) or when the MapPoint assigned
     * to it earlier in this call has Observations() > 0 (:77-79, tested per point: mp_has_obs, NULL =
     * every map point has observations); an assignment by a point without observations can be
     * overwritten, and every assignment counts in nmatches (:108-110). */
    std::vector<std::vector<int>> grid = build_grid(F);
    for (int i = 0; i < F->n; ++i) match_mp[i] = -1;
    int nmatches = 0;
    const bool bFactor = th != 1.0;
    std::vector<int> vIndices;
    for (int iMP = 0; iMP < nmp; iMP++) {
        if (!mp_inview[iMP]) continue;
        if (bFarPoints && mp_depth[iMP] > thFarPoints) continue;
        const int nPredictedLevel = mp_level[iMP];
        float r = (mp_viewcos[iMP] > 0.998) ? 2.5f : 4.0f; /* RadiusByViewingCos :125-131 */
        if (bFactor) r *= th;
        r *= F->scale[nPredictedLevel];
        features_in_area(F, grid, mp_x[iMP], mp_y[iMP], r, nPredictedLevel - 1, nPredictedLevel, vIndices);
        if (vIndices.empty()) continue;
        const uint8_t* MPdescriptor = mp_desc + (size_t)iMP * 32;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
            if (claimed_in && claimed_in[idx]) continue;
            if (match_mp[idx] >= 0 && (!mp_has_obs || mp_has_obs[match_mp[idx]])) continue;
            /* :81-86  F->Nleft == -1 && F->mvuRight[idx] > 0: the right-image coordinate must agree with
             * the projection mTrackProjXR (src/Frame.cc:325).  F->uright == NULL: monocular frame. */
            if (F->uright && F->uright[idx] > 0) {
                const float er = fabsf(mp_proj_xr[iMP] - F->uright[idx]);
                if (er > r) continue;
            }
            const int dist = orc_descriptor_distance(MPdescriptor, F->desc + (size_t)idx * 32);
            if (dist < bestDist) {
                bestDist2 = bestDist; bestDist = dist;
                bestLevel2 = bestLevel; bestLevel = F->koct[idx];
                bestIdx = idx;
            } else if (dist < bestDist2) {
                bestLevel2 = F->koct[idx];
                bestDist2 = dist;
            }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > nnRatio * bestDist2) continue;
            if (bestLevel != bestLevel2 || bestDist <= nnRatio * bestDist2) {
                match_mp[bestIdx] = iMP;
                nmatches++;
            }
        }
    }
    return nmatches;
}

int orc_search_by_projection_kf(const orc_frame_view* F, const float* fangle, int nmp,
                                const uint8_t* mp_valid, const float* mp_u, const float* mp_v,
                                const int* mp_level, const float* mp_angle, const uint8_t* mp_desc,
                                const uint8_t* occupied_in, float th, int checkOrientation,
                                int* match_mp) {
    /* src/ORBmatcher.cc:1202-1326 (projection, frustum and distance gates done by the caller) */
    std::vector<std::vector<int>> grid = build_grid(F);
    std::vector<uint8_t> occupied(F->n, 0);
    for (int i = 0; i < F->n; ++i) { occupied[i] = occupied_in ? occupied_in[i] : 0; match_mp[i] = -1; }
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<int> vIndices2;
    for (int i = 0; i < nmp; i++) {
        if (!mp_valid[i]) continue;
        const int nPredictedLevel = mp_level[i];
        const float radius = th * F->scale[nPredictedLevel];
        features_in_area(F, grid, mp_u[i], mp_v[i], radius, nPredictedLevel - 1, nPredictedLevel + 1, vIndices2);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = mp_desc + (size_t)i * 32;
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            if (occupied[i2]) continue;
            const int dist = orc_descriptor_distance(dMP, F->desc + (size_t)i2 * 32);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {
            match_mp[bestIdx2] = i;
            occupied[bestIdx2] = 1;
            nmatches++;
            if (checkOrientation) {
                float rot = mp_angle[i] - fangle[bestIdx2];
                if (rot < 0.0) rot += 360.0f;
                int bin = (int)roundf(rot * factor);
                if (bin == HISTO_LENGTH) bin = 0;
                rotHist[bin].push_back(bestIdx2);
            }
        }
    }
    if (checkOrientation) {
        int counts[HISTO_LENGTH], ind1, ind2, ind3;
        for (int i = 0; i < HISTO_LENGTH; ++i) counts[i] = (int)rotHist[i].size();
        orc_three_maxima(counts, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < HISTO_LENGTH; i++)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int j : rotHist[i]) { match_mp[j] = -1; nmatches--; }
    }
    return nmatches;
}

static void fuse_search_impl(const orc_frame_view* KF, const float* uright, const float* inv_sigma2, int nmp,
                             const uint8_t* valid, const float* mu, const float* mv, const float* mur, const int* mlevel,
                             const uint8_t* mdesc, float th, int variant, int accept, int* best_idx, int* best_dist) {
    /* search core of Fuse, src/ORBmatcher.cc:762-824 (variant 0) and :919-956 (variant 1);
     * KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:790-833) has no octave filter */
    const std::vector<std::vector<int>> grid = build_grid(KF);
    std::vector<int> vIndices;
    for (int i = 0; i < nmp; ++i) {
        best_idx[i] = -1; best_dist[i] = -1;
        if (!valid[i]) continue;
        const int nPredictedLevel = mlevel[i];
        const float radius = th * KF->scale[nPredictedLevel];
        features_in_area(KF, grid, mu[i], mv[i], radius, -1, -1, vIndices);
        if (vIndices.empty()) continue;
        const uint8_t* dMP = mdesc + (size_t)i * 32;
        int bestDist = variant == 0 ? 256 : INT_MAX, bestIdx = -1;
        for (int idx : vIndices) {
            const int kpLevel = KF->koct[idx];
            if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
            if (variant == 0) {
                const float kpx = KF->kx[idx], kpy = KF->ky[idx];
                if (uright && mur && uright[idx] >= 0) {
                    const float kpr = uright[idx];
                    const float ex = mu[i] - kpx, ey = mv[i] - kpy, er = mur[i] - kpr;
                    const float e2 = ex * ex + ey * ey + er * er;
                    if (e2 * inv_sigma2[kpLevel] > 7.8) continue;
                } else {
                    const float ex = mu[i] - kpx, ey = mv[i] - kpy;
                    const float e2 = ex * ex + ey * ey;
                    if (e2 * inv_sigma2[kpLevel] > 5.99) continue;
                }
            }
            const int dist = orc_descriptor_distance(dMP, KF->desc + (size_t)idx * 32);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (bestIdx >= 0) best_dist[i] = bestDist;
        if (bestDist <= accept) best_idx[i] = bestIdx;
    }
}

void orc_fuse_search(const orc_frame_view* KF, const float* uright, const float* inv_sigma2, int nmp,
                     const uint8_t* valid, const float* mu, const float* mv, const float* mur, const int* mlevel,
                     const uint8_t* mdesc, float th, int variant, int* best_idx, int* best_dist) {
    fuse_search_impl(KF, uright, inv_sigma2, nmp, valid, mu, mv, mur, mlevel, mdesc, th, variant, TH_LOW, best_idx, best_dist);
}

int orc_search_by_sim3(const orc_frame_view* KF1, const orc_frame_view* KF2,
                       const uint8_t* valid1, const float* u1in2, const float* v1in2, const int* level1, const uint8_t* desc1,
                       const uint8_t* valid2, const float* u2in1, const float* v2in1, const int* level2, const uint8_t* desc2,
                       float th, int* match12) {
    /* src/ORBmatcher.cc:977-1200 after the projections: vnMatch1 / vnMatch2 with bestDist <= TH_HIGH
     * (:1076-1079, :1140-1143), then the agreement check (:1148-1163) */
    const int N1 = KF1->n, N2 = KF2->n;
    std::vector<int> vnMatch1(std::max(N1, 1), -1), vnMatch2(std::max(N2, 1), -1), d1(std::max(N1, 1)), d2(std::max(N2, 1));
    fuse_search_impl(KF2, nullptr, nullptr, N1, valid1, u1in2, v1in2, nullptr, level1, desc1, th, 1, TH_HIGH, vnMatch1.data(), d1.data());
    fuse_search_impl(KF1, nullptr, nullptr, N2, valid2, u2in1, v2in1, nullptr, level2, desc2, th, 1, TH_HIGH, vnMatch2.data(), d2.data());
    int nFound = 0;
    for (int i1 = 0; i1 < N1; i1++) {
        match12[i1] = -1;
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0) {
            const int idx1 = vnMatch2[idx2];
            if (idx1 == i1) { match12[i1] = idx2; nFound++; }
        }
    }
    return nFound;
}

int orc_search_for_initialization(const orc_frame_view* F1, const float* angle1, const orc_frame_view* F2,
                                  const float* angle2, int windowSize, float nnRatio, int checkOrientation,
                                  int* vnMatches12) {
    /* src/ORBmatcher.cc:329-439 */
    int nmatches = 0;
    for (int i = 0; i < F1->n; ++i) vnMatches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    std::vector<int> vMatchedDistance(F2->n, INT_MAX), vnMatches21(F2->n, -1);
    const std::vector<std::vector<int>> grid = build_grid(F2);
    std::vector<int> vIndices2;
    for (int i1 = 0; i1 < F1->n; i1++) {
        const int level1 = F1->koct[i1];
        if (level1 > 0) continue;
        features_in_area(F2, grid, F1->kx[i1], F1->ky[i1], (float)windowSize, level1, level1, vIndices2);
        if (vIndices2.empty()) continue;
        const uint8_t* d1 = F1->desc + (size_t)i1 * 32;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            const int dist = orc_descriptor_distance(d1, F2->desc + (size_t)i2 * 32);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * nnRatio) {
                if (vnMatches21[bestIdx2] >= 0) { vnMatches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
                vnMatches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                nmatches++;
                if (checkOrientation) {
                    float rot = angle1[i1] - angle2[bestIdx2];
                    if (rot < 0.0) rot += 360.0f;
                    int bin = (int)roundf(rot * factor);
                    if (bin == HISTO_LENGTH) bin = 0;
                    rotHist[bin].push_back(i1);
                }
            }
        }
    }
    if (checkOrientation) {
        int counts[HISTO_LENGTH], ind1, ind2, ind3;
        for (int i = 0; i < HISTO_LENGTH; i++) counts[i] = (int)rotHist[i].size();
        orc_three_maxima(counts, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); j++) {
                const int idx1 = rotHist[i][j];
                if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; nmatches--; }
            }
        }
    }
    return nmatches;
}

int orc_search_by_bow(const orc_kf_view* KF, const orc_kf_view* F, int Nleft, float nnRatio,
                      int checkOrientation, int* match_kf) {
    /* src/ORBmatcher.cc:133-327.  match_kf[i] = KeyFrame keypoint whose MapPoint lands in
     * vpMapPointMatches[i], or -1; KF->has_mp[i] = (pMP && !pMP->isBad()). */
    for (int i = 0; i < F->n; ++i) match_kf[i] = -1;
    int nmatches = 0;
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int KFit = 0, Fit = 0;
    while (KFit < KF->nnodes && Fit < F->nnodes) {
        if (KF->node_id[KFit] == F->node_id[Fit]) {
            for (int iKF = KF->node_off[KFit]; iKF < KF->node_off[KFit + 1]; iKF++) {
                const int realIdxKF = KF->node_idx[iKF];
                if (!KF->has_mp[realIdxKF]) continue;
                const uint8_t* dKF = KF->desc + (size_t)realIdxKF * 32;
                int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
                int bestDist1R = 256, bestIdxFR = -1, bestDist2R = 256;
                for (int iF = F->node_off[Fit]; iF < F->node_off[Fit + 1]; iF++) {
                    const int realIdxF = F->node_idx[iF];
                    if (match_kf[realIdxF] >= 0) continue;
                    const int dist = orc_descriptor_distance(dKF, F->desc + (size_t)realIdxF * 32);
                    if (Nleft == -1) {
                        if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
                        else if (dist < bestDist2) bestDist2 = dist;
                    } else {
                        if (realIdxF < Nleft && dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
                        else if (realIdxF < Nleft && dist < bestDist2) bestDist2 = dist;
                        if (realIdxF >= Nleft && dist < bestDist1R) { bestDist2R = bestDist1R; bestDist1R = dist; bestIdxFR = realIdxF; }
                        else if (realIdxF >= Nleft && dist < bestDist2R) bestDist2R = dist;
                    }
                }
                if (bestDist1 <= TH_LOW) {
                    if ((float)bestDist1 < nnRatio * (float)bestDist2) {
                        match_kf[bestIdxF] = realIdxKF;
                        if (checkOrientation) {
                            float rot = KF->kangle[realIdxKF] - F->kangle[bestIdxF];
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)roundf(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back(bestIdxF);
                        }
                        nmatches++;
                    }
                    if (bestDist1R <= TH_LOW) {
                        match_kf[bestIdxFR] = realIdxKF;         /* ratio test is `|| true` in the fork */
                        if (checkOrientation) {
                            float rot = KF->kangle[realIdxKF] - F->kangle[bestIdxFR];
                            if (rot < 0.0) rot += 360.0f;
                            int bin = (int)roundf(rot * factor);
                            if (bin == HISTO_LENGTH) bin = 0;
                            rotHist[bin].push_back(bestIdxFR);
                        }
                        nmatches++;
                    }
                }
            }
            KFit++; Fit++;
        } else if (KF->node_id[KFit] < F->node_id[Fit]) {
            while (KFit < KF->nnodes && KF->node_id[KFit] < F->node_id[Fit]) KFit++;   /* lower_bound */
        } else {
            while (Fit < F->nnodes && F->node_id[Fit] < KF->node_id[KFit]) Fit++;
        }
    }
    if (checkOrientation) {
        int counts[HISTO_LENGTH], ind1, ind2, ind3;
        for (int i = 0; i < HISTO_LENGTH; i++) counts[i] = (int)rotHist[i].size();
        orc_three_maxima(counts, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (size_t j = 0; j < rotHist[i].size(); j++) { match_kf[rotHist[i][j]] = -1; nmatches--; }
        }
    }
    return nmatches;
}

int orc_search_for_triangulation(const orc_kf_view* k1, const orc_kf_view* k2, const float* F12,
                                 float epx, float epy, const float* scale2, int bOnlyStereo,
                                 int bCoarse, int checkOrientation, int* pairs, int* npairs) {
    /* src/ORBmatcher.cc:441-676; Pinhole::epipolarConstrain src/CameraModels/Pinhole.cpp:104-131 */
    int nmatches = 0;
    std::vector<int> vMatches12(k1->n, -1);
    std::vector<int> rotHist[HISTO_LENGTH];
    const float factor = 1.0f / HISTO_LENGTH;
    int f1 = 0, f2 = 0;
    while (f1 < k1->nnodes && f2 < k2->nnodes) {
        if (k1->node_id[f1] == k2->node_id[f2]) {
            for (int i1 = k1->node_off[f1]; i1 < k1->node_off[f1 + 1]; i1++) {
                const int idx1 = k1->node_idx[i1];
                if (k1->has_mp && k1->has_mp[idx1]) continue;
                const bool bStereo1 = k1->uright && k1->uright[idx1] >= 0;
                if (bOnlyStereo && !bStereo1) continue;
                const uint8_t* d1 = k1->desc + (size_t)idx1 * 32;
                int bestDist = TH_LOW, bestIdx2 = -1;
                for (int i2 = k2->node_off[f2]; i2 < k2->node_off[f2 + 1]; i2++) {
                    const int idx2 = k2->node_idx[i2];
                    if (k2->has_mp && k2->has_mp[idx2]) continue;  /* vbMatched2 is never set (:485,537) */
                    const bool bStereo2 = k2->uright && k2->uright[idx2] >= 0;
                    if (bOnlyStereo && !bStereo2) continue;
                    const int dist = orc_descriptor_distance(d1, k2->desc + (size_t)idx2 * 32);
                    if (dist > TH_LOW || dist > bestDist) continue;
                    if (!bStereo1 && !bStereo2) {
                        const float distex = epx - k2->kx[idx2], distey = epy - k2->ky[idx2];
                        const float err = distex * distex + distey * distey;
                        if (err < 100 * scale2[k2->koct[idx2]]) continue;
                    }
                    /* epipolarConstrain(..., unc = 1.0) */
                    const float x1 = k1->kx[idx1], y1 = k1->ky[idx1];
                    const float a = x1 * F12[0] + y1 * F12[3] + F12[6];
                    const float b = x1 * F12[1] + y1 * F12[4] + F12[7];
                    const float c = x1 * F12[2] + y1 * F12[5] + F12[8];
                    const float num = a * k2->kx[idx2] + b * k2->ky[idx2] + c;
                    const float den = a * a + b * b;
                    bool ok = false;
                    if (den != 0) {
                        const float dsqr = num * num / den;
                        ok = dsqr < (float)(3.84 * 1.0);
                    }
                    if (bCoarse || ok) { bestIdx2 = idx2; bestDist = dist; }
                }
                if (bestIdx2 >= 0) {
                    vMatches12[idx1] = bestIdx2;
                    nmatches++;
                    if (checkOrientation) {
                        float rot = k1->kangle[idx1] - k2->kangle[bestIdx2];
                        if (rot < 0.0) rot += 360.0f;
                        int bin = (int)roundf(rot * factor);
                        if (bin == HISTO_LENGTH) bin = 0;
                        rotHist[bin].push_back(idx1);
                    }
                }
            }
            f1++; f2++;
        } else if (k1->node_id[f1] < k2->node_id[f2]) {
            while (f1 < k1->nnodes && k1->node_id[f1] < k2->node_id[f2]) f1++;   /* lower_bound */
        } else {
            while (f2 < k2->nnodes && k2->node_id[f2] < k1->node_id[f1]) f2++;
        }
    }
    if (checkOrientation) {
        int counts[HISTO_LENGTH], ind1, ind2, ind3;
        for (int i = 0; i < HISTO_LENGTH; ++i) counts[i] = (int)rotHist[i].size();
        orc_three_maxima(counts, HISTO_LENGTH, &ind1, &ind2, &ind3);
        for (int i = 0; i < HISTO_LENGTH; i++) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int j : rotHist[i]) { vMatches12[j] = -1; nmatches--; }
        }
    }
    int np = 0;
    for (int i = 0; i < k1->n; ++i) {
        if (vMatches12[i] < 0) continue;
        pairs[2 * np] = i; pairs[2 * np + 1] = vMatches12[i];
        np++;
    }
    *npairs = np;
    return nmatches;
}

void orc_stereo_matches(const orc_stereo_side* L, const orc_stereo_side* R, const float* scale,
                        const float* inv_scale, float mbf, float mb, float* uRight, float* depth) {
    /* Frame::ComputeStereoMatches, upstream v1.0 (absent from the fork: include/Frame.h:53-62);
     * SURVEY.md Appendix A5. SAD on raw u8 (no centre subtraction). */
    const int N = L->n;
    for (int i = 0; i < N; ++i) { uRight[i] = -1.0f; depth[i] = -1.0f; }
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const int nRows = L->lvl_h[0];
    std::vector<std::vector<int>> vRowIndices(nRows);
    for (int iR = 0; iR < R->n; iR++) {
        const float kpY = R->ky[iR];
        const float r = 2.0f * scale[R->koct[iR]];
        const int maxr = (int)ceilf(kpY + r), minr = (int)floorf(kpY - r);
        for (int yi = minr; yi <= maxr; yi++)
            if (yi >= 0 && yi < nRows) vRowIndices[yi].push_back(iR);
    }
    const float minZ = mb, minD = 0, maxD = mbf / minZ;
    std::vector<std::pair<int, int>> vDistIdx;
    for (int iL = 0; iL < N; iL++) {
        const int levelL = L->koct[iL];
        const float vL = L->ky[iL], uL = L->kx[iL];
        const int row = (int)vL;
        if (row < 0 || row >= nRows) continue;
        const std::vector<int>& vCandidates = vRowIndices[row];
        if (vCandidates.empty()) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH, bestIdxR = 0;
        const uint8_t* dL = L->desc + (size_t)iL * 32;
        for (int iR : vCandidates) {
            if (R->koct[iR] < levelL - 1 || R->koct[iR] > levelL + 1) continue;
            const float uR = R->kx[iR];
            if (uR >= minU && uR <= maxU) {
                const int dist = orc_descriptor_distance(dL, R->desc + (size_t)iR * 32);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist < thOrbDist) {
            const float uR0 = R->kx[bestIdxR];
            const float scaleFactor = inv_scale[levelL];
            const float scaleduL = roundf(uL * scaleFactor);
            const float scaledvL = roundf(vL * scaleFactor);
            const float scaleduR0 = roundf(uR0 * scaleFactor);
            const int w = 5, Lw = 5;
            const uint8_t* pl = L->lvl[levelL]; const int sl = L->lvl_stride[levelL];
            const uint8_t* pr = R->lvl[levelL]; const int sr = R->lvl_stride[levelL];
            int bestSAD = INT_MAX, bestincR = 0;
            float vDists[11];
            const float iniu = scaleduR0 + Lw - w, endu = scaleduR0 + Lw + w + 1;
            if (iniu < 0 || endu >= R->lvl_w[levelL]) continue;
            const int y0 = (int)(scaledvL - w), xl0 = (int)(scaleduL - w);
            for (int incR = -Lw; incR <= +Lw; incR++) {
                const int xr0 = (int)(scaleduR0 + incR - w);
                int sad = 0;
                for (int yy = 0; yy < 2 * w + 1; ++yy)
                    for (int xx = 0; xx < 2 * w + 1; ++xx)
                        sad += std::abs((int)pl[(size_t)(y0 + yy) * sl + xl0 + xx] - (int)pr[(size_t)(y0 + yy) * sr + xr0 + xx]);
                const float dist = (float)sad;
                if (dist < bestSAD) { bestSAD = (int)dist; bestincR = incR; }
                vDists[Lw + incR] = dist;
            }
            if (bestincR == -Lw || bestincR == Lw) continue;
            const float dist1 = vDists[Lw + bestincR - 1], dist2 = vDists[Lw + bestincR], dist3 = vDists[Lw + bestincR + 1];
            const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
            if (deltaR < -1 || deltaR > 1) continue;
            float bestuR = scale[levelL] * ((float)scaleduR0 + (float)bestincR + deltaR);
            float disparity = (uL - bestuR);
            if (disparity >= minD && disparity < maxD) {
                if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
                depth[iL] = mbf / disparity;
                uRight[iL] = bestuR;
                vDistIdx.push_back(std::pair<int, int>(bestSAD, iL));
            }
        }
    }
    if (vDistIdx.empty()) return;   /* upstream indexes vDistIdx[0] here (UB); defined as no-op */
    std::sort(vDistIdx.begin(), vDistIdx.end());
    const float median = vDistIdx[vDistIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
        if (vDistIdx[i].first < thDist) break;
        uRight[vDistIdx[i].second] = -1; depth[vDistIdx[i].second] = -1;
    }
}

void orc_knn2(const uint8_t* q, int nq, const uint8_t* db, int ndb, int* d1, int* i1, int* d2,
              int* i2, int nthreads) {
    auto work = [&](int lo, int hi) {
        for (int a = lo; a < hi; ++a) {
            uint64_t qa[4]; memcpy(qa, q + (size_t)a * 32, 32);
            int b1 = 257, b2 = 257, j1 = -1, j2 = -1;
            for (int b = 0; b < ndb; ++b) {
                uint64_t db_[4]; memcpy(db_, db + (size_t)b * 32, 32);
                const int d = __builtin_popcountll(qa[0] ^ db_[0]) + __builtin_popcountll(qa[1] ^ db_[1]) +
                              __builtin_popcountll(qa[2] ^ db_[2]) + __builtin_popcountll(qa[3] ^ db_[3]);
                if (d < b1) { b2 = b1; j2 = j1; b1 = d; j1 = b; }
                else if (d < b2) { b2 = d; j2 = b; }
            }
            d1[a] = b1; i1[a] = j1; d2[a] = b2; i2[a] = j2;
        }
    };
    if (nthreads <= 1) { work(0, nq); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t) th.emplace_back(work, (int)((long)nq * t / nthreads), (int)((long)nq * (t + 1) / nthreads));
    for (auto& t : th) t.join();
}

double orc_bench_extract(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh,
                         const uint8_t* images, int nimages, int width, int height,
                         int iters_per_thread, int nthreads, long* total_kps) {
    std::atomic<long> kps{0};
    auto work = [&](int t) {
        orc_extractor* e = orc_extractor_create(nfeatures, scaleFactor, nlevels, iniTh, minTh);
        std::vector<orc_keypoint> k(nfeatures * 2 + 64);
        std::vector<uint8_t> d((size_t)(nfeatures * 2 + 64) * 32);
        const int lap[2] = {0, 0};
        long s = 0;
        for (int it = 0; it < iters_per_thread; ++it) {
            const uint8_t* img = images + (size_t)((t * iters_per_thread + it) % nimages) * width * height;
            int mono;
            s += orc_extract(e, img, width, height, width, lap, k.data(), d.data(), (int)k.size(), &mono);
        }
        kps += s;
        orc_extractor_destroy(e);
    };
    auto t0 = std::chrono::steady_clock::now();
    if (nthreads <= 1) work(0);
    else {
        std::vector<std::thread> th;
        for (int t = 0; t < nthreads; ++t) th.emplace_back(work, t);
        for (auto& t : th) t.join();
    }
    auto t1 = std::chrono::steady_clock::now();
    if (total_kps) *total_kps = kps.load();
    return std::chrono::duration<double>(t1 - t0).count();
}

} // extern "C"
