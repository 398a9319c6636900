// The cv::Mat / cv::KeyPoint form of the drop-in (include/orbx/ORBextractor.h under ORBX_WITH_OPENCV), compiled
// against tests/cpp/opencv_stub and run the way upstream Frame's stereo constructor drives it: two extractors
// (left / right, as upstream's two threads), operator() on each, then ComputeStereoMatches, reading mvImagePyramid
// as std::vector<cv::Mat>-like.  usage: test_shim_opencv <left.raw> <right.raw> <w> <h> <nfeatures> <bf> <b> <out.blob>
// With six more arguments  <bgr.raw> <map1.raw> <map2.raw> <src_w> <src_h> <out2.blob>  it then runs
// ros2_orbslam3::ImageGrabber (include/orbx/ImageGrabber.h): ConvertImageToGPU(cv::Mat) to w x h, and the chained
// ConvertAndExtract into exL.
#include <cstdio>
#include <cstdlib>
#include <thread>
#include "../../include/orbx/ORBextractor.h"
#include "../../include/orbx/ImageGrabber.h"
#include "blob.h"

static std::vector<uint8_t> load(const char* p, size_t n) {
    std::vector<uint8_t> v(n);
    FILE* f = fopen(p, "rb");
    if (!f || fread(v.data(), 1, n, f) != n) throw std::runtime_error("cannot read image");
    fclose(f);
    return v;
}

int main(int argc, char** argv) {
    if (argc < 9) return 2;
    const int w = atoi(argv[3]), h = atoi(argv[4]), nf = atoi(argv[5]);
    const float bf = (float)atof(argv[6]), b = (float)atof(argv[7]);
    try {
        std::vector<uint8_t> lbuf = load(argv[1], (size_t)w * h), rbuf = load(argv[2], (size_t)w * h);
        cv::Mat imL(h, w, CV_8UC1, lbuf.data()), imR(h, w, CV_8UC1, rbuf.data());
        ORB_SLAM3::ORBextractor exL(nf, 1.2f, 8, 20, 7), exR(nf, 1.2f, 8, 20, 7);
        std::vector<cv::KeyPoint> kL, kR;
        cv::Mat dL, dR;
        std::vector<int> lap = {0, 0};
        int monoL = 0, monoR = 0;
        /* upstream Frame::Frame (stereo): thread threadLeft(&Frame::ExtractORB, this, 0, imLeft, 0, 0); thread threadRight(...) */
        std::thread tl([&] { monoL = exL(imL, cv::Mat(), kL, dL, lap); });
        std::thread tr([&] { monoR = exR(imR, cv::Mat(), kR, dR, lap); });
        tl.join(); tr.join();
        static_assert(sizeof(cv::KeyPoint) == 28, "cv::KeyPoint layout");
        std::vector<float> uR, depth;
        ORB_SLAM3::ComputeStereoMatches(exL, exR, bf, b, uR, depth);
        /* mvImagePyramid as upstream's ComputeStereoMatches reads it: a Mat ROI with a valid 19-px border around it */
        const cv::Mat& L1 = exL.mvImagePyramid[1];
        const cv::Mat patch = L1.rowRange(20, 31).colRange(30, 41);           /* 11 x 11 window */
        int psum = 0;
        for (int y = 0; y < patch.rows; ++y) for (int x = 0; x < patch.cols; ++x) psum += patch.at<uint8_t>(y, x);
        const int border_px = L1.ptr(0)[-19 - 19 * (ptrdiff_t)L1.step];       /* top-left corner of the bordered buffer */
        Blob out;
        out.put_int("monoL", monoL); out.put_int("monoR", monoR);
        out.put("kL", (const uint8_t*)kL.data(), kL.size() * sizeof(cv::KeyPoint));
        out.put("kR", (const uint8_t*)kR.data(), kR.size() * sizeof(cv::KeyPoint));
        out.put("dL", dL.data, (size_t)dL.rows * 32); out.put("dR", dR.data, (size_t)dR.rows * 32);
        out.put("uR", uR); out.put("depth", depth);
        int lv[6] = {L1.rows, L1.cols, (int)L1.step, psum, border_px, (int)exL.mvImagePyramid.size()};
        out.put("L1", lv, 6);
        std::vector<uint8_t> roi((size_t)L1.rows * L1.cols);
        for (int y = 0; y < L1.rows; ++y) memcpy(&roi[(size_t)y * L1.cols], L1.ptr(y), L1.cols);
        out.put("L1roi", roi);
        out.save(argv[8]);
        if (argc >= 15) {
            const int sw = atoi(argv[12]), sh = atoi(argv[13]);
            std::vector<uint8_t> bgr = load(argv[9], (size_t)sw * sh * 3), m1 = load(argv[10], (size_t)sw * sh * 4),
                                 m2 = load(argv[11], (size_t)sw * sh * 4);
            cv::Mat im(sh, sw, CV_8UC3, bgr.data()), map1(sh, sw, CV_32FC1, m1.data()), map2(sh, sw, CV_32FC1, m2.data());
            ros2_orbslam3::ImageGrabber igb(cv::Size(sw, sh), map1, map2, w, h);
            cv::Mat grey = igb.ConvertImageToGPU(im);
            Blob o2;
            o2.put("grey", grey.data, (size_t)grey.rows * grey.cols);
            std::vector<orbx_keypoint> k;
            std::vector<uint8_t> d;
            const int mono = igb.ConvertAndExtract(im.data, im.step, exL, k, d, lap);
            o2.put_int("mono", mono);
            o2.put("k", (const uint8_t*)k.data(), k.size() * sizeof(orbx_keypoint));
            o2.put("d", d);
            o2.save(argv[14]);
        }
        cv::Mat none;
        if (exL(none, cv::Mat(), kL, dL, lap) != -1) return 4;                /* empty image -> -1 */
    } catch (const std::exception& e) {
        fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    printf("OK\n");
    return 0;
}
